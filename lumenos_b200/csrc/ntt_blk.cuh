// 2048-point block core of the negacyclic NTT / INTT (sm_100a).
//
// A ring of N = 2048*S coefficients (S = 2, 4, 8  <->  log N = 12, 13, 14) is transformed as S independent block transforms
// of M = 2048 points; the log2(S) levels that couple the blocks are applied by the elementwise kernel that produces
// (forward) or consumes (inverse) the data.  One CTA of 256 threads owns one block: every thread carries 8 residues in
// registers through 3 butterfly levels per pass (3 + 3 + 3 + 2), with three trips through padded shared memory in between;
// all 2047 twiddles of the block sit in shared memory.  (A 4096-point radix-16 core was measured slower: fewer resident
// warps per scheduler.)
//
// Arithmetic (measured with tools/int_pipe_probe.cu): IMAD.WIDE.U32 occupies the fmaheavy pipe for 4 cycles per warp, IMAD
// for 2, the ALU pipe is separate.  The butterflies therefore use a Shoup quotient estimate that drops the low x low
// partial product (3 wide multiplies instead of 4): the estimate is at most 1 below the exact one, so the lazy product
// lies in [0, 3q) and forward values grow by 3q per level.  With every modulus below 2^64 / (3 log N + 8) (checked in
// lmn_ctx_create) nothing overflows, and one Barrett step at the end gives the canonical residue - the same number the exact
// transform gives.  1167 vs 940 Gbutterfly/s register-resident.
//
// Twiddles are Lattigo's tables (ring.SubRing.RootsForward / RootsBackward): index m + i at the level with m groups.
#pragma once
#include "device.cuh"

namespace lmn {

// Loads that must be ISSUED where they are written (prefetch into registers one or two stages ahead of their use):
// ptxas sinks ld.global.nc towards the first use to shorten live ranges - even across bar.sync - which puts an L2 round
// trip in front of the consumer when only 4 warps per scheduler are resident.  A plain (coherent) ld.global may not move
// across the CTA barrier that follows it, so it stays in the stage where it is written.  (ld.volatile also stays put, but
// is issued as LDG.STRONG.SYS and measured 10 % slower end to end.)
__device__ __forceinline__ ulonglong2 ld_pin2(const ulonglong2 *p) {
    ulonglong2 v;
    asm volatile("ld.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ u64 ld_pin(const u64 *p) {
    u64 v;
    asm volatile("ld.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// modulus, 2^64 - q, 3q, and z = a zero that ptxas cannot see (read from ModDev.pad[0]).  ptxas lowers a two-operand add-with-carry
// to either IADD3.X (ALU pipe) or IMAD.X (fmaheavy pipe) to "balance" the pipes; in these kernels fmaheavy is the saturated pipe
// (IMAD.WIDE holds it 4 cycles), so the carry adds of the butterflies take z as a third register operand, a form only IADD3.X has.
struct FwdMod { u64 q, nq, q3; unsigned z; };
// a + b, the high half as a three-register add-with-carry (always IADD3.X)
__device__ __forceinline__ u64 add64_alu(u64 a, u64 b, unsigned z) {
    unsigned lo, hi;
    asm("add.cc.u32 %0, %2, %3;\n\taddc.u32 %1, %4, %5;" : "=r"(lo), "=r"(hi) : "r"((unsigned)a), "r"((unsigned)b), "r"((unsigned)(a >> 32)), "r"((unsigned)(b >> 32)));
    hi += z;
    return ((u64)hi << 32) | lo;
}

// y * w - hi' * q  with  hi' in { floor(y*ws/2^64) - 1, floor(y*ws/2^64) } : result in [0, 3q) for any y < 2^64
__device__ __forceinline__ u64 mul_shoup_approx(u64 y, u64 w, u64 ws, u64 nq, unsigned z = 0) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 yl, yh, wl, wh, sl, sh, nl, nh, alo, ahi, clo, chi, mlo, mhi, mc, hl, hh, tlo, thi;\n\t"
        ".reg .u64 a, c, ad, hi, t;\n\t"
        "mov.b64 {yl, yh}, %1;\n\t" "mov.b64 {wl, wh}, %2;\n\t" "mov.b64 {sl, sh}, %3;\n\t" "mov.b64 {nl, nh}, %4;\n\t"
        "mul.wide.u32 a, yl, sh;\n\t" "mul.wide.u32 c, yh, sl;\n\t"
        "mov.b64 {alo, ahi}, a;\n\t" "mov.b64 {clo, chi}, c;\n\t"
        "add.cc.u32 mlo, alo, clo;\n\t" "addc.cc.u32 mhi, ahi, chi;\n\t" "addc.u32 mc, 0, 0;\n\t"
        "mul.wide.u32 hi, yh, sh;\n\t"
        "mov.b64 {hl, hh}, hi;\n\t"
        "add.cc.u32 hl, hl, mhi;\n\t" "addc.u32 hh, hh, mc;\n\t" "add.u32 hh, hh, %5;\n\t"
        "mul.wide.u32 t, yl, wl;\n\t" "mad.wide.u32 t, hl, nl, t;\n\t"
        "mov.b64 {tlo, thi}, t;\n\t"
        "mad.lo.u32 thi, yl, wh, thi;\n\t" "mad.lo.u32 thi, yh, wl, thi;\n\t"
        "mad.lo.u32 thi, hl, nh, thi;\n\t" "mad.lo.u32 thi, hh, nl, thi;\n\t"
        "mov.b64 %0, {tlo, thi};\n\t"
        "}" : "=l"(r) : "l"(y), "l"(w), "l"(ws), "l"(nq), "r"(z));
    return r;
}
// The same lazy product in the formulation of tools/int_pipe_probe.cu (k_bfly<1>): the sum of the two middle products rides into the
// high product as the 64-bit addend of IMAD.WIDE, and nothing is pinned to the ALU pipe.  5 IMAD.WIDE + 4 IMAD per butterfly; it is
// the fastest register-resident butterfly measured on this part (1169 Gbutterfly/s) and therefore the ceiling lmn_bench_int_peak
// reports.  Inside the block kernels it was A/B-tested against the ALU-pinned form above (build with -DLMN_BFLY_MADWIDE and / or
// -DLMN_BFLY_PLAINADD; results in DESIGN.md section 4).
__device__ __forceinline__ u64 mul_shoup_approx_mw(u64 y, u64 w, u64 ws, u64 nq) {
    u64 r;
    asm("{\n\t"
        ".reg .u32 yl, yh, wl, wh, sl, sh, nl, nh, alo, ahi, clo, chi, mlo, mhi, mc, hl, hh, tlo, thi;\n\t"
        ".reg .u64 a, c, ad, hi, t;\n\t"
        "mov.b64 {yl, yh}, %1;\n\t" "mov.b64 {wl, wh}, %2;\n\t" "mov.b64 {sl, sh}, %3;\n\t" "mov.b64 {nl, nh}, %4;\n\t"
        "mul.wide.u32 a, yl, sh;\n\t" "mul.wide.u32 c, yh, sl;\n\t"
        "mov.b64 {alo, ahi}, a;\n\t" "mov.b64 {clo, chi}, c;\n\t"
        "add.cc.u32 mlo, alo, clo;\n\t" "addc.cc.u32 mhi, ahi, chi;\n\t" "addc.u32 mc, 0, 0;\n\t"
        "mov.b64 ad, {mhi, mc};\n\t"
        "mad.wide.u32 hi, yh, sh, ad;\n\t"
        "mov.b64 {hl, hh}, hi;\n\t"
        "mul.wide.u32 t, yl, wl;\n\t" "mad.wide.u32 t, hl, nl, t;\n\t"
        "mov.b64 {tlo, thi}, t;\n\t"
        "mad.lo.u32 thi, yl, wh, thi;\n\t" "mad.lo.u32 thi, yh, wl, thi;\n\t"
        "mad.lo.u32 thi, hl, nh, thi;\n\t" "mad.lo.u32 thi, hh, nl, thi;\n\t"
        "mov.b64 %0, {tlo, thi};\n\t"
        "}" : "=l"(r) : "l"(y), "l"(w), "l"(ws), "l"(nq));
    return r;
}
#ifdef LMN_BFLY_MADWIDE
#define LMN_MUL_APPROX(y, w, ws, m) mul_shoup_approx_mw(y, w, ws, (m).nq)
#else
#define LMN_MUL_APPROX(y, w, ws, m) mul_shoup_approx(y, w, ws, (m).nq, (m).z)
#endif
__device__ __forceinline__ void bfly_ct3(u64 &x, u64 &y, const ulonglong2 w, const FwdMod &m) {
    const u64 t = LMN_MUL_APPROX(y, w.x, w.y, m);
    y = x - t + m.q3;
#ifdef LMN_BFLY_PLAINADD
    x = x + t;
#else
    x = add64_alu(x, t, m.z);
#endif
}
// the probe's butterfly, whatever the build flags say (k_peak_bfly)
__device__ __forceinline__ void bfly_ct3_probe(u64 &x, u64 &y, const ulonglong2 w, const FwdMod &m) {
    const u64 t = mul_shoup_approx_mw(y, w.x, w.y, m.nq);
    y = x - t + m.q3;
    x = x + t;
}

constexpr int RA_THREADS = 256;
constexpr int RA_LOGM = 11;
constexpr int RA_M = 1 << RA_LOGM;
constexpr int RA_WORDS = RA_M + (RA_M >> 4);

// table index of the twiddle of (block-local level l, group g); levels 9 and 10 are stored thread-major ([b][u]) so that
// the last pass reads them conflict-free
__device__ __forceinline__ void ra_fill_stw(ulonglong2 *stw, const ulonglong2 *__restrict__ tw, unsigned SH) {
    for (unsigned idx = threadIdx.x; idx < RA_M; idx += RA_THREADS) {
        unsigned l, g;
        if (idx < 512) { if (idx == 0) continue; l = 31 - __clz(idx); g = idx - (1u << l); }
        else if (idx < 1024) { l = 9; g = 2 * (idx & 255) + ((idx - 512) >> 8); }
        else { l = 10; g = 4 * (idx & 255) + ((idx - 1024) >> 8); }
        stw[idx] = __ldg(&tw[(SH << l) + g]);
    }
}
// 3 Cooley-Tukey levels on 8 registers (a[r] <-> offset r*stride); twiddle of (level t, sub-group b) = tw_of(t, b)
template <class F>
__device__ __forceinline__ void r8_fwd_f(u64 (&a)[8], F tw_of, const FwdMod &m) {
#pragma unroll
    for (int t = 0; t < 3; t++) {
        const int half = 4 >> t;
#pragma unroll
        for (int b = 0; b < (1 << t); b++) {
            const ulonglong2 w = tw_of(t, b);
#pragma unroll
            for (int j = 0; j < half; j++) bfly_ct3(a[b * 2 * half + j], a[b * 2 * half + j + half], w, m);
        }
    }
}
// forward transform of one 2048-point block: in a[r] = point u + 256 r, out a[r] = point 8u + r (lazy, entry bound + 33 q).
// `s` = RA_WORDS words; the caller guarantees the previous readers of `s` are done (double buffer or barrier).
// Hook0 runs after the first register stage and before the first store to `s` (a barrier there makes one buffer enough).
template <class Hook0, class Hook>
__device__ __forceinline__ void ntt2048_fwd(u64 (&a)[8], u64 *s, const ulonglong2 *stw, const FwdMod &m, Hook0 before_store_a, Hook before_pass_c) {
    const unsigned u = threadIdx.x;
    r8_fwd_f(a, [&](int t, int b) { return stw[(1 << t) + b]; }, m);                       // levels 0-2
    before_store_a();
#pragma unroll
    for (int r = 0; r < 8; r++) s[PAD(u + 256 * r)] = a[r];
    __syncthreads();
    { const unsigned i = u >> 5, j = u & 31;
#pragma unroll
      for (int r = 0; r < 8; r++) a[r] = s[PAD(i * 256 + j + 32 * r)];
      r8_fwd_f(a, [&](int t, int b) { return stw[((8 + i) << t) + b]; }, m);               // levels 3-5
#pragma unroll
      for (int r = 0; r < 8; r++) s[PAD(i * 256 + j + 32 * r)] = a[r]; }
    // from here on a warp only touches its own 256-point segment [256 w, 256 w + 255] (levels 3-5 wrote it, levels 6-8 and 9-10 read and
    // write inside it), so the two remaining exchanges need a warp barrier, not a CTA barrier
    __syncwarp();
    { const unsigned i = u >> 2, j = u & 3;
#pragma unroll
      for (int r = 0; r < 8; r++) a[r] = s[PAD(i * 32 + j + 4 * r)];
      before_pass_c();   // the caller's prefetches for its epilogue
      r8_fwd_f(a, [&](int t, int b) { return stw[((64 + i) << t) + b]; }, m);              // levels 6-8
#pragma unroll
      for (int r = 0; r < 8; r++) s[PAD(i * 32 + j + 4 * r)] = a[r]; }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8; r++) a[r] = s[PAD(8 * u + r)];
#pragma unroll
    for (int b = 0; b < 2; b++) {                                                           // level 9: groups 2u + b
        const ulonglong2 w = stw[512 + b * 256 + u];
        bfly_ct3(a[4 * b], a[4 * b + 2], w, m); bfly_ct3(a[4 * b + 1], a[4 * b + 3], w, m);
    }
#pragma unroll
    for (int b = 0; b < 4; b++) {                                                           // level 10: groups 4u + b
        const ulonglong2 w = stw[1024 + b * 256 + u];
        bfly_ct3(a[2 * b], a[2 * b + 1], w, m);
    }
}
template <class Hook>
__device__ __forceinline__ void ntt2048_fwd(u64 (&a)[8], u64 *s, const ulonglong2 *stw, const FwdMod &m, Hook before_pass_c) {
    ntt2048_fwd(a, s, stw, m, []() {}, before_pass_c);
}

// ---------------------------------------------------------------------------------------------------
// inverse of a 2048-point block (Gentleman-Sande).  Invariant [0, 4q): the sum is corrected by 4q once, the difference
// (x - y + 4q < 8q) is multiplied with the 3-product quotient estimate, whose lazy result lies in [0, 3q).  The exact
// transform mod q is unchanged; consumers finish with an exact (canonical) multiplication by N^-1.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bfly_gs3(u64 &x, u64 &y, const ulonglong2 w, const FwdMod &m, u64 q4) {
    const u64 s = x + y;
    const u64 d = x - y + q4;
    x = s >= q4 ? s - q4 : s;
    y = LMN_MUL_APPROX(d, w.x, w.y, m);
}
template <class F>
__device__ __forceinline__ void r8_inv_f(u64 (&a)[8], F tw_of, const FwdMod &m, u64 q4) {
#pragma unroll
    for (int t = 2; t >= 0; t--) {
        const int half = 4 >> t;
#pragma unroll
        for (int b = 0; b < (1 << t); b++) {
            const ulonglong2 w = tw_of(t, b);
#pragma unroll
            for (int j = 0; j < half; j++) bfly_gs3(a[b * 2 * half + j], a[b * 2 * half + j + half], w, m, q4);
        }
    }
}
// in a[r] = point 8u + r of the block (values in [0, 4q)); out a[r] = point u + 256 r, in [0, 4q), NOT scaled by N^-1;
// the log2(S) levels that couple the S blocks of a limb remain to be applied by the consumer.  stw = the block's table of
// INVERSE twiddles (ra_fill_stw on RootsBackward).  `s` as in ntt2048_fwd.
template <class Hook0>
__device__ __forceinline__ void ntt2048_inv(u64 (&a)[8], u64 *s, const ulonglong2 *stw, const FwdMod &m, Hook0 before_store) {
    const unsigned u = threadIdx.x;
    const u64 q4 = 4 * m.q;
#pragma unroll
    for (int b = 0; b < 4; b++) {                                                           // level 10: groups 4u + b
        const ulonglong2 w = stw[1024 + b * 256 + u];
        bfly_gs3(a[2 * b], a[2 * b + 1], w, m, q4);
    }
#pragma unroll
    for (int b = 0; b < 2; b++) {                                                           // level 9: groups 2u + b
        const ulonglong2 w = stw[512 + b * 256 + u];
        bfly_gs3(a[4 * b], a[4 * b + 2], w, m, q4); bfly_gs3(a[4 * b + 1], a[4 * b + 3], w, m, q4);
    }
    before_store();   // runs after the first register stage and before the first store to `s`
#pragma unroll
    for (int r = 0; r < 8; r++) s[PAD(8 * u + r)] = a[r];
    __syncwarp();      // levels 10-6 stay inside the warp's own 256-point segment: warp barriers; the CTA barrier comes before levels 2-0
    { const unsigned i = u >> 2, j = u & 3;
#pragma unroll
      for (int r = 0; r < 8; r++) a[r] = s[PAD(i * 32 + j + 4 * r)];
      r8_inv_f(a, [&](int t, int b) { return stw[((64 + i) << t) + b]; }, m, q4);           // levels 8-6
#pragma unroll
      for (int r = 0; r < 8; r++) s[PAD(i * 32 + j + 4 * r)] = a[r]; }
    __syncwarp();
    { const unsigned i = u >> 5, j = u & 31;
#pragma unroll
      for (int r = 0; r < 8; r++) a[r] = s[PAD(i * 256 + j + 32 * r)];
      r8_inv_f(a, [&](int t, int b) { return stw[((8 + i) << t) + b]; }, m, q4);            // levels 5-3
#pragma unroll
      for (int r = 0; r < 8; r++) s[PAD(i * 256 + j + 32 * r)] = a[r]; }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 8; r++) a[r] = s[PAD(u + 256 * r)];
    r8_inv_f(a, [&](int t, int b) { return stw[(1 << t) + b]; }, m, q4);                    // levels 2-0
}

// the log2(S) Gentleman-Sande levels that couple the S blocks of a limb, on the S values v[r] = point n + r*M of the limb
// (inverse twiddle table of the limb's modulus: index (1 << t) + b at the level with 2^t groups); [0, 4q) in and out
template <int S>
__device__ __forceinline__ void couple_inv(u64 (&v)[S], const ulonglong2 *__restrict__ twi, const FwdMod &m) {
    const u64 q4 = 4 * m.q;
#pragma unroll
    for (int half = 1; half < S; half <<= 1) {          // level with S / (2 half) groups
        const int groups = S / (2 * half);
#pragma unroll
        for (int b = 0; b < groups; b++) {
            const ulonglong2 w = __ldg(&twi[groups + b]);
#pragma unroll
            for (int j = 0; j < half; j++) bfly_gs3(v[b * 2 * half + j], v[b * 2 * half + j + half], w, m, q4);
        }
    }
}
// the log2(S) Cooley-Tukey levels that couple the S blocks (forward twiddle table of the limb's modulus); values grow by 3q
// per level
template <int S>
__device__ __forceinline__ void couple_fwd(u64 (&v)[S], const ulonglong2 *__restrict__ twf, const FwdMod &m) {
#pragma unroll
    for (int half = S / 2; half >= 1; half >>= 1) {     // level with S / (2 half) groups
        const int groups = S / (2 * half);
#pragma unroll
        for (int b = 0; b < groups; b++) {
            const ulonglong2 w = __ldg(&twf[groups + b]);
#pragma unroll
            for (int j = 0; j < half; j++) bfly_ct3(v[b * 2 * half + j], v[b * 2 * half + j + half], w, m);
        }
    }
}
}  // namespace lmn
