// Block NTT + gadget MAC of the hybrid key switch (log N >= 12): for one target modulus and one 2048-point block,
//     acc_c = sum_d evk[d][c] (.) NTT(ModUp_d(y))            (lattigo DecomposeSingleNTT + gadgetProductMultiplePLazy,
//                                                             called from InnerSum at fhe/ligero.go:325)
// The digits arrive base-converted with the coupling levels applied (k_modup_blk), the digit's own limbs come straight from
// the NTT-domain input.  Key rows are in the block read order (key_slot, kernels.cu): pair j of thread u at slot j*256 + u.
//
//   k_mac_ra   register-fed: 64-bit lazy accumulators, operands prefetched into registers one stage ahead
//   k_mac_rd   TMA-fed: residues and key rows arrive by bulk copies (cp.async.bulk + mbarrier) one stage ahead without
//              holding registers, which makes room for 128-bit accumulators (4 IMAD.WIDE per product, one reduction per
//              output instead of one per product)
#include "kernels.cuh"
#include "ntt_r16.cuh"

#include <cstdlib>

namespace lmn {

template <int S> struct LogS2 { static constexpr int v = S == 8 ? 3 : S == 4 ? 2 : S == 2 ? 1 : 0; };

template <int S, int OCC, int PF>
__global__ void __launch_bounds__(RA_THREADS, OCC) k_mac_ra(const ModDev *mods, const ulonglong2 *tw_fwd, const KsConsts *C,
                                                           const KsJobs J, const u64 *__restrict__ evk) {
    extern __shared__ u64 sm[];
    constexpr int M = RA_M;
    constexpr int LOGN = RA_LOGM + LogS2<S>::v;
    ulonglong2 *stw = reinterpret_cast<ulonglong2 *>(sm);
    u64 *buf0 = sm + 2 * RA_M, *buf1 = buf0 + RA_WORDS;
    const int tix = blockIdx.x / S, h = blockIdx.x % S, job = blockIdx.y;
    const int nq = C->nq, k = C->k, KP = k + C->np;
    const int gm = ks_mod_index(tix, nq, k);
    const ModDev md = mods[gm];
    const u64 p = md.q;
    const FwdMod fm = {p, 0 - p, md.q3};
    const unsigned u = threadIdx.x;
    const size_t blk = (size_t)h * M;
    const ulonglong2 *c1 = reinterpret_cast<const ulonglong2 *>(J.col + (size_t)job * J.col_words + ((size_t)(k + tix) << LOGN) + blk + 8 * u);
    const u64 *mu = J.mu + (size_t)job * J.mu_words + ((size_t)gm << LOGN) + blk + u;
    const int beta = C->beta;
    int own = -1;   // digit whose own limb this target is (copied from the NTT-domain input, DecomposeSingleNTT)
    for (int d = 0; d < beta; d++) if (tix < nq && tix >= C->dig_start[d] && tix < C->dig_start[d] + C->dig_size[d]) own = d;
    ra_fill_stw(stw, tw_fwd + ((size_t)gm << LOGN), S + h);

    auto load_digit = [&](int d, u64 (&x)[8]) {
        if (d == own) {
#pragma unroll
            for (int r = 0; r < 4; r++) { const ulonglong2 v = __ldg(c1 + r); x[2 * r] = v.x; x[2 * r + 1] = v.y; }
        } else {
            const u64 *src = mu + (((size_t)d * KP) << LOGN);
#pragma unroll
            for (int r = 0; r < 8; r++) x[r] = __ldg(src + 256 * r);
        }
    };
    u64 acc0[8], acc1[8], a[8];
#pragma unroll
    for (int r = 0; r < 8; r++) { acc0[r] = 0; acc1[r] = 0; }
    load_digit(0, a);
    __syncthreads();   // twiddle table visible
    int flip = 0;
    for (int d = 0; d < beta; d++) {
        const ulonglong2 *e0 = reinterpret_cast<const ulonglong2 *>(evk + ((((size_t)d * 2 + 0) * KP + gm) << LOGN) + blk) + u;
        const ulonglong2 *e1 = reinterpret_cast<const ulonglong2 *>(evk + ((((size_t)d * 2 + 1) * KP + gm) << LOGN) + blk) + u;
        // key rows: component 0 is requested before the transform, component 1 before its third pass; both are plain loads
        // pinned in front of the next barrier (ld_pin2), a full pass or more ahead of the MAC that consumes them
        ulonglong2 k0[4], k1[4];
#pragma unroll
        for (int r = 0; r < 4; r++) k0[r] = ld_pin2(e0 + 256 * r);
        auto load_k1 = [&]() {
#pragma unroll
            for (int r = 0; r < 4; r++) k1[r] = ld_pin2(e1 + 256 * r);
        };
        if (d != own) {
            if (PF & 2) ntt2048_fwd(a, flip ? buf1 : buf0, stw, fm, load_k1);
            else { ntt2048_fwd(a, flip ? buf1 : buf0, stw, fm, []() {}); load_k1(); }
            flip ^= 1;
        } else load_k1();
        u64 nx[8];
        if ((PF & 1) && d + 1 < beta) load_digit(d + 1, nx);
#pragma unroll
        for (int r = 0; r < 4; r++) {
            acc0[2 * r] += mont_mul_lazy(a[2 * r], k0[r].x, p, md.qinv);
            acc0[2 * r + 1] += mont_mul_lazy(a[2 * r + 1], k0[r].y, p, md.qinv);
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
            acc1[2 * r] += mont_mul_lazy(a[2 * r], k1[r].x, p, md.qinv);
            acc1[2 * r + 1] += mont_mul_lazy(a[2 * r + 1], k1[r].y, p, md.qinv);
        }
        if (d + 1 < beta) {
            if (PF & 1) {
#pragma unroll
                for (int r = 0; r < 8; r++) a[r] = nx[r];
            } else load_digit(d + 1, a);
        }
    }
    ulonglong2 *o0 = reinterpret_cast<ulonglong2 *>(J.acc + (size_t)job * J.acc_words + ((size_t)gm << LOGN) + blk + 8 * u);
    ulonglong2 *o1 = o0 + ((size_t)KP << (LOGN - 1));
#pragma unroll
    for (int r = 0; r < 4; r++) {
        ulonglong2 r0, r1;
        r0.x = barrett(acc0[2 * r], p, md.mu); r0.y = barrett(acc0[2 * r + 1], p, md.mu);
        r1.x = barrett(acc1[2 * r], p, md.mu); r1.y = barrett(acc1[2 * r + 1], p, md.mu);
        o0[r] = r0; o1[r] = r1;
    }
}

// ---------------------------------------------------------------------------------------------------
// TMA-fed variant.  One CTA owns (target modulus, block h) for jb consecutive jobs; its work is a sequence of items
// (job, digit).  Thread 0 requests with bulk copies (cp.async.bulk, completion on an mbarrier) the 16 KB of residues of
// item i+1 and the 2 x 16 KB of key rows of item i right after the one CTA barrier of item i that proves the landing
// buffers free; the copies land while the other warps transform.  Shared memory: 32 KB twiddles + 17 KB transposition
// buffer + 16 KB residues + 32 KB key rows = 97 KB, two CTAs per SM.
// ---------------------------------------------------------------------------------------------------
typedef unsigned __int128 u128;
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *bar, unsigned parity) {
    asm volatile("{\n\t.reg .pred p;\n\tLMN_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra LMN_DONE;\n\tbra LMN_WAIT;\n\tLMN_DONE:\n\t}"
                 ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

constexpr size_t MAC_RD_SMEM = (2 * (size_t)RA_M + (size_t)RA_WORDS + (size_t)RA_M + 2 * (size_t)RA_M + 2) * 8;
template <int S>
__global__ void __launch_bounds__(RA_THREADS, 2) k_mac_rd(const ModDev *mods, const ulonglong2 *tw_fwd, const KsConsts *C,
                                                           const KsJobs J, const u64 *__restrict__ evk, int jb) {
    extern __shared__ __align__(16) u64 sm[];
    constexpr int M = RA_M;
    constexpr int LOGN = RA_LOGM + LogS2<S>::v;
    constexpr unsigned BLK_BYTES = M * 8;
    ulonglong2 *stw = reinterpret_cast<ulonglong2 *>(sm);
    u64 *buf = sm + 2 * RA_M;
    u64 *stage = buf + RA_WORDS;          // residues of the next item: words u + 256 r (own limb: 8u + r)
    u64 *keys = stage + RA_M;             // [2][2048] key rows of the current item, block read order
    u64 *bars = keys + 2 * RA_M;          // [0] residues landed, [1] key rows landed
    const int tix = blockIdx.x / S, h = blockIdx.x % S;
    const int nq = C->nq, k = C->k, KP = k + C->np;
    const int gm = ks_mod_index(tix, nq, k);
    const ModDev md = mods[gm];
    const u64 p = md.q;
    const FwdMod fm = {p, 0 - p, md.q3};
    const unsigned u = threadIdx.x;
    const size_t blk = (size_t)h * M;
    const int beta = C->beta;
    int own = -1;   // digit whose own limb this target is (copied from the NTT-domain input, DecomposeSingleNTT)
    for (int d = 0; d < beta; d++) if (tix < nq && tix >= C->dig_start[d] && tix < C->dig_start[d] + C->dig_size[d]) own = d;
    const int j0 = blockIdx.y * jb, j1 = min(J.jobs, j0 + jb);
    const int nit = (j1 - j0) * beta;

    auto issue_data = [&](int job, int d) {
        const u64 *src = d == own ? J.col + (size_t)job * J.col_words + ((size_t)(k + tix) << LOGN) + blk
                                  : J.mu + (size_t)job * J.mu_words + (((size_t)d * KP + gm) << LOGN) + blk;
        mbar_expect_tx(&bars[0], BLK_BYTES);
        bulk_g2s(stage, src, BLK_BYTES, &bars[0]);
    };
    auto issue_keys = [&](int d) {
        const u64 *e0 = evk + ((((size_t)d * 2 + 0) * KP + gm) << LOGN) + blk;
        const u64 *e1 = evk + ((((size_t)d * 2 + 1) * KP + gm) << LOGN) + blk;
        mbar_expect_tx(&bars[1], 2 * BLK_BYTES);
        bulk_g2s(keys, e0, BLK_BYTES, &bars[1]);
        bulk_g2s(keys + M, e1, BLK_BYTES, &bars[1]);
    };
    if (u == 0) {
        mbar_init(&bars[0], 1); mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    ra_fill_stw(stw, tw_fwd + ((size_t)gm << LOGN), S + h);
    __syncthreads();   // twiddle table and barriers visible
    if (u == 0) issue_data(j0, 0);

    u128 acc0[8], acc1[8];
#pragma unroll
    for (int r = 0; r < 8; r++) { acc0[r] = 0; acc1[r] = 0; }
    const u64 p2 = 2 * p, p4 = 4 * p, p8 = 8 * p;
    auto fin = [&](u128 v) {
        // (hi:lo) 2^-64 mod p = hi - floor(m p / 2^64) + p, in (0, hi + p], hi < 14 p
        u64 t = mont_reduce_lazy((u64)v, (u64)(v >> 64), p, md.qinv);
        t = t >= p8 ? t - p8 : t; t = t >= p4 ? t - p4 : t; t = t >= p2 ? t - p2 : t;
        return csub(t, p);
    };
    int d = 0, job = j0;
    for (int it = 0; it < nit; it++) {
        u64 a[8];
        mbar_wait(&bars[0], it & 1);
        auto after_read = [&]() {
            // The bulk copies requested below overwrite `stage` and `keys` through the async proxy: every thread's reads of them
            // must be complete (not merely issued) and ordered before it (measured: without this a copy occasionally lands
            // before a queued LDS of the own-limb path has executed).
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();   // all threads hold their residues, are past the previous MAC and the previous pass D
            if (u == 0) {
                issue_keys(d);
                if (it + 1 < nit) { const int dn = d + 1 == beta ? 0 : d + 1; issue_data(dn ? job : job + 1, dn); }
            }
        };
        if (d == own) {
#pragma unroll
            for (int r = 0; r < 4; r++) { const ulonglong2 v = reinterpret_cast<const ulonglong2 *>(stage)[4 * u + r]; a[2 * r] = v.x; a[2 * r + 1] = v.y; }
            {   // no transform consumes these loads before the barrier: make their completion a register dependency
                u64 x = 0;
#pragma unroll
                for (int r = 0; r < 8; r++) x |= a[r];
                if (x == ~0ull) stage[u] = x;   // residues are below 2^59: never taken
            }
            after_read();
        } else {
#pragma unroll
            for (int r = 0; r < 8; r++) a[r] = stage[u + 256 * r];
            ntt2048_fwd(a, buf, stw, fm, after_read, []() {});
        }
        mbar_wait(&bars[1], it & 1);
        const ulonglong2 *kp = reinterpret_cast<const ulonglong2 *>(keys) + u;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const ulonglong2 k0 = kp[256 * r], k1 = kp[1024 + 256 * r];
            acc0[2 * r] += (u128)a[2 * r] * k0.x;
            acc0[2 * r + 1] += (u128)a[2 * r + 1] * k0.y;
            acc1[2 * r] += (u128)a[2 * r] * k1.x;
            acc1[2 * r + 1] += (u128)a[2 * r + 1] * k1.y;
        }
        if (++d == beta) {
            ulonglong2 *o0 = reinterpret_cast<ulonglong2 *>(J.acc + (size_t)job * J.acc_words + ((size_t)gm << LOGN) + blk + 8 * u);
            ulonglong2 *o1 = o0 + ((size_t)KP << (LOGN - 1));
#pragma unroll
            for (int r = 0; r < 4; r++) {
                ulonglong2 r0, r1;
                r0.x = fin(acc0[2 * r]); r0.y = fin(acc0[2 * r + 1]);
                r1.x = fin(acc1[2 * r]); r1.y = fin(acc1[2 * r + 1]);
                o0[r] = r0; o1[r] = r1;
                acc0[2 * r] = 0; acc0[2 * r + 1] = 0; acc1[2 * r] = 0; acc1[2 * r + 1] = 0;
            }
            d = 0; job++;
        }
    }
}

// =====================================================================================================
void launch_mac_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J, const u64 *evk) {
    LaunchScope scope(L, KID_MAC);
    const int SB = blk_blocks(T.logn);
    static const int var = getenv("LMN_MAC_VAR") ? atoi(getenv("LMN_MAC_VAR")) : -1;   // experiment switch: -1 TMA-fed, 0 register-fed
    if (var >= 0) {
        const size_t smm = (2 * (size_t)RA_M + 2 * (size_t)RA_WORDS) * 8;
        const dim3 g((Ch.nq + Ch.np) * SB, J.jobs);
        if (SB == 8) k_mac_ra<8, 2, 3><<<g, RA_THREADS, smm, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk);
        else if (SB == 4) k_mac_ra<4, 2, 3><<<g, RA_THREADS, smm, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk);
        else k_mac_ra<2, 2, 3><<<g, RA_THREADS, smm, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk);
        return;
    }
    static const int jbe = getenv("LMN_MAC_JB") ? atoi(getenv("LMN_MAC_JB")) : 0;
    const int ctas = (Ch.nq + Ch.np) * SB;
    int jb = jbe > 0 ? jbe : (ctas * ((J.jobs + 1) / 2) >= 8 * 148 * 2 ? 2 : 1);
    const dim3 g(ctas, (J.jobs + jb - 1) / jb);
    if (SB == 8) k_mac_rd<8><<<g, RA_THREADS, MAC_RD_SMEM, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk, jb);
    else if (SB == 4) k_mac_rd<4><<<g, RA_THREADS, MAC_RD_SMEM, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk, jb);
    else k_mac_rd<2><<<g, RA_THREADS, MAC_RD_SMEM, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk, jb);
}

const char *kernels_mac_init() {
    cudaError_t e = cudaSuccess;
    const int maxsm = 227 * 1024;
#define SETSM(K) do { if (e == cudaSuccess) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, maxsm); } while (0)
    SETSM((k_mac_ra<2, 2, 3>)); SETSM((k_mac_ra<4, 2, 3>)); SETSM((k_mac_ra<8, 2, 3>));
    SETSM(k_mac_rd<2>); SETSM(k_mac_rd<4>); SETSM(k_mac_rd<8>);
#undef SETSM
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace lmn
