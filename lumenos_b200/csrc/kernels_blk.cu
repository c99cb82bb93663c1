// 2048-point block pipeline of the hybrid key switch and the rescale for rings of N = 2048*S coefficients (S = 2, 4, 8
// <-> log N = 12, 13, 14; configs A/B, C, D of BASELINE.json).
//
// Every length-N transform is S independent 2048-point block transforms (256 threads x 8 register-resident residues, all
// of the block's twiddles in shared memory, ntt_blk.cuh) plus log2(S) levels that couple the blocks.  The coupling levels
// are pointwise over the S points n + r*M of a limb, so they live in the elementwise kernels that sit between an inverse
// and a forward transform anyway (base conversions):
//
//   k_blk_intt   raw block INTT                      (DecomposeNTT's INTT of c1; ModDown's INTT of the P part; Rescale's INTT)
//   k_modup_blk  coupling + N^-1 (Q_d/q_i)^-1, float64 v, digit -> every other modulus, forward coupling
//   k_mac_ra     block NTT + gadget MAC              (kernels.cu)
//   k_down_pre   coupling + N^-1 (P/p_j)^-1 + P/2, float64 v, P -> q_i, forward coupling
//   k_blk_fwd<0> block NTT, (acc - ext) P^-1 (+c0), automorphism, out = in + pi(u)
//   k_rs_pre / k_blk_fwd<2>   the same shape for DivRoundByLastModulusNTT
//
// References: lattigo rlwe.Evaluator.DecomposeNTT / GadgetProductHoisted / ModDownQPtoQNTT / AutomorphismHoisted and
// ring.DivRoundByLastModulusNTT as called from fhe/ligero.go:325 (InnerSum) and :331-333, :271-273 (Rescale); restated
// in oracle/lumenos_oracle.c.
#include <cstdlib>
#include "kernels.cuh"
#include "ntt_blk.cuh"
#include "tma.cuh"

namespace lmn {

// resident CTAs per SM each kernel is compiled for (register cap 65536 / (threads * n)).  Measured on B200 at config D with A/B builds:
// the pointwise base-conversion kernels compiled without a cap were slower than with one (256 jobs per launch: k_modup_blk 27.0 -> 20.1 ms,
// k_down_pre 10.5 -> 7.8 ms; at the benchmark's 592 jobs per launch 676 -> 637 ms and 248 -> 243 ms per step); the block-transform
// kernels showed no gain from other values than the ones below, and k_mac_rd with one CTA per SM lost 23 %.
#ifndef LMN_MODUP_MINB
#define LMN_MODUP_MINB 4
#endif
#ifndef LMN_DOWNPRE_MINB
#define LMN_DOWNPRE_MINB 3
#endif
#ifndef LMN_BLKINTT_MINB
#define LMN_BLKINTT_MINB 3
#endif
#ifndef LMN_BLKFWD_MINB
#define LMN_BLKFWD_MINB 2
#endif
#ifndef LMN_CHAIN_MINB
#define LMN_CHAIN_MINB 1
#endif
constexpr size_t BLK_INTT_SMEM = (2 * (size_t)RA_M + (size_t)RA_WORDS + (size_t)RA_M + 2) * 8;       // twiddles, buffer, one landing block, barriers
constexpr size_t BLK_FWD_SMEM = (2 * (size_t)RA_M + (size_t)RA_WORDS + 3 * (size_t)RA_M + 4) * 8;    // twiddles, buffer, three landing blocks, barriers
constexpr unsigned BLK_BYTES = RA_M * 8;
template <int S> struct LogS { static constexpr int v = S == 8 ? 3 : S == 4 ? 2 : S == 2 ? 1 : 0; };

// block twiddle tables in the shared-memory layout of ra_fill_stw, one 32 KB table per (modulus, block): a CTA fetches its
// table with one bulk copy.  grid (S, moduli)
__global__ void k_build_blk_tw(const ulonglong2 *__restrict__ tw, ulonglong2 *__restrict__ out, int logn, int S) {
    const int h = blockIdx.x, gm = blockIdx.y;
    ulonglong2 *o = out + (((size_t)gm * S + h) << RA_LOGM);
    if (threadIdx.x == 0) o[0] = make_ulonglong2(0, 0);
    ra_fill_stw(o, tw + ((size_t)gm << logn), (unsigned)(S + h));
}
void launch_build_blk_tables(cudaStream_t stream, const DeviceTables &T, ulonglong2 *fwd, ulonglong2 *inv) {
    const int S = blk_blocks(T.logn);
    k_build_blk_tw<<<dim3(S, T.k + T.np), RA_THREADS, 0, stream>>>(T.tw_fwd, fwd, T.logn, S);
    k_build_blk_tw<<<dim3(S, T.k + T.np), RA_THREADS, 0, stream>>>(T.tw_inv, inv, T.logn, S);
}

// ---------------------------------------------------------------------------------------------------
// raw block INTT of `nslots` limbs of `jobs` jobs.  grid (S*nslots, ceil(jobs/jb)); one CTA owns block h of slot s for jb
// consecutive jobs.  The block's twiddle table and every job's 16 KB block arrive by bulk copies (TMA), the next job's while
// the current one is transformed.  In-place use (src == dst) is safe: the block of a job has landed in shared memory
// before the CTA writes it, and other jobs' blocks are other addresses.
// ---------------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(RA_THREADS, LMN_BLKINTT_MINB) k_blk_intt(const ModDev *mods, const ulonglong2 *twb_inv, const BlkSlots A, int jobs, int jb) {
    extern __shared__ __align__(16) u64 sm[];
    constexpr int LOGN = RA_LOGM + LogS<S>::v;
    ulonglong2 *stw = reinterpret_cast<ulonglong2 *>(sm);
    u64 *buf = sm + 2 * RA_M;
    u64 *land = buf + RA_WORDS;       // the job's block, natural order
    u64 *bars = land + RA_M;          // [0] block landed, [1] table landed
    const int slot = blockIdx.x / S, h = blockIdx.x % S;
    const int gm = A.mod[slot];
    const ModDev md = mods[gm];
    const FwdMod fm = {md.q, 0 - md.q, md.q3, (unsigned)md.pad[0]};
    const unsigned u = threadIdx.x;
    const size_t soff = ((size_t)A.src_limb[slot] << LOGN) + (size_t)h * RA_M;
    const size_t doff = ((size_t)A.dst_limb[slot] << LOGN) + (size_t)h * RA_M + u;
    const int j0 = blockIdx.y * jb, j1 = min(jobs, j0 + jb);
    if (u == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init_fence(); }
    __syncthreads();
    if (u == 0) {
        mbar_expect_tx(&bars[1], 2 * BLK_BYTES);
        bulk_g2s(stw, twb_inv + (((size_t)gm * S + h) << RA_LOGM), 2 * BLK_BYTES, &bars[1]);
        mbar_expect_tx(&bars[0], BLK_BYTES);
        bulk_g2s(land, A.src + (size_t)j0 * A.src_job_words + soff, BLK_BYTES, &bars[0]);
    }
    mbar_wait(&bars[1], 0);
    for (int job = j0; job < j1; job++) {
        u64 a[8];
        mbar_wait(&bars[0], (job - j0) & 1);
#pragma unroll
        for (int r = 0; r < 4; r++) { const ulonglong2 v = reinterpret_cast<const ulonglong2 *>(land)[4 * u + r]; a[2 * r] = v.x; a[2 * r + 1] = v.y; }
        ntt2048_inv(a, buf, stw, fm, [&]() {
            release_smem_to_async_proxy();
            __syncthreads();   // every thread holds its residues (consumed by the first butterflies) and is past the previous job's last pass
            if (u == 0 && job + 1 < j1) {
                mbar_expect_tx(&bars[0], BLK_BYTES);
                bulk_g2s(land, A.src + (size_t)(job + 1) * A.src_job_words + soff, BLK_BYTES, &bars[0]);
            }
        });
        u64 *o = A.dst + (size_t)job * A.dst_job_words + doff;
#pragma unroll
        for (int r = 0; r < 8; r++) o[256 * r] = a[r];
    }
}


// ---------------------------------------------------------------------------------------------------
// digit d of y (raw block INTTs of c1) -> every other modulus (lattigo DecomposeAndSplit: reconstructRNS + multSum).
// One thread owns the S points n + r*M of the limb: it finishes the INTT (coupling levels, then the exact product with
// N^-1 (Q_d/q_a)^-1), folds the float64 overflow estimate v = floor(sum_a float64(y_a)/float64(q_a)) (same operation order
// as Go) into the first residue, z_0 = y_0 + (nd-1-v) q_0, converts to each target modulus and applies the forward
// coupling levels, so that k_mac_ra runs independent block transforms.  ND = digit size (0: read at run time).
// grid (M/256, digits, jobs)
// ---------------------------------------------------------------------------------------------------
template <int S, int ND>
__global__ void __launch_bounds__(256, LMN_MODUP_MINB) k_modup_blk(const ModDev *mods, const ulonglong2 *tw_fwd, const ulonglong2 *tw_inv,
                                                    const KsConsts *C, const KsJobs J, int d0) {
    constexpr int LOGN = RA_LOGM + LogS<S>::v;
    constexpr int M = RA_M;
    constexpr int NDM = ND ? ND : MAX_P;
    const int d = d0 + blockIdx.y, job = blockIdx.z;
    const int nd = ND ? ND : C->dig_size[d], s0 = C->dig_start[d];
    const int nq = C->nq, k = C->k, KP = k + C->np;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const u64 *y = J.y + (size_t)job * J.y_words;
    u64 *mu = J.mu + (size_t)job * J.mu_words + (((size_t)d * KP) << LOGN);
    u64 z[NDM][S];
#pragma unroll
    for (int a = 0; a < NDM; a++) {
        if (a < nd) {
            const int gm = s0 + a;
            const u64 q = mods[gm].q;
            const FwdMod fm = {q, 0 - q, mods[gm].q3, (unsigned)mods[gm].pad[0]};
            const u64 *src = y + ((size_t)gm << LOGN) + n;
#pragma unroll
            for (int r = 0; r < S; r++) z[a][r] = src[(size_t)r * M];
            couple_inv<S>(z[a], tw_inv + ((size_t)gm << LOGN), fm);
            const u64 w = C->intt_w[gm], ws = C->intt_ws[gm];
#pragma unroll
            for (int r = 0; r < S; r++) z[a][r] = mul_shoup(z[a][r], w, ws, q);
        } else {
#pragma unroll
            for (int r = 0; r < S; r++) z[a][r] = 0;
        }
    }
    if (nd >= 2) {
        const u64 q0 = mods[s0].q;
#pragma unroll
        for (int r = 0; r < S; r++) {
            double vf = 0.0;
#pragma unroll
            for (int a = 0; a < NDM; a++)
                if (a < nd) vf = __dadd_rn(vf, __ddiv_rn(__ull2double_rn(z[a][r]), __ull2double_rn(mods[s0 + a].q)));
            z[0][r] += (u64)(nd - 1 - (int)__double2ull_rz(vf)) * q0;
        }
    }
    const int ntix = nq + C->np;
#pragma unroll 1
    for (int tix = 0; tix < ntix; tix++) {
        if (tix < nq && tix >= s0 && tix < s0 + nd) continue;   // own limbs are copied from the NTT-domain input
        const int gm = ks_mod_index(tix, nq, k);
        const u64 p = mods[gm].q, pinv = mods[gm].qinv;
        const FwdMod fm = {p, 0 - p, mods[gm].q3, (unsigned)mods[gm].pad[0]};
        u64 vals[S];
        if (nd == 1) {
            // single-prime digit: centred lift (DecomposeAndSplit, decompLvl < 0), then the key-row scale 2^(e-64)
            const u64 qd = mods[s0].q, hq = qd >> 1, gw = C->mu_g_w[d][tix], gws = C->mu_g_ws[d][tix], pmu = mods[gm].mu;
#pragma unroll
            for (int r = 0; r < S; r++) {
                const u64 c = z[0][r];
                const bool neg = c >= hq;
                const u64 rr = barrett(neg ? qd - c : c, p, pmu);
                vals[r] = mul_shoup_lazy(neg ? (rr ? p - rr : 0) : rr, gw, gws, p);
            }
        } else if (C->mu_fast[d][tix]) {
            const uint32_t sgn = C->mu_sgn[d][tix];
            uint32_t mag[NDM];
#pragma unroll
            for (int a = 0; a < NDM; a++) mag[a] = a < nd ? C->mu_mag[d][tix][a] : 0;
            const u64 klo = C->mu_k_lo[d][tix], khi = C->mu_k_hi[d][tix];
#pragma unroll
            for (int r = 0; r < S; r++) {
                u64 lo = klo, hi = khi;
#pragma unroll
                for (int a = 0; a < NDM; a++) {
                    if (a < nd) {
                        // 64 x 32 -> 96-bit product with two 32x32 multiplies
                        const u64 p0 = (u64)(uint32_t)z[a][r] * mag[a];
                        const u64 p1 = (u64)(uint32_t)(z[a][r] >> 32) * mag[a];
                        const u64 plo = p0 + (p1 << 32);
                        const u64 phi = (p1 >> 32) + (plo < p0);
                        if ((sgn >> a) & 1) { const u64 t = lo; lo -= plo; hi -= phi + (t < plo); }
                        else { lo += plo; hi += phi + (lo < plo); }
                    }
                }
                vals[r] = mont_reduce_lazy(lo, hi, p, pinv);
            }
        } else {
            u64 cm[NDM];
#pragma unroll
            for (int a = 0; a < NDM; a++) cm[a] = a < nd ? C->mu_c[d][tix][a] : 0;
            const u64 klo = C->mu_k_lo[d][tix], khi = C->mu_k_hi[d][tix];
#pragma unroll
            for (int r = 0; r < S; r++) {
                u64 lo = klo, hi = khi;
#pragma unroll
                for (int a = 0; a < NDM; a++)
                    if (a < nd) mac128(lo, hi, z[a][r], cm[a]);
                vals[r] = csub(mont_reduce_lazy(lo, hi, p, pinv), 2 * p);
            }
        }
        couple_fwd<S>(vals, tw_fwd + ((size_t)gm << LOGN), fm);
        u64 *o = mu + ((size_t)gm << LOGN) + n;
#pragma unroll
        for (int r = 0; r < S; r++) o[(size_t)r * M] = vals[r];
    }
}

// ---------------------------------------------------------------------------------------------------
// ModDown, pointwise part (lattigo BasisExtender.ModDownQPtoQNTT): the P limbs of the accumulators arrive as raw block
// INTTs; finish them (coupling levels, N^-1 (P/p_j)^-1, + P/2), fold the float64 v of ModUpExact into the first residue,
// convert to every q_i and apply the forward coupling levels.  Output: mu[c][i][N].  grid (M/256, 2, jobs)
// ---------------------------------------------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(256, LMN_DOWNPRE_MINB) k_down_pre(const ModDev *mods, const ulonglong2 *tw_fwd, const ulonglong2 *tw_inv,
                                                   const KsConsts *C, const KsJobs J) {
    constexpr int LOGN = RA_LOGM + LogS<S>::v;
    constexpr int M = RA_M;
    const int c = blockIdx.y, job = blockIdx.z;
    const int k = C->k, np = C->np, nq = C->nq, KP = k + np;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const u64 *accp = J.acc + (size_t)job * J.acc_words + (((size_t)c * KP + k) << LOGN) + n;
    u64 z[MAX_P][S];
#pragma unroll
    for (int a = 0; a < MAX_P; a++) {
        if (a < np) {
            const int gm = k + a;
            const u64 p = mods[gm].q;
            const FwdMod fm = {p, 0 - p, mods[gm].q3, (unsigned)mods[gm].pad[0]};
#pragma unroll
            for (int r = 0; r < S; r++) z[a][r] = accp[((size_t)a << LOGN) + (size_t)r * M];
            couple_inv<S>(z[a], tw_inv + ((size_t)gm << LOGN), fm);
            const u64 w = C->pin_w[a], ws = C->pin_ws[a], ad = C->pin_add[a];
#pragma unroll
            for (int r = 0; r < S; r++) z[a][r] = add_mod(mul_shoup(z[a][r], w, ws, p), ad, p);
        } else {
#pragma unroll
            for (int r = 0; r < S; r++) z[a][r] = 0;
        }
    }
    if (np >= 2) {
        const u64 p0 = mods[k].q;
#pragma unroll
        for (int r = 0; r < S; r++) {
            double vf = 0.0;
#pragma unroll
            for (int a = 0; a < MAX_P; a++)
                if (a < np) vf = __dadd_rn(vf, __ddiv_rn(__ull2double_rn(z[a][r]), __ull2double_rn(mods[k + a].q)));
            z[0][r] += (u64)(np - 1 - (int)__double2ull_rz(vf)) * p0;
        }
    }
    u64 *tmp = J.mu + (size_t)job * J.mu_words + (((size_t)c * nq) << LOGN) + n;
#pragma unroll 1
    for (int i = 0; i < nq; i++) {
        const u64 q = mods[i].q, qinv = mods[i].qinv;
        const FwdMod fm = {q, 0 - q, mods[i].q3, (unsigned)mods[i].pad[0]};
        u64 cm[MAX_P];
#pragma unroll
        for (int a = 0; a < MAX_P; a++) cm[a] = a < np ? C->md_c[i][a] : 0;
        const u64 neg = C->md_neg[i];
        u64 vals[S];
#pragma unroll
        for (int r = 0; r < S; r++) {
            u64 lo = 0, hi = neg;
#pragma unroll
            for (int a = 0; a < MAX_P; a++)
                if (a < np) mac128(lo, hi, z[a][r], cm[a]);
            vals[r] = csub(mont_reduce_lazy(lo, hi, q, qinv), 2 * q);
        }
        couple_fwd<S>(vals, tw_fwd + ((size_t)i << LOGN), fm);
        u64 *o = tmp + ((size_t)i << LOGN);
#pragma unroll
        for (int r = 0; r < S; r++) o[(size_t)r * M] = vals[r];
    }
}

// Rescale, pointwise part (ring.DivRoundByLastModulusNTT): acc slot c holds the raw block INTT of limb `level`; finish it
// (coupling, N^-1), add q_l/2, and for every remaining limb reduce, subtract q_l/2 mod q_i and apply the forward coupling
// levels.  Output mu[c][i][N].  grid (M/256, 2, jobs)
template <int S>
__global__ void __launch_bounds__(256) k_rs_pre(const ModDev *mods, const ulonglong2 *tw_fwd, const ulonglong2 *tw_inv,
                                                 const RescaleConsts *R, const KsJobs J, int level) {
    constexpr int LOGN = RA_LOGM + LogS<S>::v;
    constexpr int M = RA_M;
    const int c = blockIdx.y, job = blockIdx.z;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const ModDev ml = mods[level];
    const FwdMod fl = {ml.q, 0 - ml.q, ml.q3, (unsigned)ml.pad[0]};
    const u64 *raw = J.acc + (size_t)job * J.acc_words + ((size_t)c << LOGN) + n;
    u64 v[S];
#pragma unroll
    for (int r = 0; r < S; r++) v[r] = raw[(size_t)r * M];
    couple_inv<S>(v, tw_inv + ((size_t)level << LOGN), fl);
    const u64 half = R->half;
#pragma unroll
    for (int r = 0; r < S; r++) v[r] = add_mod(mul_shoup(v[r], ml.ninv, ml.ninvs, ml.q), half, ml.q);
    u64 *tmp = J.mu + (size_t)job * J.mu_words + (((size_t)c * level) << LOGN) + n;
#pragma unroll 1
    for (int i = 0; i < level; i++) {
        const u64 q = mods[i].q, qmu = mods[i].mu, hm = R->half_mod[i];
        const FwdMod fm = {q, 0 - q, mods[i].q3, (unsigned)mods[i].pad[0]};
        u64 vals[S];
#pragma unroll
        for (int r = 0; r < S; r++) vals[r] = sub_mod(barrett(v[r], q, qmu), hm, q);
        couple_fwd<S>(vals, tw_fwd + ((size_t)i << LOGN), fm);
        u64 *o = tmp + ((size_t)i << LOGN);
#pragma unroll
        for (int r = 0; r < S; r++) o[(size_t)r * M] = vals[r];
    }
}

// ---------------------------------------------------------------------------------------------------
// Chained rescale, pointwise part.  acc slot (c, i) holds the raw block INTT of limb i (i <= from) of poly c.  One thread
// owns the S points n + r*M of the polynomial: limbs are visited from the top; limb i is finished (coupling, N^-1), then every
// step l = from .. i+1 that drops a limb above it is applied, x_i <- (x_i - ((r_l mod q_i) - half_l mod q_i)) q_l^-1 with
// r_l = x_l + half_l mod q_l the rounded residue of the dropped limb (ring.DivRoundByLastModulusNTT); a limb that will itself
// be dropped leaves its r in shared memory, a surviving limb (i <= to) gets the forward coupling levels and goes to mu[c][i]
// for the block NTT.  grid (M/128, 2, jobs), 128 threads, (from - to) * S * 128 words of shared memory.
// ---------------------------------------------------------------------------------------------------
constexpr int CHAIN_THREADS = 128;
template <int S>
__global__ void __launch_bounds__(CHAIN_THREADS, LMN_CHAIN_MINB) k_rs_chain(const ModDev *mods, const ulonglong2 *tw_fwd, const ulonglong2 *tw_inv,
                                                            const RescaleChain *R, const KsJobs J) {
    extern __shared__ __align__(16) u64 sm[];   // rr[l - to - 1][r][thread]
    constexpr int LOGN = RA_LOGM + LogS<S>::v;
    constexpr int M = RA_M;
    const int from = R->from, to = R->to;
    const int c = blockIdx.y, job = blockIdx.z;
    const int n = blockIdx.x * CHAIN_THREADS + threadIdx.x;
    const u64 *raw = J.acc + (size_t)job * J.acc_words + (((size_t)c * (from + 1)) << LOGN) + n;
    u64 *out = J.mu + (size_t)job * J.mu_words + (((size_t)c * (to + 1)) << LOGN) + n;
    // the S raw values of limb i - 1 are requested while limb i is processed: with two warps per scheduler (the dropped limbs' residues
    // fill the shared memory) nothing else hides a global load (ncu r2d: long_scoreboard was the top stall, 1.3 cycles per issue)
    u64 nx[S];
#pragma unroll
    for (int r = 0; r < S; r++) nx[r] = ld_pin(raw + ((size_t)from << LOGN) + (size_t)r * M);
#pragma unroll 1
    for (int i = from; i >= 0; i--) {
        const ModDev mi = mods[i];
        const u64 q = mi.q;
        const FwdMod fm = {q, 0 - q, mi.q3, (unsigned)mi.pad[0]};
        u64 v[S];
#pragma unroll
        for (int r = 0; r < S; r++) v[r] = nx[r];
        if (i > 0) {
#pragma unroll
            for (int r = 0; r < S; r++) nx[r] = ld_pin(raw + ((size_t)(i - 1) << LOGN) + (size_t)r * M);
        }
        couple_inv<S>(v, tw_inv + ((size_t)i << LOGN), fm);
#pragma unroll
        for (int r = 0; r < S; r++) v[r] = mul_shoup(v[r], mi.ninv, mi.ninvs, q);
        const int lmin = i > to ? i + 1 : to + 1;
#pragma unroll 1
        for (int l = from; l >= lmin; l--) {
            // x_i <- (x_i - ((r_l mod q_i) - (half_l mod q_i))) q_l^-1 = (x_i + cadd - r_l) q_l^-1 mod q_i: r_l is not reduced first
            // (the 3-product Shoup estimate takes any 64-bit value to [0, 3q)), and x_i stays lazy in [0, 3q) from step to step
            const u64 ca = R->cadd[l][i], w = R->qlinv_w[l][i], ws = R->qlinv_ws[l][i];
            const u64 *rr = sm + ((size_t)(l - to - 1) * S) * CHAIN_THREADS + threadIdx.x;
#pragma unroll
            for (int r = 0; r < S; r++) v[r] = mul_shoup_approx(v[r] + ca - rr[r * CHAIN_THREADS], w, ws, fm.nq, fm.z);
        }
        if (lmin <= from) {
#pragma unroll
            for (int r = 0; r < S; r++) v[r] = csub(csub(v[r], 2 * q), q);
        }
        if (i > to) {
            const u64 half = R->half[i];
            u64 *rr = sm + ((size_t)(i - to - 1) * S) * CHAIN_THREADS + threadIdx.x;
#pragma unroll
            for (int r = 0; r < S; r++) rr[r * CHAIN_THREADS] = add_mod(v[r], half, q);
        } else {
            couple_fwd<S>(v, tw_fwd + ((size_t)i << LOGN), fm);
#pragma unroll
            for (int r = 0; r < S; r++) out[((size_t)i << LOGN) + (size_t)r * M] = v[r];
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// block NTT of mu[c][i] + epilogue.  grid (S*nlimbs, 2, ceil(jobs/jb)); one CTA owns block h of limb i of poly c for jb jobs.
//   MODE 0  u = (acc - ext) P^-1 (k_mac has already added c0 P to acc_0), then AutomorphismHoisted + Add: col2 = col + pi(u).
//           The permutation out[n] = in[index(n)] of the NTT domain maps blocks onto blocks (the top log2(S) bits of
//           index(n) depend only on those of n), so the CTA stages its block of u in shared memory and gathers from it
//           while it adds the one destination block of col and stores it to col2 (coalesced, out of place).
//   MODE 1  col2 = u  (no automorphism: the ring-switch key switch)
//   MODE 2  Rescale: col_i = (col_i - ext) q_l^-1, in place
//   MODE 3  col_i = NTT(mu[c][i]) (canonical), no epilogue operands: the tail of the chained rescale
// All operands are 16 KB blocks that arrive by bulk copies: the transform's input of job j+1, and the epilogue's operands of
// job j (acc block; destination block of col), requested right after the one barrier of job j that proves the landing
// buffers free, so they land while the CTA transforms.
// ---------------------------------------------------------------------------------------------------
template <int S, int MODE>
__global__ void __launch_bounds__(RA_THREADS, LMN_BLKFWD_MINB) k_blk_fwd(const ModDev *mods, const ulonglong2 *twb_fwd, const KsConsts *C,
                                                            const RescaleConsts *R, const KsJobs J, uint32_t gal_el, int nlimbs,
                                                            int k, int np, int jobs, int jb) {
    extern __shared__ __align__(16) u64 sm[];
    constexpr int LOGN = RA_LOGM + LogS<S>::v;
    constexpr int M = RA_M;
    ulonglong2 *stw = reinterpret_cast<ulonglong2 *>(sm);
    u64 *buf = sm + 2 * RA_M;
    u64 *landD = buf + RA_WORDS;      // transform input of the next job (words u + 256 r)
    u64 *landA = landD + RA_M;        // acc block (MODE 2: the col block), natural order
    u64 *landX = landA + RA_M;        // MODE 0: destination block of col, natural order
    u64 *bars = landX + RA_M;         // [0] input landed, [1] epilogue operands landed, [2] table landed
    const int i = blockIdx.x / S, h = blockIdx.x % S, c = blockIdx.y;
    const ModDev md = mods[i];
    const u64 q = md.q;
    const FwdMod fm = {q, 0 - q, md.q3, (unsigned)md.pad[0]};
    const unsigned u = threadIdx.x;
    const int KP = k + np;   // k = limb stride of col (Q limbs of the context), KP = limb stride of acc
    const u64 w = MODE == 3 ? 0 : MODE == 2 ? R->qlinv_w[i] : C->pinv_w[i], ws = MODE == 3 ? 0 : MODE == 2 ? R->qlinv_ws[i] : C->pinv_ws[i];
    const size_t blk = (size_t)h * M;
    const u64 qK = (u64)(3 * LOGN + 2) * q;
    int hd = 0;   // destination block of the automorphism
    if (MODE == 0) {
#pragma unroll
        for (int hp = 0; hp < S; hp++)
            if ((int)(auto_index((unsigned)hp * M, gal_el, LOGN) >> RA_LOGM) == h) hd = hp;
    }
    const int j0 = blockIdx.z * jb, j1 = min(jobs, j0 + jb);
    const size_t d_off = (((size_t)c * nlimbs + i) << LOGN) + blk;
    const size_t col_off = ((size_t)c * k + i) << LOGN;
    const size_t a_off = MODE == 2 ? col_off + blk : (((size_t)c * KP + i) << LOGN) + blk;
    const u64 *a_base = MODE == 2 ? J.col : J.acc;
    const size_t a_words = MODE == 2 ? J.col_words : J.acc_words;
    if (u == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); mbar_init_fence(); }
    __syncthreads();
    if (u == 0) {
        mbar_expect_tx(&bars[2], 2 * BLK_BYTES);
        bulk_g2s(stw, twb_fwd + (((size_t)i * S + h) << RA_LOGM), 2 * BLK_BYTES, &bars[2]);
        mbar_expect_tx(&bars[0], BLK_BYTES);
        bulk_g2s(landD, J.mu + (size_t)j0 * J.mu_words + d_off, BLK_BYTES, &bars[0]);
    }
    mbar_wait(&bars[2], 0);
    for (int job = j0; job < j1; job++) {
        const unsigned ph = (unsigned)(job - j0) & 1u;
        u64 a[8];
        mbar_wait(&bars[0], ph);
#pragma unroll
        for (int r = 0; r < 8; r++) a[r] = landD[u + 256 * r];
        ntt2048_fwd(a, buf, stw, fm, [&]() {
            release_smem_to_async_proxy();
            __syncthreads();   // every thread holds its input and is past the previous job's epilogue (buf, landA, landX free)
            if (u == 0) {
                if (job + 1 < j1) {
                    mbar_expect_tx(&bars[0], BLK_BYTES);
                    bulk_g2s(landD, J.mu + (size_t)(job + 1) * J.mu_words + d_off, BLK_BYTES, &bars[0]);
                }
                if (MODE != 3) {
                    mbar_expect_tx(&bars[1], MODE == 0 ? 2 * BLK_BYTES : BLK_BYTES);
                    bulk_g2s(landA, a_base + (size_t)job * a_words + a_off, BLK_BYTES, &bars[1]);
                }
                if (MODE == 0) bulk_g2s(landX, J.col + (size_t)job * J.col_words + col_off + (size_t)hd * M, BLK_BYTES, &bars[1]);
            }
        }, []() {});
        // a[r] = point 8u + r of the block, lazy
        u64 *gout = (MODE >= 2 ? J.col : J.col2) + (size_t)job * J.col_words + col_off;
        if (MODE == 3) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
                ulonglong2 v;
                v.x = barrett(a[2 * r], q, md.mu); v.y = barrett(a[2 * r + 1], q, md.mu);
                reinterpret_cast<ulonglong2 *>(gout + blk + 8 * u)[r] = v;
            }
            continue;
        }
        mbar_wait(&bars[1], ph);
        const ulonglong2 *av = reinterpret_cast<const ulonglong2 *>(landA) + 4 * u;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            // (v - a) w mod q without reducing a first: the transform's input is below 2q (k_down_pre / k_rs_pre reduce before their
            // coupling levels), every level adds at most 3q, so a < (3 log N + 2) q =: qK and v + qK - a lies in (0, qK + q), below
            // 2^64 by the modulus check of lmn_ctx_create; the 3-product Shoup estimate takes any 64-bit value to [0, 3q)
            ulonglong2 v = av[r];
            v.x = csub(csub(mul_shoup_approx(v.x + qK - a[2 * r], w, ws, fm.nq, fm.z), 2 * q), q);
            v.y = csub(csub(mul_shoup_approx(v.y + qK - a[2 * r + 1], w, ws, fm.nq, fm.z), 2 * q), q);
            if (MODE != 0) reinterpret_cast<ulonglong2 *>(gout + blk + 8 * u)[r] = v;
            else { buf[PAD(8 * u + 2 * r)] = v.x; buf[PAD(8 * u + 2 * r + 1)] = v.y; }   // the thread's own last-pass slots
        }
        if (MODE == 0) {
            __syncthreads();
            const unsigned nb = (unsigned)hd * M + 8 * u;
            const ulonglong2 *xv = reinterpret_cast<const ulonglong2 *>(landX) + 4 * u;
            ulonglong2 *co = reinterpret_cast<ulonglong2 *>(gout + nb);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                ulonglong2 v = xv[r];
                const unsigned sa = auto_index(nb + 2 * r, gal_el, LOGN) & (M - 1);
                const unsigned sb = auto_index(nb + 2 * r + 1, gal_el, LOGN) & (M - 1);
                v.x = add_mod(v.x, buf[PAD(sa)], q);
                v.y = add_mod(v.y, buf[PAD(sb)], q);
                co[r] = v;
            }
        }
    }
}

// =====================================================================================================
// launchers
// =====================================================================================================
// jobs per CTA: as many as keep at least four waves of CTAs in flight (the twiddle table is filled once per CTA)
static int pick_jb(int ctas_per_job, int jobs) {
    const int want = 4 * 148 * 3;
    static const int jmax = getenv("LMN_BLK_JB") ? atoi(getenv("LMN_BLK_JB")) : 16;   // tuning knob (config D, 1184 jobs: 4 -> 454 ms of k_down per step, 16 -> 450)
    for (int jb = jmax; jb > 1; jb >>= 1)
        if ((long)ctas_per_job * ((jobs + jb - 1) / jb) >= want) return jb;
    return 1;
}
#define LMN_DISPATCH_S(SB, ...)                                  \
    switch (SB) {                                                \
        case 2: { constexpr int SS = 2; __VA_ARGS__; } break;    \
        case 4: { constexpr int SS = 4; __VA_ARGS__; } break;    \
        case 8: { constexpr int SS = 8; __VA_ARGS__; } break;    \
        default: break;                                          \
    }

void launch_blk_intt(const Launch &L, const DeviceTables &T, const BlkSlots &A, int jobs, int kid) {
    LaunchScope scope(L, kid, (uint64_t)jobs);
    const int SB = blk_blocks(T.logn);
    const int jb = pick_jb(SB * A.nslots, jobs);
    const dim3 g(SB * A.nslots, (jobs + jb - 1) / jb);
    LMN_DISPATCH_S(SB, k_blk_intt<SS><<<g, RA_THREADS, BLK_INTT_SMEM, L.stream>>>(T.mods, T.twb_inv, A, jobs, jb));
}

void launch_modup_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J) {
    LaunchScope scope(L, KID_VFIX, (uint64_t)J.jobs);
    const int SB = blk_blocks(T.logn);
    // digits of the common size np run with the digit size as a template parameter; a shorter last digit separately
    int full = 0;
    while (full < Ch.beta && Ch.dig_size[full] == Ch.np) full++;
    const int xb = RA_M / 256;
    if (full) {
        const dim3 g(xb, full, J.jobs);
        if (Ch.np == 2) { LMN_DISPATCH_S(SB, k_modup_blk<SS, 2><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J, 0)); }
        else if (Ch.np == 1) { LMN_DISPATCH_S(SB, k_modup_blk<SS, 1><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J, 0)); }
        else { LMN_DISPATCH_S(SB, k_modup_blk<SS, 0><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J, 0)); }
    }
    if (full < Ch.beta) {
        const dim3 g(xb, Ch.beta - full, J.jobs);
        if (Ch.dig_size[full] == 1 && Ch.beta - full == 1) { LMN_DISPATCH_S(SB, k_modup_blk<SS, 1><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J, full)); }
        else { LMN_DISPATCH_S(SB, k_modup_blk<SS, 0><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J, full)); }
    }
}

void launch_down_pre(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J) {
    LaunchScope scope(L, KID_VPFIX, (uint64_t)J.jobs);
    const int SB = blk_blocks(T.logn);
    const dim3 g(RA_M / 256, 2, J.jobs);
    LMN_DISPATCH_S(SB, k_down_pre<SS><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J));
}

void launch_down_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J,
                     uint32_t gal_el, int mode) {
    LaunchScope scope(L, KID_DOWN, (uint64_t)J.jobs);
    const int SB = blk_blocks(T.logn);
    const int jb = pick_jb(SB * Ch.nq * 2, J.jobs);
    const dim3 g(SB * Ch.nq, 2, (J.jobs + jb - 1) / jb);
    if (mode == 1) { LMN_DISPATCH_S(SB, k_blk_fwd<SS, 1><<<g, RA_THREADS, BLK_FWD_SMEM, L.stream>>>(T.mods, T.twb_fwd, Cd, nullptr, J, gal_el, Ch.nq, T.k, T.np, J.jobs, jb)); }
    else { LMN_DISPATCH_S(SB, k_blk_fwd<SS, 0><<<g, RA_THREADS, BLK_FWD_SMEM, L.stream>>>(T.mods, T.twb_fwd, Cd, nullptr, J, gal_el, Ch.nq, T.k, T.np, J.jobs, jb)); }
}

void launch_rs_pre(const Launch &L, const DeviceTables &T, const RescaleConsts *Rd, int level, const KsJobs &J) {
    LaunchScope scope(L, KID_RS_PRE, (uint64_t)J.jobs);
    const int SB = blk_blocks(T.logn);
    const dim3 g(RA_M / 256, 2, J.jobs);
    LMN_DISPATCH_S(SB, k_rs_pre<SS><<<g, 256, 0, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Rd, J, level));
}

void launch_rs_blk(const Launch &L, const DeviceTables &T, const RescaleConsts *Rd, int level, const KsJobs &J) {
    LaunchScope scope(L, KID_RS_APPLY, (uint64_t)J.jobs);
    const int SB = blk_blocks(T.logn);
    const int jb = pick_jb(SB * level * 2, J.jobs);
    const dim3 g(SB * level, 2, (J.jobs + jb - 1) / jb);
    LMN_DISPATCH_S(SB, k_blk_fwd<SS, 2><<<g, RA_THREADS, BLK_FWD_SMEM, L.stream>>>(T.mods, T.twb_fwd, nullptr, Rd, J, 0u, level, T.k, T.np, J.jobs, jb));
}

void launch_rescale_chain(const Launch &L, const DeviceTables &T, const RescaleChain *Cd, const RescaleChain &Ch, const KsJobs &J,
                          const u64 *src, uint64_t src_job_words) {
    const int SB = blk_blocks(T.logn);
    const int nl = Ch.from + 1;
    BlkSlots A;
    // only the first transform reads the input ciphertexts: they may stay where they are (a store), the results go to J.col
    A.src = src ? src : J.col; A.src_job_words = src ? src_job_words : J.col_words; A.dst = J.acc; A.dst_job_words = J.acc_words; A.nslots = 2 * nl;
    for (int c = 0; c < 2; c++)
        for (int i = 0; i < nl; i++) { A.mod[c * nl + i] = i; A.src_limb[c * nl + i] = (uint32_t)(c * T.k + i); A.dst_limb[c * nl + i] = (uint32_t)(c * nl + i); }
    launch_blk_intt(L, T, A, J.jobs, KID_RS_INTT);
    {
        LaunchScope scope(L, KID_RS_PRE, (uint64_t)J.jobs);
        const dim3 g(RA_M / CHAIN_THREADS, 2, J.jobs);
        const size_t smem = (size_t)(Ch.from - Ch.to) * SB * CHAIN_THREADS * 8;
        LMN_DISPATCH_S(SB, k_rs_chain<SS><<<g, CHAIN_THREADS, smem, L.stream>>>(T.mods, T.tw_fwd, T.tw_inv, Cd, J));
    }
    {
        LaunchScope scope(L, KID_RS_APPLY, (uint64_t)J.jobs);
        const int keep = Ch.to + 1;
        const int jb = pick_jb(SB * keep * 2, J.jobs);
        const dim3 g(SB * keep, 2, (J.jobs + jb - 1) / jb);
        LMN_DISPATCH_S(SB, k_blk_fwd<SS, 3><<<g, RA_THREADS, BLK_FWD_SMEM, L.stream>>>(T.mods, T.twb_fwd, nullptr, nullptr, J, 0u, keep, T.k, T.np, J.jobs, jb));
    }
}

const char *kernels_blk_init() {
    cudaError_t e = cudaSuccess;
#define SETSM(K) do { if (e == cudaSuccess) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BLK_FWD_SMEM); } while (0)
#define SETS(SS) SETSM(k_blk_intt<SS>); SETSM((k_blk_fwd<SS, 0>)); SETSM((k_blk_fwd<SS, 1>)); SETSM((k_blk_fwd<SS, 2>)); SETSM((k_blk_fwd<SS, 3>)); SETSM(k_rs_chain<SS>);
    SETS(2) SETS(4) SETS(8)
#undef SETS
#undef SETSM
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace lmn
