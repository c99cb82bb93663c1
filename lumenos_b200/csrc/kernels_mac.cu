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
#include "ntt_blk.cuh"
#include "tma.cuh"

#include <cstdlib>

namespace lmn {

template <int S> struct LogS2 { static constexpr int v = S == 8 ? 3 : S == 4 ? 2 : S == 2 ? 1 : 0; };

// ---------------------------------------------------------------------------------------------------
// TMA-fed variant.  One CTA owns (target modulus, block h) for jb consecutive jobs; its work is a sequence of items
// (job, digit).  Thread 0 requests with bulk copies (cp.async.bulk, completion on an mbarrier) the 16 KB of residues of
// item i+1 and the 2 x 16 KB of key rows of item i right after the one CTA barrier of item i that proves the landing
// buffers free; the copies land while the other warps transform.  Shared memory: 32 KB twiddles + 17 KB transposition
// buffer + 16 KB residues + 32 KB key rows = 97 KB, two CTAs per SM.
// ---------------------------------------------------------------------------------------------------
#ifndef LMN_MACRD_MINB
#define LMN_MACRD_MINB 2
#endif
typedef unsigned __int128 u128;
constexpr size_t MAC_RD_SMEM = (2 * (size_t)RA_M + (size_t)RA_WORDS + (size_t)RA_M + 2 * (size_t)RA_M + 4) * 8;
template <int S>
__global__ void __launch_bounds__(RA_THREADS, LMN_MACRD_MINB) k_mac_rd(const ModDev *mods, const ulonglong2 *twb_fwd, const KsConsts *C,
                                                           const KsJobs J, const u64 *__restrict__ evk, int jb) {
    extern __shared__ __align__(16) u64 sm[];
    constexpr int M = RA_M;
    constexpr int LOGN = RA_LOGM + LogS2<S>::v;
    constexpr unsigned BLK_BYTES = M * 8;
    ulonglong2 *stw = reinterpret_cast<ulonglong2 *>(sm);
    u64 *buf = sm + 2 * RA_M;
    u64 *stage = buf + RA_WORDS;          // residues of the next item: words u + 256 r (own limb: 8u + r)
    u64 *keys = stage + RA_M;             // [2][2048] key rows of the current item, block read order
    u64 *bars = keys + 2 * RA_M;          // [0] residues landed, [1] key rows landed, [2] twiddle table landed
    const int tix = blockIdx.x / S, h = blockIdx.x % S;
    const int nq = C->nq, k = C->k, KP = k + C->np;
    const int gm = ks_mod_index(tix, nq, k);
    const ModDev md = mods[gm];
    const u64 p = md.q;
    const FwdMod fm = {p, 0 - p, md.q3, (unsigned)md.pad[0]};
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
    if (u == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); mbar_init_fence(); }
    __syncthreads();   // barriers visible
    if (u == 0) {
        mbar_expect_tx(&bars[2], 2 * BLK_BYTES);
        bulk_g2s(stw, twb_fwd + (((size_t)gm * S + h) << RA_LOGM), 2 * BLK_BYTES, &bars[2]);
        issue_data(j0, 0);
    }

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
    // (acc_0 - ext) P^-1 + c0 = (acc_0 + c0 P - ext) P^-1: the "+ c0" of the key switch is folded into component 0 here, so
    // that k_blk_fwd needs no block of c0
    const bool addc0 = tix < nq;
    const u64 pw = addc0 ? C->pmod_w[tix] : 0, pws = addc0 ? C->pmod_ws[tix] : 0;
    mbar_wait(&bars[2], 0);
    int d = 0, job = j0;
    for (int it = 0; it < nit; it++) {
        u64 a[8];
        mbar_wait(&bars[0], it & 1);
        auto after_read = [&]() {
            // The bulk copies requested below overwrite `stage` and `keys` through the async proxy: every thread's reads of them
            // must be complete (not merely issued) and ordered before it (measured: without this a copy occasionally lands
            // before a queued LDS of the own-limb path has executed).
            release_smem_to_async_proxy();
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
            ulonglong2 c0[4];
            if (addc0) {
                const ulonglong2 *cp = reinterpret_cast<const ulonglong2 *>(J.col + (size_t)job * J.col_words + ((size_t)tix << LOGN) + blk + 8 * u);
#pragma unroll
                for (int r = 0; r < 4; r++) c0[r] = ld_pin2(cp + r);   // requested before the reductions that hide their latency
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                ulonglong2 r0, r1;
                r0.x = fin(acc0[2 * r]); r0.y = fin(acc0[2 * r + 1]);
                r1.x = fin(acc1[2 * r]); r1.y = fin(acc1[2 * r + 1]);
                if (addc0) {
                    r0.x = add_mod(r0.x, mul_shoup(c0[r].x, pw, pws, p), p);
                    r0.y = add_mod(r0.y, mul_shoup(c0[r].y, pw, pws, p), p);
                }
                o0[r] = r0; o1[r] = r1;
                acc0[2 * r] = 0; acc0[2 * r + 1] = 0; acc1[2 * r] = 0; acc1[2 * r + 1] = 0;
            }
            d = 0; job++;
        }
    }
}

// =====================================================================================================
void launch_mac_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, const KsJobs &J, const u64 *evk) {
    LaunchScope scope(L, KID_MAC, (uint64_t)J.jobs);
    const int SB = blk_blocks(T.logn);
    static const int jbe = getenv("LMN_MAC_JB") ? atoi(getenv("LMN_MAC_JB")) : 0;
    const int ctas = (Ch.nq + Ch.np) * SB;
    // jobs per CTA: the twiddle table and the pipeline start-up are paid once per CTA; as many as leave eight waves of CTAs
    // (measured at config D, 1184 jobs per launch: 1 -> 1296 ms per step, 2 -> 1270, 4 -> 1247, 8 -> 1243)
    int jb = 1;
    if (jbe > 0) jb = jbe;
    else for (int c = 8; c > 1; c >>= 1) if ((long)ctas * ((J.jobs + c - 1) / c) >= 8 * 148 * 2) { jb = c; break; }
    const dim3 g(ctas, (J.jobs + jb - 1) / jb);
    if (SB == 8) k_mac_rd<8><<<g, RA_THREADS, MAC_RD_SMEM, L.stream>>>(T.mods, T.twb_fwd, Cd, J, evk, jb);
    else if (SB == 4) k_mac_rd<4><<<g, RA_THREADS, MAC_RD_SMEM, L.stream>>>(T.mods, T.twb_fwd, Cd, J, evk, jb);
    else k_mac_rd<2><<<g, RA_THREADS, MAC_RD_SMEM, L.stream>>>(T.mods, T.twb_fwd, Cd, J, evk, jb);
}

const char *kernels_mac_init() {
    cudaError_t e = cudaSuccess;
    const int maxsm = 227 * 1024;
#define SETSM(K) do { if (e == cudaSuccess) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, maxsm); } while (0)
    SETSM(k_mac_rd<2>); SETSM(k_mac_rd<4>); SETSM(k_mac_rd<8>);
#undef SETSM
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace lmn
