// Device-side 64-bit modular arithmetic and the shared-memory negacyclic NTT / INTT used by every
// integer-bound kernel.  sm_100a: the 64x64 products run on the IMAD pipe (IMAD.WIDE.U32 chains), everything
// else on the ALU pipe; there is no tensor-core formulation of this (modular integer) work.
//
// Moduli are < 2^59 (58/56-bit Q chain, 55-bit P; fhe/bfv.go:160-178).  The forward transform is a Harvey
// lazy Cooley-Tukey network: with B_0 = 1 the bound on values grows by 2q per level, (1 + 2*16) q < 2^64, so no
// correction is needed inside the transform for log N <= 16 and a single Barrett step makes the result
// canonical.  Twiddles are (w, floor(w*2^64/q)) pairs in Lattigo's table order RootsForward[bitrev(j)] = psi^j.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lmn {
typedef unsigned long long u64;

struct ModDev {
    u64 q;
    u64 qinv;   // q^-1 mod 2^64 (Montgomery)
    u64 mu;     // floor(2^64 / q)  (single-word Barrett for x < 2^64)
    u64 ninv;   // N^-1 mod q
    u64 ninvs;  // Shoup companion of ninv
    u64 q3;     // 3q (lazy forward butterfly with the 3-product quotient estimate, ntt_r16.cuh)
    u64 pad[2]; // pad[0] stays 0: kernels read it as the "opaque zero" operand of FwdMod::z (ntt_blk.cuh)
};

__device__ __forceinline__ u64 csub(u64 x, u64 q) { return x >= q ? x - q : x; }
__device__ __forceinline__ u64 add_mod(u64 a, u64 b, u64 q) { return csub(a + b, q); }
__device__ __forceinline__ u64 sub_mod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
// x * w mod q, lazy: result in [0, 2q) for any x < 2^64
__device__ __forceinline__ u64 mul_shoup_lazy(u64 x, u64 w, u64 ws, u64 q) {
    u64 hi = __umul64hi(x, ws);
    return x * w - hi * q;
}
__device__ __forceinline__ u64 mul_shoup(u64 x, u64 w, u64 ws, u64 q) { return csub(mul_shoup_lazy(x, w, ws, q), q); }
// (hi:lo) * 2^-64 mod q for hi:lo < q * 2^64; result in (0, 2q)
__device__ __forceinline__ u64 mont_reduce_lazy(u64 lo, u64 hi, u64 q, u64 qinv) {
    u64 m = lo * qinv;
    u64 h = __umul64hi(m, q);
    return hi - h + q;
}
// a * b_mont * 2^-64 mod q, a < 2^64, b_mont < q; lazy (0, 2q)
__device__ __forceinline__ u64 mont_mul_lazy(u64 a, u64 bm, u64 q, u64 qinv) {
    return mont_reduce_lazy(a * bm, __umul64hi(a, bm), q, qinv);
}
__device__ __forceinline__ u64 mont_mul(u64 a, u64 bm, u64 q, u64 qinv) { return csub(mont_mul_lazy(a, bm, q, qinv), q); }
// x mod q for any x < 2^64
__device__ __forceinline__ u64 barrett(u64 x, u64 q, u64 mu) {
    u64 qh = __umul64hi(x, mu);
    return csub(x - qh * q, q);
}
// 128-bit accumulate
__device__ __forceinline__ void mac128(u64 &lo, u64 &hi, u64 a, u64 b) {
    u64 pl = a * b, ph = __umul64hi(a, b);
    lo += pl;
    hi += ph + (lo < pl);
}

// shared-memory index padding: one spare word every 16 keeps the 64-bit accesses of the contiguous
// (stride-1/2/4) radix-8 passes at the 2-wavefront minimum
__device__ __forceinline__ int PAD(int i) { return i + (i >> 4); }
__host__ __device__ constexpr int padded_words(int n) { return n + (n >> 4) + 1; }

// ---------------------------------------------------------------------------------------------------
// forward NTT of M = 2^LOGM words held in shared memory (padded), values < 2^64 - 2*LOGM*2q on entry.
// SH = S + h selects the sub-transform h of an S-way split of a length S*M transform whose first log2(S)
// levels were already applied: the twiddle of local level m, local group i is tw[m*SH + i].
// Output values are lazy (< input bound + 2*LOGM*q); callers reduce with barrett().
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bfly_ct(u64 &x, u64 &y, const ulonglong2 w, u64 q, u64 q2) {
    u64 t = mul_shoup_lazy(y, w.x, w.y, q);
    y = x - t + q2;
    x = x + t;
}

template <int LOGM>
__device__ __forceinline__ void ntt_fwd_smem(u64 *s, const ulonglong2 *__restrict__ tw, u64 q, int SH) {
    constexpr int M = 1 << LOGM;
    const u64 q2 = 2 * q;
    int m = 1;  // local number of groups at the current level
    constexpr int REM = LOGM % 3;
    if (REM == 1) {
        const int t = M >> 1;
        for (int u = threadIdx.x; u < M / 2; u += blockDim.x) {
            ulonglong2 w = __ldg(&tw[SH]);
            u64 a = s[PAD(u)], b = s[PAD(u + t)];
            bfly_ct(a, b, w, q, q2);
            s[PAD(u)] = a; s[PAD(u + t)] = b;
        }
        m = 2;
        __syncthreads();
    } else if (REM == 2) {
        const int t2 = M >> 2;
        for (int u = threadIdx.x; u < M / 4; u += blockDim.x) {
            u64 a0 = s[PAD(u)], a1 = s[PAD(u + t2)], a2 = s[PAD(u + 2 * t2)], a3 = s[PAD(u + 3 * t2)];
            ulonglong2 w = __ldg(&tw[SH]);
            bfly_ct(a0, a2, w, q, q2); bfly_ct(a1, a3, w, q, q2);
            ulonglong2 wa = __ldg(&tw[2 * SH]), wb = __ldg(&tw[2 * SH + 1]);
            bfly_ct(a0, a1, wa, q, q2); bfly_ct(a2, a3, wb, q, q2);
            s[PAD(u)] = a0; s[PAD(u + t2)] = a1; s[PAD(u + 2 * t2)] = a2; s[PAD(u + 3 * t2)] = a3;
        }
        m = 4;
        __syncthreads();
    }
#pragma unroll 1
    for (; m < M; m <<= 3) {
        const int t3 = M / (8 * m);  // distance between the 8 points of one radix-8 task
        for (int u = threadIdx.x; u < M / 8; u += blockDim.x) {
            const int i = u / t3, j = u - i * t3;
            const int base = i * 8 * t3 + j;
            u64 a[8];
#pragma unroll
            for (int r = 0; r < 8; r++) a[r] = s[PAD(base + r * t3)];
            ulonglong2 w = __ldg(&tw[m * SH + i]);
#pragma unroll
            for (int r = 0; r < 4; r++) bfly_ct(a[r], a[r + 4], w, q, q2);
            ulonglong2 wa = __ldg(&tw[2 * m * SH + 2 * i]), wb = __ldg(&tw[2 * m * SH + 2 * i + 1]);
            bfly_ct(a[0], a[2], wa, q, q2); bfly_ct(a[1], a[3], wa, q, q2);
            bfly_ct(a[4], a[6], wb, q, q2); bfly_ct(a[5], a[7], wb, q, q2);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                ulonglong2 wc = __ldg(&tw[4 * m * SH + 4 * i + r]);
                bfly_ct(a[2 * r], a[2 * r + 1], wc, q, q2);
            }
#pragma unroll
            for (int r = 0; r < 8; r++) s[PAD(base + r * t3)] = a[r];
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// inverse NTT (Gentleman-Sande) of N = 2^LOGN words in shared memory; entry values in [0, 2q), values stay
// in [0, 2q) between levels (Harvey), exit values in [0, 2q) and NOT yet scaled by N^-1 (callers fold N^-1
// into their own Shoup constant).  Twiddles: RootsBackward[bitrev(j)] = psi^-j.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bfly_gs(u64 &x, u64 &y, const ulonglong2 w, u64 q, u64 q2) {
    u64 u = x + y;
    u64 d = x - y + q2;
    x = csub(u, q2);
    y = mul_shoup_lazy(d, w.x, w.y, q);
}

template <int LOGN>
__device__ __forceinline__ void ntt_inv_smem(u64 *s, const ulonglong2 *__restrict__ tw, u64 q) {
    constexpr int N = 1 << LOGN;
    const u64 q2 = 2 * q;
    int h = N >> 1;  // groups at the current level (N/2, N/4, ..., 1)
    constexpr int REM = LOGN % 3;
    constexpr int FULL = LOGN / 3;
#pragma unroll 1
    for (int pass = 0; pass < FULL; pass++, h >>= 3) {
        const int t = N / (2 * h);  // butterfly span of the first level of this pass
        for (int u = threadIdx.x; u < N / 8; u += blockDim.x) {
            const int i = u / t, j = u - i * t;
            const int base = i * 8 * t + j;
            u64 a[8];
#pragma unroll
            for (int r = 0; r < 8; r++) a[r] = s[PAD(base + r * t)];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                ulonglong2 wc = __ldg(&tw[h + 4 * i + r]);
                bfly_gs(a[2 * r], a[2 * r + 1], wc, q, q2);
            }
            ulonglong2 wa = __ldg(&tw[(h >> 1) + 2 * i]), wb = __ldg(&tw[(h >> 1) + 2 * i + 1]);
            bfly_gs(a[0], a[2], wa, q, q2); bfly_gs(a[1], a[3], wa, q, q2);
            bfly_gs(a[4], a[6], wb, q, q2); bfly_gs(a[5], a[7], wb, q, q2);
            ulonglong2 w = __ldg(&tw[(h >> 2) + i]);
#pragma unroll
            for (int r = 0; r < 4; r++) bfly_gs(a[r], a[r + 4], w, q, q2);
#pragma unroll
            for (int r = 0; r < 8; r++) s[PAD(base + r * t)] = a[r];
        }
        __syncthreads();
    }
    if (REM == 2) {  // levels h = 2, 1 : spans N/4, N/2
        const int t = N >> 2;
        for (int u = threadIdx.x; u < N / 4; u += blockDim.x) {
            u64 a0 = s[PAD(u)], a1 = s[PAD(u + t)], a2 = s[PAD(u + 2 * t)], a3 = s[PAD(u + 3 * t)];
            ulonglong2 wa = __ldg(&tw[2]), wb = __ldg(&tw[3]);
            bfly_gs(a0, a1, wa, q, q2); bfly_gs(a2, a3, wb, q, q2);
            ulonglong2 w = __ldg(&tw[1]);
            bfly_gs(a0, a2, w, q, q2); bfly_gs(a1, a3, w, q, q2);
            s[PAD(u)] = a0; s[PAD(u + t)] = a1; s[PAD(u + 2 * t)] = a2; s[PAD(u + 3 * t)] = a3;
        }
        __syncthreads();
    } else if (REM == 1) {  // level h = 1 : span N/2
        const int t = N >> 1;
        for (int u = threadIdx.x; u < N / 2; u += blockDim.x) {
            u64 a = s[PAD(u)], b = s[PAD(u + t)];
            ulonglong2 w = __ldg(&tw[1]);
            bfly_gs(a, b, w, q, q2);
            s[PAD(u)] = a; s[PAD(u + t)] = b;
        }
        __syncthreads();
    }
}

// automorphism index in the NTT domain (lattigo: ring.AutomorphismNTTIndex): out[i] = in[auto_index(i)]
__device__ __forceinline__ unsigned auto_index(unsigned i, unsigned gal_el, int logn) {
    unsigned mask = (2u << logn) - 1u;
    unsigned t1 = 2u * (__brev(i) >> (32 - logn)) + 1u;
    unsigned t2 = (((gal_el * t1) & mask) - 1u) >> 1;
    return __brev(t2) >> (32 - logn);
}

}  // namespace lmn
