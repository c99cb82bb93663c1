// CUDA kernels (sm_100a) of the lumenos hot path.  See DESIGN.md for the kernel list and rooflines.
//
//   Encode eval      k_encode_pass                      HBM-bound   (fhe/code.go, fhe/ntt.go)
//   Inner product    k_mul_pt, k_intt_c1, k_vfix, k_mac, k_pintt, k_vpfix, k_down   integer-bound
//                    (fhe/ligero.go:319-325 = MulNew + InnerSum; lattigo DecomposeNTT / GadgetProduct / ModDown)
//   Rescale          k_rs_intt, k_rs_apply              integer-bound (fhe/ligero.go:331-333, 271-273)
#include "kernels.cuh"
#include "ntt_blk.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace lmn {

#define LMN_DISPATCH_LOGN(LOGN_RT, ...)                                     \
    switch (LOGN_RT) {                                                      \
        case 6:  { constexpr int LG = 6;  __VA_ARGS__; } break;             \
        case 8:  { constexpr int LG = 8;  __VA_ARGS__; } break;             \
        case 10: { constexpr int LG = 10; __VA_ARGS__; } break;             \
        case 12: { constexpr int LG = 12; __VA_ARGS__; } break;             \
        case 13: { constexpr int LG = 13; __VA_ARGS__; } break;             \
        case 14: { constexpr int LG = 14; __VA_ARGS__; } break;             \
        default: break;                                                     \
    }

static inline int ntt_threads(int m) { int t = m / 8; return t < 32 ? 32 : (t > 1024 ? 1024 : t); }
const char *kernel_name(int id) {
    static const char *names[KID_COUNT] = {"k_encode_pass", "k_mul_pt", "k_intt_c1", "k_modup", "k_mac", "k_pintt", "k_down_pre",
                                           "k_down", "k_rs_intt", "k_rs_pre", "k_rs_apply", "k_pack_out", "k_gather_in", "k_ring_decimate", "k_leaf_serialize", "k_sha256_leaves", "k_sha256_nodes", "other"};
    return id >= 0 && id < KID_COUNT ? names[id] : "?";
}
cudaEvent_t Profiler::get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
LaunchScope::LaunchScope(const Launch &l, int kid, uint64_t jobs) : L(l), id(kid) {
    if (L.counter) (*L.counter)++;
    if (L.prof) {
        L.prof->launches[id]++;
        L.prof->jobs[id] += jobs;
        if (L.prof->mask & (1u << id)) {
            cudaEvent_t e0 = L.prof->get(); e1 = L.prof->get();
            cudaEventRecord(e0, L.stream);
            L.prof->spans[id].emplace_back(e0, e1);
        }
    }
}
LaunchScope::~LaunchScope() { if (e1) cudaEventRecord(e1, L.stream); }

// =====================================================================================================
// elementwise kernels
// =====================================================================================================
__global__ void k_addsub(const ModDev *mods, int nlimbs, int logn, const u64 *__restrict__ a, const u64 *__restrict__ b,
                         u64 *__restrict__ o, int sub) {
    // grid.y = poly*nlimbs + limb ; grid.x covers N/2 (ulonglong2 vectorised)
    const int limb = blockIdx.y % nlimbs;
    const u64 q = mods[limb].q;
    const size_t off = ((size_t)blockIdx.y << logn);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << (logn - 1))) return;
    ulonglong2 x = reinterpret_cast<const ulonglong2 *>(a + off)[i];
    ulonglong2 y = reinterpret_cast<const ulonglong2 *>(b + off)[i];
    ulonglong2 r;
    if (sub) { r.x = sub_mod(x.x, y.x, q); r.y = sub_mod(x.y, y.y, q); }
    else     { r.x = add_mod(x.x, y.x, q); r.y = add_mod(x.y, y.y, q); }
    reinterpret_cast<ulonglong2 *>(o + off)[i] = r;
}
void launch_addsub(const Launch &L, const DeviceTables &T, int level, const u64 *a, const u64 *b, u64 *o, int sub) {
    LaunchScope scope(L, KID_OTHER);
    int nl = level + 1, half = T.n / 2, bs = half < 256 ? half : 256;
    dim3 g((half + bs - 1) / bs, 2 * nl);
    k_addsub<<<g, bs, 0, L.stream>>>(T.mods, nl, T.logn, a, b, o, sub);
}

__global__ void k_mul_scalar(const ModDev *mods, int nlimbs, int logn, const u64 *__restrict__ a,
                             const ulonglong2 *__restrict__ w, u64 *__restrict__ o) {
    const int limb = blockIdx.y % nlimbs;
    const u64 q = mods[limb].q;
    const ulonglong2 ww = w[limb];
    const size_t off = ((size_t)blockIdx.y << logn);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << (logn - 1))) return;
    ulonglong2 x = reinterpret_cast<const ulonglong2 *>(a + off)[i];
    x.x = mul_shoup(x.x, ww.x, ww.y, q);
    x.y = mul_shoup(x.y, ww.x, ww.y, q);
    reinterpret_cast<ulonglong2 *>(o + off)[i] = x;
}
void launch_mul_scalar(const Launch &L, const DeviceTables &T, int level, const u64 *a, const ulonglong2 *w, u64 *o) {
    LaunchScope scope(L, KID_OTHER);
    int nl = level + 1, half = T.n / 2, bs = half < 256 ? half : 256;
    dim3 g((half + bs - 1) / bs, 2 * nl);
    k_mul_scalar<<<g, bs, 0, L.stream>>>(T.mods, nl, T.logn, a, w, o);
}

// ptm = MForm(pt * (T mod q_i))  : folds the factor T of bgv tensorStandard (degree-0 operand) and the Montgomery
// form into the plaintext once per Prove (ref: fhe/ligero.go:204-205,218-219 -> MulNew at :319)
__global__ void k_prep_pt(const ModDev *mods, int logn, const u64 *__restrict__ pt, u64 t_mod, u64 *__restrict__ ptm) {
    const int limb = blockIdx.y;
    const ModDev md = mods[limb];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << logn)) return;
    const size_t off = ((size_t)limb << logn) + i;
    // (pt * T) mod q, then * 2^64 mod q, by 128-bit products reduced with Barrett on the high/low words
    u64 x = pt[off];
    u64 tq = t_mod % md.q;
    // x*tq mod q via Montgomery with R^2: avoid needing R^2 -- use two-step: y = x*tq mod q by 128-bit long reduction
    u64 lo = x * tq, hi = __umul64hi(x, tq);
    // reduce hi:lo mod q  (hi < q since x,tq < q): shift-subtract 64 steps (setup kernel, not on the hot path)
    u64 r = hi % md.q;
    for (int b = 63; b >= 0; b--) {
        r = (r << 1) | ((lo >> b) & 1);   // r < 2q < 2^60
        r = csub(r, md.q);
    }
    // MForm: r * 2^64 mod q
    u64 m = r;
    for (int b = 0; b < 64; b++) m = csub(m << 1, md.q);
    ptm[off] = m;
}
void launch_prep_pt(const Launch &L, const DeviceTables &T, int level, const u64 *pt, u64 t_mod, u64 *ptm) {
    LaunchScope scope(L, KID_OTHER);
    int bs = T.n < 256 ? T.n : 256;
    dim3 g((T.n + bs - 1) / bs, level + 1);
    k_prep_pt<<<g, bs, 0, L.stream>>>(T.mods, T.logn, pt, t_mod, ptm);
}

// ct x pt for one or two plaintexts, reading the matrix ciphertext once
__global__ void k_mul_pt(const ModDev *mods, int nlimbs, int logn, const u64 *__restrict__ in, uint64_t in_words,
                         const u64 *__restrict__ ptm0, const u64 *__restrict__ ptm1, u64 *__restrict__ out0,
                         u64 *__restrict__ out1, uint64_t col_words, int kstride) {
    // grid: x over N/2, y = poly*nlimbs + limb, z = column
    const int poly = blockIdx.y / nlimbs, limb = blockIdx.y % nlimbs;
    const ModDev md = mods[limb];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << (logn - 1))) return;
    const size_t src = (size_t)blockIdx.z * in_words + (((size_t)poly * nlimbs + limb) << logn);
    const size_t dst = (size_t)blockIdx.z * col_words + (((size_t)poly * kstride + limb) << logn);
    ulonglong2 x = reinterpret_cast<const ulonglong2 *>(in + src)[i];
    ulonglong2 p0 = reinterpret_cast<const ulonglong2 *>(ptm0 + ((size_t)limb << logn))[i];
    ulonglong2 r;
    r.x = mont_mul(x.x, p0.x, md.q, md.qinv); r.y = mont_mul(x.y, p0.y, md.q, md.qinv);
    reinterpret_cast<ulonglong2 *>(out0 + dst)[i] = r;
    if (ptm1) {
        ulonglong2 p1 = reinterpret_cast<const ulonglong2 *>(ptm1 + ((size_t)limb << logn))[i];
        r.x = mont_mul(x.x, p1.x, md.q, md.qinv); r.y = mont_mul(x.y, p1.y, md.q, md.qinv);
        reinterpret_cast<ulonglong2 *>(out1 + dst)[i] = r;
    }
}
void launch_mul_pt(const Launch &L, const DeviceTables &T, int level, const u64 *in, uint64_t in_words, int ncols,
                   const u64 *ptm0, const u64 *ptm1, u64 *out0, u64 *out1, uint64_t col_words) {
    LaunchScope scope(L, KID_MULPT);
    int nl = level + 1, half = T.n / 2, bs = half < 256 ? half : 256;
    dim3 g((half + bs - 1) / bs, 2 * nl, ncols);
    // the job col keeps the context-wide limb stride T.k when it is a workspace col, or nl when it is a store entry
    int kstride = (int)(col_words / 2 >> T.logn);
    k_mul_pt<<<g, bs, 0, L.stream>>>(T.mods, nl, T.logn, in, in_words, ptm0, ptm1, out0, out1, col_words, kstride);
}

// key rows (digit d, component c, modulus gm) times a per-(d, gm) constant given in Montgomery form (0 = copy the row), out
// of place.  For the block pipeline (log N >= 12) the words of every 2048-block are stored in the order the block kernels
// read them: thread u of 256 owns words 8u..8u+7 and reads them as four 16-byte pairs j = 0..3, so pair (u, j) goes to
// slot j*256 + u - a warp then reads 512 contiguous bytes per j, from global or shared memory alike.
__host__ __device__ inline uint32_t key_slot(uint32_t i, int permute) {
    if (!permute) return i;
    const uint32_t e = i & 2047u, u = e >> 3, r = e & 7u;
    return (i & ~2047u) | ((((r >> 1) * 256u + u) << 1) | (r & 1u));
}
__global__ void k_scale_rows(const u64 *__restrict__ in, u64 *__restrict__ out, const ModDev *mods, const u64 *__restrict__ fm, int KP,
                             int logn, int permute) {
    const int row = blockIdx.y, gm = row % KP, d = row / (2 * KP);
    const u64 f = fm[d * KP + gm];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << logn)) return;
    const ModDev md = mods[gm];
    const u64 x = in[((size_t)row << logn) + i];
    out[((size_t)row << logn) + key_slot((uint32_t)i, permute)] = f ? mont_mul(x, f, md.q, md.qinv) : x;
}
void launch_scale_rows(const Launch &L, const DeviceTables &T, const u64 *in, u64 *evk, const u64 *fm_dev, int beta) {
    LaunchScope scope(L, KID_OTHER);
    int bs = T.n < 256 ? T.n : 256;
    const int KP = T.k + T.np;
    k_scale_rows<<<dim3((T.n + bs - 1) / bs, beta * 2 * KP), bs, 0, L.stream>>>(in, evk, T.mods, fm_dev, KP, T.logn, blk_blocks(T.logn) > 0);
}

// =====================================================================================================
// single-limb NTT / INTT (test hook; the production kernels fuse these transforms with their neighbours)
// =====================================================================================================
template <int LOGN>
__global__ void __launch_bounds__(1024) k_limb_ntt(const ModDev *mods, const ulonglong2 *tw, int mod_idx, u64 *limb, int inverse) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const ModDev md = mods[mod_idx];
    const ulonglong2 *t = tw + ((size_t)mod_idx << LOGN);
    for (int i = threadIdx.x; i < N; i += blockDim.x) sm[PAD(i)] = limb[i];
    __syncthreads();
    if (!inverse) {
        ntt_fwd_smem<LOGN>(sm, t, md.q, 1);
        for (int i = threadIdx.x; i < N; i += blockDim.x) limb[i] = barrett(sm[PAD(i)], md.q, md.mu);
    } else {
        ntt_inv_smem<LOGN>(sm, t, md.q);
        for (int i = threadIdx.x; i < N; i += blockDim.x) limb[i] = mul_shoup(sm[PAD(i)], md.ninv, md.ninvs, md.q);
    }
}
void launch_limb_ntt(const Launch &L, const DeviceTables &T, int mod_idx, u64 *limb, int inverse) {
    LaunchScope scope(L, KID_OTHER);
    size_t smem = (size_t)padded_words(T.n) * 8;
    LMN_DISPATCH_LOGN(T.logn, k_limb_ntt<LG><<<1, ntt_threads(T.n), smem, L.stream>>>(
                                  T.mods, inverse ? T.tw_inv : T.tw_fwd, mod_idx, limb, inverse));
}

// =====================================================================================================
// Encode eval: one pass of the compiled butterfly plan over a tile of ciphertext slots x 32 coefficients
// =====================================================================================================
__global__ void __launch_bounds__(256) k_encode_pass(const EncPassArgs A) {
    extern __shared__ u64 sm[];  // [tile][ENC_C]
    const int tile_slots = 1 << A.log_tile;
    const uint32_t tile = blockIdx.y;
    const uint64_t lane0 = A.lane_off + (uint64_t)blockIdx.x * ENC_C;       // first lane (word offset inside a ciphertext)
    const int limb = (int)((lane0 >> A.logn) % A.k);
    const int lane = threadIdx.x & (ENC_C - 1);
    const int row0 = threadIdx.x / ENC_C, rows_per_it = blockDim.x / ENC_C;
    const uint32_t slot_base = tile * A.tile_mul;

    // ---- load (slots >= in_count read the shared zero ciphertext: fhe/code.go:24-26) ----
    for (int l = row0; l < tile_slots; l += rows_per_it) {
        const uint32_t slot = slot_base + (uint32_t)l * A.stride;
        const u64 *src;
        if (A.in_by_pos) src = A.in + (uint64_t)A.out_pos[slot] * A.ct_words;
        else src = (slot < A.in_count) ? A.in + (uint64_t)slot * A.ct_words : A.zero;
        sm[l * ENC_C + lane] = src[lane0 + lane];
    }
    __syncthreads();

    // ---- butterfly levels ----
    const ulonglong2 *scal = A.scal + (size_t)limb * A.nscal;
    const u64 qq = A.mods[limb].q;
    for (int s = A.s0; s < A.s1; s++) {
        const int lb = (int)((A.lbits >> (4 * (s - A.s0))) & 15);
        const uint16_t *sc = A.scale_after + (size_t)s * A.n;
        for (int p = row0; p < tile_slots / 2; p += rows_per_it) {
            const int lo = ((p >> lb) << (lb + 1)) | (p & ((1 << lb) - 1));
            const int hi = lo | (1 << lb);
            u64 a = sm[lo * ENC_C + lane], b = sm[hi * ENC_C + lane];
            u64 su = add_mod(a, b, qq), di = sub_mod(a, b, qq);
            const uint32_t ia = sc[slot_base + (uint32_t)lo * A.stride], ib = sc[slot_base + (uint32_t)hi * A.stride];
            if (ia) { ulonglong2 w = __ldg(&scal[ia]); su = mul_shoup(su, w.x, w.y, qq); }
            if (ib) { ulonglong2 w = __ldg(&scal[ib]); di = mul_shoup(di, w.x, w.y, qq); }
            sm[lo * ENC_C + lane] = su; sm[hi * ENC_C + lane] = di;
        }
        __syncthreads();
    }

    // ---- store to the final position of each slot ----
    for (int l = row0; l < tile_slots; l += rows_per_it) {
        const uint32_t slot = slot_base + (uint32_t)l * A.stride;
        A.out[(uint64_t)A.out_pos[slot] * A.ct_words + lane0 + lane] = sm[l * ENC_C + lane];
    }
}
// Register-resident variant for tiles of T = 8 V slots (V = 4 .. 32): the CTA's 8 warps split the tile, a warp's 32 threads are the
// 32 coefficients of the chunk (every global access is one 256 B row), and each thread carries V slots through up to log2 V
// butterfly levels per stage without touching shared memory; between stages the tile is re-dealt through shared memory so
// that the next stage's index bits are register-local.  The scalars that follow the butterflies depend on (level, slot) only:
// the CTA gathers its [levels][T] table of Shoup pairs once and reuses it for all its chunks (same limb).  Results are canonical
// after every operation, exactly as in the tile kernel above.
// x - q if x >= q else x, as a branch-free unsigned minimum (x - q wraps above x when x < q)
__device__ __forceinline__ u64 csub_min(u64 x, u64 q) { const u64 t = x - q; return t < x ? t : x; }
__device__ __forceinline__ int nth_unused_bit(unsigned used, int nth, int nbits) {
    int r = 0, c = 0;
    for (int b = 0; b < nbits; b++)
        if (!((used >> b) & 1u)) { if (c == nth) r = b; c++; }
    return r;
}
#ifndef LMN_ENC3_MINB
#define LMN_ENC3_MINB 4   // measured at config D: 3 -> 54.5 ms per Encode, 4 -> 53.6 ms, 2 -> 59.1 ms (pass kernels, both passes)
#endif
#ifndef LMN_ENC4_MINB
#define LMN_ENC4_MINB 2   // 1 -> 70.2 ms, 3 -> 62.2 ms
#endif
template <int LOGV>
__global__ void __launch_bounds__(256, (LOGV <= 2 ? 3 : LOGV == 3 ? LMN_ENC3_MINB : LOGV == 4 ? LMN_ENC4_MINB : 1)) k_encode_pass_r(const EncPassArgs A, int chunks_per_cta) {
    constexpr int V = 1 << LOGV, LT = LOGV + 3, T = 1 << LT;
    constexpr int NST = (LT + LOGV - 1) / LOGV;
    extern __shared__ __align__(16) u64 sm[];
    ulonglong2 *stab = reinterpret_cast<ulonglong2 *>(sm);     // [LT][T]
    u64 *xbuf = sm + 2 * LT * T;                                // [T][ENC_C]
    const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t slot_base = blockIdx.y * A.tile_mul;
    uint64_t lane0 = A.lane_off + (uint64_t)blockIdx.x * chunks_per_cta * ENC_C;
    const int limb = (int)((lane0 >> A.logn) % A.k);
    const u64 q = A.mods[limb].q, q2 = 2 * q, q3 = 3 * q, nq = 0 - q;
    {
        const ulonglong2 *scal = A.scal + (size_t)limb * A.nscal;
        for (int e = tid; e < LT * T; e += 256) {
            const int lv = e >> LT, l = e & (T - 1);
            const uint32_t id = A.scale_after[(size_t)(A.s0 + lv) * A.n + slot_base + (uint32_t)l * A.stride];
            stab[e] = id ? __ldg(&scal[id]) : make_ulonglong2(0, 0);
        }
    }
    // per stage: off[t] = 1 << (local index bit carried by register-index bit t): the stage's level bits first, then filler bits
    // (lowest unused) when the stage has fewer than log2 V levels; wbase = the warp's 3 remaining bits
    int off[NST][LOGV], wbase[NST];
#pragma unroll
    for (int st = 0; st < NST; st++) {
        const int m = (LT - st * LOGV) < LOGV ? (LT - st * LOGV) : LOGV;
        unsigned used = 0;
#pragma unroll
        for (int t = 0; t < LOGV; t++) {
            if (t < m) {
                const int b = (int)((A.lbits >> (4 * (st * LOGV + t))) & 15);
                off[st][t] = 1 << b; used |= 1u << b;
            }
        }
#pragma unroll
        for (int t = 0; t < LOGV; t++)
            if (t >= m) off[st][t] = 1 << nth_unused_bit(used, t - m, LT);
        int wb = 0;
#pragma unroll
        for (int u = 0; u < 3; u++) wb |= (int)((wid >> u) & 1u) << nth_unused_bit(used, LOGV - m + u, LT);
        wbase[st] = wb;
    }
    __syncthreads();
#pragma unroll 1
    for (int ch = 0; ch < chunks_per_cta; ch++, lane0 += ENC_C) {
        u64 x[V];
        int ls[V];
#pragma unroll
        for (int j = 0; j < V; j++) {
            int l = wbase[0];
#pragma unroll
            for (int t = 0; t < LOGV; t++) if (j & (1 << t)) l += off[0][t];
            ls[j] = l;
            const uint32_t slot = slot_base + (uint32_t)l * A.stride;
            const u64 *src;
            if (A.in_by_pos) src = A.in + (uint64_t)A.out_pos[slot] * A.ct_words;
            else src = (slot < A.in_count) ? A.in + (uint64_t)slot * A.ct_words : A.zero;
            x[j] = src[lane0 + lane];
        }
#pragma unroll
        for (int st = 0; st < NST; st++) {
            if (st > 0) {
                if (st > 1 || ch > 0) __syncthreads();     // readers of the previous deal are done
#pragma unroll
                for (int j = 0; j < V; j++) xbuf[ls[j] * ENC_C + lane] = x[j];
                __syncthreads();
#pragma unroll
                for (int j = 0; j < V; j++) {
                    int l = wbase[st];
#pragma unroll
                    for (int t = 0; t < LOGV; t++) if (j & (1 << t)) l += off[st][t];
                    ls[j] = l;
                    x[j] = xbuf[l * ENC_C + lane];
                }
            }
#pragma unroll
            for (int t = 0; t < LOGV; t++) {
                const int lv = st * LOGV + t;
                if (lv < LT) {
                    const ulonglong2 *sc = stab + lv * T;
#pragma unroll
                    for (int j = 0; j < V; j++) {
                        if (j & (1 << t)) continue;
                        const int jj = j | (1 << t);
                        // values stay in [0, 3q) inside the pass: the 3-product Shoup estimate (ntt_blk.cuh) takes any 64-bit input to
                        // [0, 3q) by itself, an output without a scalar takes one conditional subtraction of 3q
                        const u64 a = x[j], b = x[jj];
                        u64 su = a + b, di = a - b + q3;
                        const ulonglong2 w0 = sc[ls[j]], w1 = sc[ls[jj]];
                        if (w0.x) su = mul_shoup_approx(su, w0.x, w0.y, nq); else su = csub_min(su, q3);
                        if (w1.x) di = mul_shoup_approx(di, w1.x, w1.y, nq); else di = csub_min(di, q3);
                        x[j] = su; x[jj] = di;
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < V; j++) {
            const uint32_t slot = slot_base + (uint32_t)ls[j] * A.stride;
            A.out[(uint64_t)A.out_pos[slot] * A.ct_words + lane0 + lane] = csub_min(csub_min(x[j], q2), q);   // canonical in HBM
        }
    }
}
static bool enc_tile_kernel_only() { static const bool v = getenv("LMN_ENCODE_TILE_KERNEL") != nullptr; return v; }
void launch_encode_pass(const Launch &L, const EncPassArgs &A) {
    LaunchScope scope(L, KID_ENCODE);
    const uint64_t chunks = (A.lane_words ? A.lane_words : A.ct_words) / ENC_C;   // lane ranges are whole limbs (or whole slices)
    const uint32_t tiles = A.n >> A.log_tile;
    if (A.log_tile >= 5 && A.log_tile <= 8 && A.s1 - A.s0 == A.log_tile && !enc_tile_kernel_only()) {
        const int per_limb = 1 << (A.logn - 5);
        static const int cpc_max = getenv("LMN_ENC_CPC") ? atoi(getenv("LMN_ENC_CPC")) : 8;   // chunks per CTA (one scalar-table gather per CTA)
        const int cpc = per_limb >= cpc_max ? cpc_max : per_limb;
        const int T = 1 << A.log_tile;
        const size_t smem = ((size_t)2 * A.log_tile * T + (size_t)T * ENC_C) * 8;
        const dim3 g((unsigned)(chunks / cpc), tiles);
        switch (A.log_tile) {
            case 5: k_encode_pass_r<2><<<g, 256, smem, L.stream>>>(A, cpc); break;
            case 6: k_encode_pass_r<3><<<g, 256, smem, L.stream>>>(A, cpc); break;
            case 7: k_encode_pass_r<4><<<g, 256, smem, L.stream>>>(A, cpc); break;
            default: k_encode_pass_r<5><<<g, 256, smem, L.stream>>>(A, cpc); break;
        }
        return;
    }
    dim3 g((unsigned)chunks, tiles);
    size_t smem = ((size_t)ENC_C << A.log_tile) * 8;
    k_encode_pass<<<g, 256, smem, L.stream>>>(A);
}

// =====================================================================================================
// hybrid key switch
// =====================================================================================================
// (1) y[i] = INTT(c1[i]) * (Q_d/q_i)^-1 mod q_i        (lattigo DecomposeNTT: INTT of c1, folded with reconstructRNS's y_i)
template <int LOGN>
__global__ void __launch_bounds__(1024) k_intt_c1(const ModDev *mods, const ulonglong2 *tw_inv, const KsConsts *C, const KsJobs J, int k) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int i = blockIdx.x, job = blockIdx.y;
    const ModDev md = mods[i];
    const u64 *src = J.col + (size_t)job * J.col_words + ((size_t)(k + i) << LOGN);  // poly 1, limb i
    u64 *dst = J.y + (size_t)job * J.y_words + ((size_t)i << LOGN);
    for (int n = threadIdx.x; n < N; n += blockDim.x) sm[PAD(n)] = src[n];
    __syncthreads();
    ntt_inv_smem<LOGN>(sm, tw_inv + ((size_t)i << LOGN), md.q);
    const u64 w = C->intt_w[i], ws = C->intt_ws[i];
    for (int n = threadIdx.x; n < N; n += blockDim.x) dst[n] = mul_shoup(sm[PAD(n)], w, ws, md.q);
}

// (2) base conversion of every digit to every other modulus, once per coefficient (lattigo DecomposeAndSplit:
//     reconstructRNS + multSum).  The float64 overflow estimate v = floor(sum_a float64(y_a)/float64(q_a)) (same operation
//     order as Go) is folded into the first residue: z_0 = y_0 + (nd-1-v) q_0, so that the value is sum_a z_a (Q_d/q_a)
//     - (nd-1) Q_d with no data-dependent term.  S = 2: the first Cooley-Tukey level of the length-N transform is applied
//     here (X + wY, X - wY), so that k_mac can run two independent half-length transforms.
template <int S>
__global__ void __launch_bounds__(256) k_modup(const ModDev *mods, const ulonglong2 *tw_fwd, const KsConsts *C, const KsJobs J,
                                                int logn, int beta_max) {
    const int d = blockIdx.y, job = blockIdx.z;
    const int nd = C->dig_size[d], s0 = C->dig_start[d];
    const int nq = C->nq, k = C->k, KP = k + C->np;
    const int M = (1 << logn) / S;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= M) return;
    const u64 *y = J.y + (size_t)job * J.y_words;
    u64 *mu = J.mu + (size_t)job * J.mu_words + (((size_t)d * KP) << logn);
    u64 z[S][MAX_P];
#pragma unroll
    for (int r = 0; r < S; r++) {
        double vf = 0.0;
#pragma unroll
        for (int a = 0; a < MAX_P; a++) {
            if (a < nd) {
                z[r][a] = y[((size_t)(s0 + a) << logn) + n + r * M];
                vf = __dadd_rn(vf, __ddiv_rn(__ull2double_rn(z[r][a]), __ull2double_rn(mods[s0 + a].q)));
            } else z[r][a] = 0;
        }
        if (nd >= 2) z[r][0] += (u64)(nd - 1 - (int)__double2ull_rz(vf)) * mods[s0].q;
    }
    for (int tix = 0; tix < nq + C->np; tix++) {
        if (tix < nq && tix >= s0 && tix < s0 + nd) continue;   // own limbs are copied from the NTT-domain input
        const int gm = ks_mod_index(tix, nq, k);
        const u64 p = mods[gm].q, pinv = mods[gm].qinv, p2 = 2 * p;
        u64 vals[S];
        if (nd == 1) {
            // single-prime digit: centred lift (DecomposeAndSplit, decompLvl < 0), then the key-row scale 2^(e-64)
            const u64 qd = mods[s0].q, hq = qd >> 1, gw = C->mu_g_w[d][tix], gws = C->mu_g_ws[d][tix], pmu = mods[gm].mu;
#pragma unroll
            for (int r = 0; r < S; r++) {
                const u64 c = z[r][0];
                const bool neg = c >= hq;
                const u64 rr = barrett(neg ? qd - c : c, p, pmu);
                vals[r] = mul_shoup_lazy(neg ? (rr ? p - rr : 0) : rr, gw, gws, p);
            }
        } else if (C->mu_fast[d][tix]) {
            const uint32_t sgn = C->mu_sgn[d][tix];
            uint32_t mag[MAX_P];
#pragma unroll
            for (int a = 0; a < MAX_P; a++) mag[a] = a < nd ? C->mu_mag[d][tix][a] : 0;
            const u64 klo = C->mu_k_lo[d][tix], khi = C->mu_k_hi[d][tix];
#pragma unroll
            for (int r = 0; r < S; r++) {
                u64 lo = klo, hi = khi;
#pragma unroll
                for (int a = 0; a < MAX_P; a++) {
                    if (a < nd) {
                        // 64 x 32 -> 96-bit product with two 32x32 multiplies
                        const u64 p0 = (u64)(uint32_t)z[r][a] * mag[a];
                        const u64 p1 = (u64)(uint32_t)(z[r][a] >> 32) * mag[a];
                        const u64 plo = p0 + (p1 << 32);
                        const u64 phi = (p1 >> 32) + (plo < p0);
                        if ((sgn >> a) & 1) { const u64 t = lo; lo -= plo; hi -= phi + (t < plo); }
                        else { lo += plo; hi += phi + (lo < plo); }
                    }
                }
                vals[r] = mont_reduce_lazy(lo, hi, p, pinv);
            }
        } else {
            u64 cm[MAX_P];
#pragma unroll
            for (int a = 0; a < MAX_P; a++) cm[a] = a < nd ? C->mu_c[d][tix][a] : 0;
            const u64 klo = C->mu_k_lo[d][tix], khi = C->mu_k_hi[d][tix];
#pragma unroll
            for (int r = 0; r < S; r++) {
                u64 lo = klo, hi = khi;
#pragma unroll
                for (int a = 0; a < MAX_P; a++)
                    if (a < nd) mac128(lo, hi, z[r][a], cm[a]);
                vals[r] = csub(mont_reduce_lazy(lo, hi, p, pinv), p2);
            }
        }
        u64 *o = mu + ((size_t)gm << logn);
        if (S > 1) {
            // the log2(S) levels that couple the S block transforms (exact lazy products: +2p per level, below 8p at S = 8)
            const ulonglong2 *twm = tw_fwd + ((size_t)gm << logn);
#pragma unroll
            for (int t = 0; (1 << t) < S; t++) {
                const int half = S >> (t + 1);
#pragma unroll
                for (int b = 0; b < (1 << t); b++) {
                    const ulonglong2 w = twm[(1 << t) + b];
#pragma unroll
                    for (int jj = 0; jj < half; jj++) bfly_ct(vals[b * 2 * half + jj], vals[b * 2 * half + jj + half], w, p, p2);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < S; r++) o[n + r * M] = vals[r];
    }
}

// (3) for one target modulus (and one half of the output when the ring is larger than the block transform):
//     acc_c = sum_d evk[d][c] (.) NTT(ModUp_d(y))       (DecomposeSingleNTT + gadgetProductMultiplePLazy)
template <int LOGM, int S>
__global__ void __launch_bounds__(1024) k_mac(const ModDev *mods, const ulonglong2 *tw_fwd, const KsConsts *C,
                                               const KsJobs J, const u64 *__restrict__ evk) {
    extern __shared__ u64 sm[];
    constexpr int M = 1 << LOGM;
    constexpr int LOGN = LOGM + (S == 2 ? 1 : 0);
    u64 *data = sm;
    u64 *acc0 = sm + padded_words(M);
    u64 *acc1 = acc0 + M;
    const int tix = blockIdx.x / S, h = blockIdx.x % S, job = blockIdx.y;
    const int nq = C->nq, k = C->k, KP = k + C->np;
    const int gm = ks_mod_index(tix, nq, k);
    const ModDev md = mods[gm];
    const u64 p = md.q;
    const ulonglong2 *tw = tw_fwd + ((size_t)gm << LOGN);
    const u64 *c1 = J.col + (size_t)job * J.col_words + ((size_t)k << LOGN);  // poly 1 (NTT domain)
    for (int n = threadIdx.x; n < M; n += blockDim.x) { acc0[n] = 0; acc1[n] = 0; }

    const int beta = C->beta;
    for (int d = 0; d < beta; d++) {
        const int s0 = C->dig_start[d], nd = C->dig_size[d];
        const u64 *e0 = evk + ((((size_t)d * 2 + 0) * KP + gm) << LOGN) + (size_t)h * M;
        const u64 *e1 = evk + ((((size_t)d * 2 + 1) * KP + gm) << LOGN) + (size_t)h * M;
        if (tix < nq && tix >= s0 && tix < s0 + nd) {
            // the digit's own limbs are copied from the NTT-domain input (DecomposeSingleNTT)
            const u64 *x = c1 + ((size_t)tix << LOGN) + (size_t)h * M;
            for (int n = threadIdx.x; n < M; n += blockDim.x) {
                const u64 xv = x[n];
                acc0[n] += mont_mul_lazy(xv, e0[n], p, md.qinv);
                acc1[n] += mont_mul_lazy(xv, e1[n], p, md.qinv);
            }
            continue;
        }
        __syncthreads();  // previous digit's readers of `data` are done
        {
            const u64 *src = J.mu + (size_t)job * J.mu_words + ((((size_t)d * KP) + gm) << LOGN) + (size_t)h * M;
            for (int n = threadIdx.x; n < M; n += blockDim.x) data[PAD(n)] = src[n];
        }
        __syncthreads();
        ntt_fwd_smem<LOGM>(data, tw, p, S + h);
        for (int n = threadIdx.x; n < M; n += blockDim.x) {
            const u64 xv = data[PAD(n)];
            acc0[n] += mont_mul_lazy(xv, e0[n], p, md.qinv);
            acc1[n] += mont_mul_lazy(xv, e1[n], p, md.qinv);
        }
    }
    u64 *o0 = J.acc + (size_t)job * J.acc_words + ((size_t)gm << LOGN) + (size_t)h * M;
    u64 *o1 = o0 + ((size_t)KP << LOGN);
    for (int n = threadIdx.x; n < M; n += blockDim.x) {
        o0[n] = barrett(acc0[n], p, md.mu);
        o1[n] = barrett(acc1[n], p, md.mu);
    }
}

// (4) P limbs of the accumulators out of the NTT domain, centred by P/2 and premultiplied by (P/p_j)^-1
//     (ModDownQPtoQNTT: INTT of the P part, ModUpPtoQ's AddScalar(PHalf) and reconstructRNS's y_j)
template <int LOGN>
__global__ void __launch_bounds__(1024) k_pintt(const ModDev *mods, const ulonglong2 *tw_inv, const KsConsts *C, const KsJobs J) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int j = blockIdx.x, c = blockIdx.y, job = blockIdx.z;
    const int k = C->k, KP = k + C->np, gm = k + j;
    const ModDev md = mods[gm];
    u64 *x = J.acc + (size_t)job * J.acc_words + (((size_t)c * KP + gm) << LOGN);
    for (int n = threadIdx.x; n < N; n += blockDim.x) sm[PAD(n)] = x[n];
    __syncthreads();
    ntt_inv_smem<LOGN>(sm, tw_inv + ((size_t)gm << LOGN), md.q);
    const u64 w = C->pin_w[j], ws = C->pin_ws[j], ad = C->pin_add[j];
    for (int n = threadIdx.x; n < N; n += blockDim.x) x[n] = add_mod(mul_shoup(sm[PAD(n)], w, ws, md.q), ad, md.q);
}

// (5) float64 overflow estimate of the P -> Q conversion folded into the first P residue
__global__ void k_vpfix(const ModDev *mods, const KsConsts *C, const KsJobs J, int logn) {
    const int c = blockIdx.y, job = blockIdx.z;
    const int np = C->np, k = C->k, KP = k + np;
    if (np < 2) return;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= (1 << logn)) return;
    u64 *yp = J.acc + (size_t)job * J.acc_words + (((size_t)c * KP + k) << logn);
    double vf = 0.0;
    for (int a = 0; a < np; a++)
        vf = __dadd_rn(vf, __ddiv_rn(__ull2double_rn(yp[((size_t)a << logn) + n]), __ull2double_rn(mods[k + a].q)));
    const u64 v = __double2ull_rz(vf);
    yp[n] += (u64)(np - 1 - (int)v) * mods[k].q;
}

// (6) ModDown by P, add c0, automorphism, accumulate into the ciphertext:
//     u_c = (acc_c - NTT(ext_c)) P^-1 (+ c0) ;  mode 0: col_c += pi(u_c)   mode 1: col_c = u_c
template <int LOGN>
__global__ void __launch_bounds__(1024) k_down(const ModDev *mods, const ulonglong2 *tw_fwd, const KsConsts *C,
                                                const KsJobs J, uint32_t gal_el, int mode) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int i = blockIdx.x, c = blockIdx.y, job = blockIdx.z;
    const int k = C->k, np = C->np, KP = k + np;
    const ModDev md = mods[i];
    const u64 q = md.q, q2 = 2 * md.q;
    const u64 *accb = J.acc + (size_t)job * J.acc_words + (((size_t)c * KP) << LOGN);
    const u64 *yp = accb + ((size_t)k << LOGN);
    u64 cm[MAX_P];
#pragma unroll
    for (int a = 0; a < MAX_P; a++) cm[a] = a < np ? C->md_c[i][a] : 0;
    const u64 neg = C->md_neg[i];
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        u64 lo = 0, hi = neg;
#pragma unroll
        for (int a = 0; a < MAX_P; a++)
            if (a < np) mac128(lo, hi, yp[((size_t)a << LOGN) + n], cm[a]);
        sm[PAD(n)] = csub(mont_reduce_lazy(lo, hi, q, md.qinv), q2);
    }
    __syncthreads();
    ntt_fwd_smem<LOGN>(sm, tw_fwd + ((size_t)i << LOGN), q, 1);
    const u64 w = C->pinv_w[i], ws = C->pinv_ws[i];
    const u64 *aq = accb + ((size_t)i << LOGN);
    u64 *col = J.col + (size_t)job * J.col_words + (((size_t)c * k + i) << LOGN);
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const u64 ext = barrett(sm[PAD(n)], q, md.mu);
        u64 u = mul_shoup(aq[n] + q - ext, w, ws, q);
        if (c == 0) u = add_mod(u, col[n], q);
        if (mode == 1) col[n] = u; else sm[PAD(n)] = u;
    }
    if (mode == 1) return;
    __syncthreads();
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const unsigned src = auto_index((unsigned)n, gal_el, LOGN);
        col[n] = add_mod(col[n], sm[PAD(src)], q);
    }
}

const char *kernels_blk_init();
static bool use_blk_pipeline(int logn) { return blk_blocks(logn) > 0; }   // log N >= 12; smaller rings: whole-limb kernels

// block pipeline (kernels_blk.cu): six launches per rotation, out of place col -> col2
static void rotate_accumulate_blk(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch, KsJobs &J,
                                  const u64 *evk, uint32_t gal_el) {
    const int nq = Ch.nq, np = Ch.np, k = T.k, KP = T.k + T.np;
    BlkSlots A;
    A.src = J.col; A.src_job_words = J.col_words; A.dst = J.y; A.dst_job_words = J.y_words; A.nslots = nq;
    for (int i = 0; i < nq; i++) { A.mod[i] = i; A.src_limb[i] = (uint32_t)(k + i); A.dst_limb[i] = (uint32_t)i; }
    launch_blk_intt(L, T, A, J.jobs, KID_INTT_C1);
    launch_modup_blk(L, T, Cd, Ch, J);
    launch_mac_blk(L, T, Cd, Ch, J, evk);
    BlkSlots B;
    B.src = J.acc; B.src_job_words = J.acc_words; B.dst = J.acc; B.dst_job_words = J.acc_words; B.nslots = 2 * np;
    for (int c = 0; c < 2; c++)
        for (int j = 0; j < np; j++) { const int s = c * np + j; B.mod[s] = k + j; B.src_limb[s] = B.dst_limb[s] = (uint32_t)(c * KP + k + j); }
    launch_blk_intt(L, T, B, J.jobs, KID_PINTT);
    launch_down_pre(L, T, Cd, Ch, J);
    launch_down_blk(L, T, Cd, Ch, J, gal_el, gal_el == 0 ? 1 : 0);
    u64 *t = J.col; J.col = J.col2; J.col2 = t;
}

void launch_rotate_accumulate(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch,
                              KsJobs &J, const u64 *evk, uint32_t gal_el) {
    if (use_blk_pipeline(T.logn)) return rotate_accumulate_blk(L, T, Cd, Ch, J, evk, gal_el);
    const int N = T.n, nq = Ch.nq, np = Ch.np, jobs = J.jobs;
    const size_t sm_full = (size_t)padded_words(N) * 8;
    const int nt_full = ntt_threads(N);
    { LaunchScope scope(L, KID_INTT_C1, (uint64_t)jobs);
    LMN_DISPATCH_LOGN(T.logn, k_intt_c1<LG><<<dim3(nq, jobs), nt_full, sm_full, L.stream>>>(T.mods, T.tw_inv, Cd, J, Ch.k)); }
    {
        LaunchScope scope(L, KID_VFIX, (uint64_t)jobs);
        const int bs = N < 256 ? N : 256;
        const int beta_max = (Ch.k + np - 1) / np;
        k_modup<1><<<dim3((N + bs - 1) / bs, Ch.beta, jobs), bs, 0, L.stream>>>(T.mods, T.tw_fwd, Cd, J, T.logn, beta_max);
    }
    {
        LaunchScope scope(L, KID_MAC, (uint64_t)jobs);
        const size_t smm = ((size_t)padded_words(N) + 2 * (size_t)N) * 8;
        LMN_DISPATCH_LOGN(T.logn, k_mac<(LG <= 13 ? LG : 13), 1><<<dim3(nq + np, jobs), nt_full, smm, L.stream>>>(T.mods, T.tw_fwd, Cd, J, evk));
    }
    { LaunchScope scope(L, KID_PINTT, (uint64_t)jobs);
    LMN_DISPATCH_LOGN(T.logn, k_pintt<LG><<<dim3(np, 2, jobs), nt_full, sm_full, L.stream>>>(T.mods, T.tw_inv, Cd, J)); }
    if (np > 1) {
        int bs = N < 256 ? N : 256;
        LaunchScope scope(L, KID_VPFIX, (uint64_t)jobs);
        k_vpfix<<<dim3((N + bs - 1) / bs, 2, jobs), bs, 0, L.stream>>>(T.mods, Cd, J, T.logn);
        }
    { LaunchScope scope(L, KID_DOWN, (uint64_t)jobs);
    LMN_DISPATCH_LOGN(T.logn, k_down<LG><<<dim3(nq, 2, jobs), nt_full, sm_full, L.stream>>>(T.mods, T.tw_fwd, Cd, J, gal_el,
                                                                                        gal_el == 0 ? 1 : 0)); }
}

void launch_keyswitch_level0(const Launch &L, const DeviceTables &T, const KsConsts *Cd, const KsConsts &Ch,
                             KsJobs &J, const u64 *evk) {
    launch_rotate_accumulate(L, T, Cd, Ch, J, evk, 0 /* mode 1: replace, no automorphism */);
}

// =====================================================================================================
// Rescale (bgv.Evaluator.Rescale -> ring.DivRoundByLastModulusNTT)
// =====================================================================================================
template <int LOGN>
__global__ void __launch_bounds__(1024) k_rs_intt(const ModDev *mods, const ulonglong2 *tw_inv, const RescaleConsts *R, const KsJobs J, int k,
                          int level) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int c = blockIdx.x, job = blockIdx.y;
    const ModDev md = mods[level];
    const u64 *x = J.col + (size_t)job * J.col_words + (((size_t)c * k + level) << LOGN);
    u64 *r = J.acc + (size_t)job * J.acc_words + ((size_t)c << LOGN);
    for (int n = threadIdx.x; n < N; n += blockDim.x) sm[PAD(n)] = x[n];
    __syncthreads();
    ntt_inv_smem<LOGN>(sm, tw_inv + ((size_t)level << LOGN), md.q);
    const u64 half = R->half;
    for (int n = threadIdx.x; n < N; n += blockDim.x)
        r[n] = add_mod(mul_shoup(sm[PAD(n)], md.ninv, md.ninvs, md.q), half, md.q);
}
template <int LOGN>
__global__ void __launch_bounds__(1024) k_rs_apply(const ModDev *mods, const ulonglong2 *tw_fwd, const RescaleConsts *R, const KsJobs J, int k) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int i = blockIdx.x, c = blockIdx.y, job = blockIdx.z;
    const ModDev md = mods[i];
    const u64 q = md.q;
    const u64 *r = J.acc + (size_t)job * J.acc_words + ((size_t)c << LOGN);
    const u64 hm = R->half_mod[i];
    for (int n = threadIdx.x; n < N; n += blockDim.x) sm[PAD(n)] = sub_mod(barrett(r[n], q, md.mu), hm, q);
    __syncthreads();
    ntt_fwd_smem<LOGN>(sm, tw_fwd + ((size_t)i << LOGN), q, 1);
    u64 *x = J.col + (size_t)job * J.col_words + (((size_t)c * k + i) << LOGN);
    const u64 w = R->qlinv_w[i], ws = R->qlinv_ws[i];
    for (int n = threadIdx.x; n < N; n += blockDim.x)
        x[n] = mul_shoup(x[n] + q - barrett(sm[PAD(n)], q, md.mu), w, ws, q);
}
void launch_rescale_step(const Launch &L, const DeviceTables &T, const RescaleConsts *Rd, int level, const KsJobs &J) {
    if (use_blk_pipeline(T.logn)) {
        BlkSlots A;
        A.src = J.col; A.src_job_words = J.col_words; A.dst = J.acc; A.dst_job_words = J.acc_words; A.nslots = 2;
        for (int c = 0; c < 2; c++) { A.mod[c] = level; A.src_limb[c] = (uint32_t)(c * T.k + level); A.dst_limb[c] = (uint32_t)c; }
        launch_blk_intt(L, T, A, J.jobs, KID_RS_INTT);
        launch_rs_pre(L, T, Rd, level, J);
        launch_rs_blk(L, T, Rd, level, J);
        return;
    }
    const size_t smf = (size_t)padded_words(T.n) * 8;
    const int nt = ntt_threads(T.n);
    { LaunchScope scope(L, KID_RS_INTT);
    LMN_DISPATCH_LOGN(T.logn, k_rs_intt<LG><<<dim3(2, J.jobs), nt, smf, L.stream>>>(T.mods, T.tw_inv, Rd, J, T.k, level)); }
    { LaunchScope scope(L, KID_RS_APPLY);
    LMN_DISPATCH_LOGN(T.logn, k_rs_apply<LG><<<dim3(level, 2, J.jobs), nt, smf, L.stream>>>(T.mods, T.tw_fwd, Rd, J, T.k)); }
}

// =====================================================================================================
// data movement between stores and the job workspace
// =====================================================================================================
__global__ void k_pack_out(const KsJobs J, int k, int nl, int logn, u64 *__restrict__ out) {
    // grid: x over N/2, y = poly*nl + limb, z = job
    const int poly = blockIdx.y / nl, limb = blockIdx.y % nl;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << (logn - 1))) return;
    const u64 *src = J.col + (size_t)blockIdx.z * J.col_words + (((size_t)poly * k + limb) << logn);
    u64 *dst = out + (((size_t)blockIdx.z * 2 * nl + (size_t)poly * nl + limb) << logn);
    reinterpret_cast<ulonglong2 *>(dst)[i] = reinterpret_cast<const ulonglong2 *>(src)[i];
}
void launch_pack_out(const Launch &L, const DeviceTables &T, const KsJobs &J, int nl, u64 *out) {
    LaunchScope scope(L, KID_PACK);
    int half = T.n / 2, bs = half < 256 ? half : 256;
    k_pack_out<<<dim3((half + bs - 1) / bs, 2 * nl, J.jobs), bs, 0, L.stream>>>(J, T.k, nl, T.logn, out);
}
__global__ void k_gather_in(const u64 *__restrict__ src, uint64_t src_words, const uint32_t *__restrict__ idx, const KsJobs J,
                            int k, int logn) {
    // grid: x over N/2, y = poly*k + limb, z = job ; source cts are at max level (k limbs)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1 << (logn - 1))) return;
    const u64 *s = src + (size_t)idx[blockIdx.z] * src_words + ((size_t)blockIdx.y << logn);
    u64 *d = J.col + (size_t)blockIdx.z * J.col_words + ((size_t)blockIdx.y << logn);
    reinterpret_cast<ulonglong2 *>(d)[i] = reinterpret_cast<const ulonglong2 *>(s)[i];
}
void launch_gather_in(const Launch &L, const DeviceTables &T, const u64 *src, uint64_t src_words, const uint32_t *idx,
                      const KsJobs &J) {
    LaunchScope scope(L, KID_GATHER);
    int half = T.n / 2, bs = half < 256 ? half : 256;
    k_gather_in<<<dim3((half + bs - 1) / bs, 2 * T.k, J.jobs), bs, 0, L.stream>>>(src, src_words, idx, J, T.k, T.logn);
}

// ring switch tail (rlwe.SwitchCiphertextRingDegreeNTT, large -> small): INTT in the large ring, keep every
// (N/Ns)-th coefficient, NTT in the small ring of the same prime
template <int LOGN, int LOGS>
__global__ void __launch_bounds__(1024) k_ring_decimate(const ModDev *mods, const ulonglong2 *tw_inv, const KsJobs J, int k, const ModDev *smod,
                                const ulonglong2 *stw, u64 *__restrict__ out) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN, NS = 1 << LOGS;
    const int c = blockIdx.x, job = blockIdx.y;
    const ModDev md = mods[0];
    const u64 *x = J.col + (size_t)job * J.col_words + (((size_t)c * k) << LOGN);
    for (int n = threadIdx.x; n < N; n += blockDim.x) sm[PAD(n)] = x[n];
    __syncthreads();
    ntt_inv_smem<LOGN>(sm, tw_inv, md.q);
    // decimate into the front of a second region (padded layout of the small transform)
    u64 *sm2 = sm + padded_words(N);
    for (int n = threadIdx.x; n < NS; n += blockDim.x)
        sm2[PAD(n)] = mul_shoup(sm[PAD(n << (LOGN - LOGS))], md.ninv, md.ninvs, md.q);
    __syncthreads();
    ntt_fwd_smem<LOGS>(sm2, stw, md.q, 1);
    u64 *o = out + (((size_t)job * 2 + c) << LOGS);
    for (int n = threadIdx.x; n < NS; n += blockDim.x) o[n] = barrett(sm2[PAD(n)], md.q, smod->mu);
}
void launch_ring_decimate(const Launch &L, const DeviceTables &T, const KsJobs &J, int logns, const ModDev *small_mod,
                          const ulonglong2 *small_tw_fwd, u64 *out) {
    LaunchScope scope(L, KID_DECIMATE);
    const size_t smem = ((size_t)padded_words(T.n) + padded_words(1 << logns)) * 8;
    const int nt = ntt_threads(T.n);
#define RD(LN, LS) k_ring_decimate<LN, LS><<<dim3(2, J.jobs), nt, smem, L.stream>>>(T.mods, T.tw_inv, J, T.k, small_mod, small_tw_fwd, out)
    if (logns == 10) { if (T.logn == 12) RD(12, 10); else if (T.logn == 13) RD(13, 10); else if (T.logn == 14) RD(14, 10); }
    else if (logns == 6) { if (T.logn == 8) RD(8, 6); else if (T.logn == 10) RD(10, 6); }
    else if (logns == 8) { if (T.logn == 10) RD(10, 8); else if (T.logn == 12) RD(12, 8); }
#undef RD
}

// =====================================================================================================
// integer-pipe ceiling: the radix-8 butterfly network of the block transforms on register-resident data (no loads, no barriers), 64
// warps per SM, in the FASTEST formulation measured on this part - tools/int_pipe_probe.cu k_bfly<1>: 3-product Shoup quotient with
// the middle sum as the 64-bit addend of IMAD.WIDE, plain 64-bit add and subtract, lazy [0, 3q) - and with nothing else in the loop
// (values wrap mod 2^64 after a few rounds; the instruction stream, which is what is timed, does not depend on them).  This is the
// denominator of the integer roofline; the product's own butterfly (bfly_ct3, ALU-pinned carries) is timed beside it.
template <int PRODUCT>
__global__ void __launch_bounds__(256) k_peak_bfly(const ModDev *mods, int rounds, u64 *sink) {
    const u64 q = mods[0].q;
    const FwdMod fm = {q, 0 - q, mods[0].q3, (unsigned)mods[0].pad[0]};
    u64 a[8];
#pragma unroll
    for (int r = 0; r < 8; r++) a[r] = (u64)(threadIdx.x * 8 + r + blockIdx.x) % q;
    ulonglong2 w[7];
#pragma unroll
    for (int r = 0; r < 7; r++) { w[r].x = (mods[0].ninv + r * 977) % q; w[r].y = mods[0].ninvs + r; }
#define LMN_PB(x, y, tw) do { if (PRODUCT) bfly_ct3(x, y, tw, fm); else bfly_ct3_probe(x, y, tw, fm); } while (0)
    for (int it = 0; it < rounds; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) LMN_PB(a[r], a[r + 4], w[0]);
        LMN_PB(a[0], a[2], w[1]); LMN_PB(a[1], a[3], w[1]);
        LMN_PB(a[4], a[6], w[2]); LMN_PB(a[5], a[7], w[2]);
#pragma unroll
        for (int r = 0; r < 4; r++) LMN_PB(a[2 * r], a[2 * r + 1], w[3 + r]);
    }
#undef LMN_PB
    u64 x = 0;
#pragma unroll
    for (int r = 0; r < 8; r++) x ^= a[r];
    if (x == 0x1234567) sink[0] = x;
}
// product = 0: the ceiling above; 1: the same network with the product's own butterfly
double measure_peak_butterflies(cudaStream_t stream, const ModDev *mods, int sm_count, int product) {
    u64 *sink = nullptr;
    if (cudaMalloc(&sink, 8) != cudaSuccess) return 0.0;
    const int rounds = 4096, blocks = sm_count * 8, threads = 256;      // 64 warps per SM
    auto launch = [&](int r) {
        if (product) k_peak_bfly<1><<<blocks, threads, 0, stream>>>(mods, r, sink);
        else k_peak_bfly<0><<<blocks, threads, 0, stream>>>(mods, r, sink);
    };
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(64);   // warm-up
    cudaEventRecord(e0, stream);
    launch(rounds);
    cudaEventRecord(e1, stream);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    if (ms <= 0) return 0.0;
    return (double)blocks * threads * rounds * 12.0 / (ms * 1e-3);
}

// =====================================================================================================
const char *kernels_init() {
    cudaError_t e = cudaSuccess;
    const int maxsm = 227 * 1024;
#define SETSM(K) do { if (e == cudaSuccess) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, maxsm); } while (0)
#define SET_LOGN(LG)                       \
    SETSM(k_limb_ntt<LG>); SETSM(k_intt_c1<LG>); SETSM(k_pintt<LG>); SETSM(k_down<LG>); \
    SETSM(k_rs_intt<LG>); SETSM(k_rs_apply<LG>);
    SET_LOGN(6) SET_LOGN(8) SET_LOGN(10) SET_LOGN(12) SET_LOGN(13) SET_LOGN(14)
    SETSM((k_mac<6, 1>)); SETSM((k_mac<8, 1>)); SETSM((k_mac<10, 1>)); SETSM((k_mac<12, 1>)); SETSM((k_mac<13, 1>));
    SETSM(k_encode_pass); SETSM(k_encode_pass_r<2>); SETSM(k_encode_pass_r<3>); SETSM(k_encode_pass_r<4>); SETSM(k_encode_pass_r<5>);
    SETSM((k_ring_decimate<12, 10>)); SETSM((k_ring_decimate<13, 10>)); SETSM((k_ring_decimate<14, 10>));
    SETSM((k_ring_decimate<8, 6>)); SETSM((k_ring_decimate<10, 6>)); SETSM((k_ring_decimate<10, 8>)); SETSM((k_ring_decimate<12, 8>));
#undef SET_LOGN
#undef SETSM
    if (e != cudaSuccess) return cudaGetErrorString(e);
    if (const char *m = kernels_blk_init()) return m;
    return kernels_mac_init();
    return cudaGetErrorString(e);
}

}  // namespace lmn
