// The stages either side of the server hot path, on the device:
//   k_witness        core.RandomMatrixRowMajor (core/utils.go:46-82): ChaCha20 (IETF) keystream words mod t, column-major out
//   k_pt_encode      bgv.Encoder.Encode, ring-T half (lattigo EncodeRingT): slot permutation + INTT over Z_t[X]/(X^N+1)
//   k_pt_lift        ... ring-Q half (RingT2Q, scaleUp): lift [0,t) into q_i, times t^-1, NTT  -> the plaintexts of Prove
//   k_encrypt        rlwe.Encryptor.EncryptNew under the public key (cmd/server/main.go:199-207; fhe/code.go:15-22 for the zero
//                    column): ct = (u pk0 + e0 + pt, u pk1 + e1), u uniform ternary, e Gaussian(sigma 3.2, bound 19)
//   k_decrypt_limb   rlwe.Decryptor (fhe/ligero.go:577-636): c0 + c1 s, INTT, times t
//   k_crt_decode     bgv.Encoder.Decode (RingQ2T with rounding, DecodeRingT): Garner to mod t, Scale^-1, NTT over Z_t, un-permute
// One CTA per (limb, ciphertext) with the whole limb in shared memory (the whole-limb transforms of device.cuh): these stages are
// 1 - 2 % of the prover's integer work, so they use the simple transform, not the block pipeline.
// Randomness: every sample is a function of (seed key, column id, stream, coefficient) through ChaCha20, so a column encrypts to the
// same ciphertext whichever rank / batch processes it.  Lattigo's own PRNG stream cannot be reproduced (not vendored): the
// contract of this stage is decrypted-plaintext parity and the noise bound, not ciphertext bits.
#include "client.cuh"

namespace lmn {

// ---------------------------------------------------------------------------------------------------
// ChaCha20 block function (RFC 8439 section 2.3): 32-bit counter, 96-bit nonce
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t rotl32(uint32_t x, int n) { return __funnelshift_l(x, x, n); }
#define LMN_QR(a, b, c, d) \
    a += b; d ^= a; d = rotl32(d, 16); c += d; b ^= c; b = rotl32(b, 12); a += b; d ^= a; d = rotl32(d, 8); c += d; b ^= c; b = rotl32(b, 7);
__device__ __forceinline__ void chacha20_block(const ChaChaKey &key, uint32_t counter, uint32_t n0, uint32_t n1, uint32_t n2, uint32_t (&out)[16]) {
    uint32_t in[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key.w[0], key.w[1], key.w[2], key.w[3],
                       key.w[4], key.w[5], key.w[6], key.w[7], counter, n0, n1, n2};
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = in[i];
#pragma unroll 1
    for (int r = 0; r < 10; r++) {
        LMN_QR(x[0], x[4], x[8], x[12]) LMN_QR(x[1], x[5], x[9], x[13]) LMN_QR(x[2], x[6], x[10], x[14]) LMN_QR(x[3], x[7], x[11], x[15])
        LMN_QR(x[0], x[5], x[10], x[15]) LMN_QR(x[1], x[6], x[11], x[12]) LMN_QR(x[2], x[7], x[8], x[13]) LMN_QR(x[3], x[4], x[9], x[14])
    }
#pragma unroll
    for (int i = 0; i < 16; i++) out[i] = x[i] + in[i];
}
#undef LMN_QR

// witness: thread per output word; W is column-major [n_cols][rows]
__global__ void k_witness(const ChaChaKey key, u64 t, u64 t_mu, uint32_t rows, uint32_t cols, uint32_t first_col, uint32_t n_cols, u64 *__restrict__ W) {
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (uint64_t)rows * n_cols) return;
    const uint32_t j = (uint32_t)(idx / rows), i = (uint32_t)(idx - (uint64_t)j * rows);
    const uint64_t word = (uint64_t)i * cols + first_col + j;          // row-major position in the keystream (core/utils.go:62-68)
    uint32_t ks[16];
    chacha20_block(key, (uint32_t)(word >> 3), 0u, 0u, 0u, ks);
    const uint32_t w = (uint32_t)(word & 7u);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int s = 0; s < 8; s++) if (s == (int)w) { lo = ks[2 * s]; hi = ks[2 * s + 1]; }
    W[idx] = barrett(((u64)hi << 32) | lo, t, t_mu);
}

// ---------------------------------------------------------------------------------------------------
#define LMN_CLIENT_DISPATCH(LOGN, ...)                                      \
    switch (LOGN) {                                                         \
        case 6: { constexpr int LG = 6; __VA_ARGS__; } break;               \
        case 8: { constexpr int LG = 8; __VA_ARGS__; } break;               \
        case 10: { constexpr int LG = 10; __VA_ARGS__; } break;             \
        case 12: { constexpr int LG = 12; __VA_ARGS__; } break;             \
        case 13: { constexpr int LG = 13; __VA_ARGS__; } break;             \
        case 14: { constexpr int LG = 14; __VA_ARGS__; } break;             \
        default: break;                                                     \
    }
static inline int client_threads(int n) { int t = n / 8; return t < 32 ? 32 : (t > 1024 ? 1024 : t); }

template <int LOGN>
__global__ void __launch_bounds__(1024) k_pt_encode(const PlainRing R, const u64 *__restrict__ values, uint32_t n_values, uint64_t value_stride,
                                                    u64 *__restrict__ m) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const u64 t = R.mod.q;
    const u64 *v = values + (size_t)blockIdx.x * value_stride;
    for (int c = threadIdx.x; c < N; c += blockDim.x) sm[PAD(c)] = 0;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n_values; i += blockDim.x) sm[PAD((int)R.perm[i])] = barrett(v[i], t, R.mod.mu);   // ringT.Reduce
    __syncthreads();
    ntt_inv_smem<LOGN>(sm, R.tw_inv, t);
    u64 *o = m + ((size_t)blockIdx.x << LOGN);
    for (int c = threadIdx.x; c < N; c += blockDim.x) o[c] = mul_shoup(sm[PAD(c)], R.mod.ninv, R.mod.ninvs, t);
}

template <int LOGN>
__global__ void __launch_bounds__(1024) k_pt_lift(const ModDev *__restrict__ mods, const ulonglong2 *__restrict__ tw_fwd, int nl,
                                                  const u64 *__restrict__ m, const ulonglong2 *__restrict__ tinv, u64 *__restrict__ pt) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int i = blockIdx.x;
    const ModDev md = mods[i];
    const ulonglong2 ti = tinv[i];
    const u64 *src = m + ((size_t)blockIdx.y << LOGN);
    for (int c = threadIdx.x; c < N; c += blockDim.x) sm[PAD(c)] = mul_shoup(src[c], ti.x, ti.y, md.q);
    __syncthreads();
    ntt_fwd_smem<LOGN>(sm, tw_fwd + ((size_t)i << LOGN), md.q, 1);
    u64 *o = pt + (((size_t)blockIdx.y * nl + i) << LOGN);
    for (int c = threadIdx.x; c < N; c += blockDim.x) o[c] = barrett(sm[PAD(c)], md.q, md.mu);
}

// 16 samples of one ChaCha20 block per thread-task; stream (low byte) 0: uniform ternary, 1 / 2: rounded Gaussian.  The upper bits of
// `stream` are the domain tag of the caller (ENC_DOMAIN_*), so that the nonce (id, tag | stream, attempt) of a matrix column, of the zero
// column and of the witness keystream (nonce 0, 0, 0) can never coincide under one 32-byte key
template <int LOGN>
__device__ __forceinline__ void fill_small(u64 *sm, const ChaChaKey &seed, uint32_t id, uint32_t stream, u64 q) {
    constexpr int N = 1 << LOGN;
    for (int b = threadIdx.x; b < N / 16; b += blockDim.x) {
        uint32_t ks[16];
        chacha20_block(seed, (uint32_t)b, id, stream, 0u, ks);
        if ((stream & 0xffu) == 0) {
#pragma unroll
            for (int s = 0; s < 16; s++) {
                const uint32_t x = (uint32_t)(((u64)ks[s] * 3u) >> 32);                    // {0, 1, 2} -> {0, +1, -1}
                sm[PAD(16 * b + s)] = x == 0 ? 0 : (x == 1 ? 1 : q - 1);
            }
        } else {
#pragma unroll 1
            for (int pr = 0; pr < 8; pr++) {
                uint32_t a = ks[2 * pr], bb = ks[2 * pr + 1];
                int v0, v1;
                for (uint32_t attempt = 1;; attempt++) {
                    const float u1 = ((float)a + 1.0f) * 2.3283064365386963e-10f;           // (0, 1]
                    const float u2 = (float)bb * 2.3283064365386963e-10f;
                    const float r = sqrtf(-2.0f * logf(u1)) * 3.2f;
                    float sn, cs;
                    sincospif(2.0f * u2, &sn, &cs);
                    v0 = __float2int_rn(r * cs); v1 = __float2int_rn(r * sn);
                    if (v0 >= -19 && v0 <= 19 && v1 >= -19 && v1 <= 19) break;             // lattigo DefaultXe: bound 19
                    uint32_t k2[16];
                    chacha20_block(seed, (uint32_t)b, id, stream, attempt, k2);             // fresh words for the rejected pair
                    a = k2[2 * pr]; bb = k2[2 * pr + 1];
                }
                sm[PAD(16 * b + 2 * pr)] = v0 >= 0 ? (u64)v0 : q - (u64)(-v0);
                sm[PAD(16 * b + 2 * pr + 1)] = v1 >= 0 ? (u64)v1 : q - (u64)(-v1);
            }
        }
    }
}

template <int LOGN>
__global__ void __launch_bounds__(1024) k_encrypt(const ModDev *__restrict__ mods, const ulonglong2 *__restrict__ tw_fwd, int k, const ChaChaKey seed,
                                                  uint32_t first_id, uint32_t domain, const u64 *__restrict__ m, const ulonglong2 *__restrict__ tinv,
                                                  const u64 *__restrict__ pk, u64 *__restrict__ cts, uint64_t ct_words) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int i = blockIdx.x;
    const uint32_t j = blockIdx.y, id = first_id + j;
    const ModDev md = mods[i];
    const u64 q = md.q;
    const ulonglong2 *tw = tw_fwd + ((size_t)i << LOGN);
    const u64 *pk0 = pk + ((size_t)i << LOGN), *pk1 = pk + ((size_t)(k + i) << LOGN);
    u64 *c0 = cts + (size_t)j * ct_words + ((size_t)i << LOGN), *c1 = cts + (size_t)j * ct_words + ((size_t)(k + i) << LOGN);
    // u pk
    fill_small<LOGN>(sm, seed, id, domain | 0u, q);
    __syncthreads();
    ntt_fwd_smem<LOGN>(sm, tw, q, 1);
    for (int c = threadIdx.x; c < N; c += blockDim.x) {
        const u64 u = barrett(sm[PAD(c)], q, md.mu);
        c0[c] = mont_mul(u, pk0[c], q, md.qinv);
        c1[c] = mont_mul(u, pk1[c], q, md.qinv);
    }
    __syncthreads();
    // + e0 + pt  (the plaintext enters in the coefficient domain: NTT(e0 + m t^-1) = NTT(e0) + pt)
    fill_small<LOGN>(sm, seed, id, domain | 1u, q);
    __syncthreads();
    if (m) {
        const ulonglong2 ti = tinv[i];
        const u64 *src = m + ((size_t)j << LOGN);
        for (int c = threadIdx.x; c < N; c += blockDim.x) sm[PAD(c)] = add_mod(sm[PAD(c)], mul_shoup(src[c], ti.x, ti.y, q), q);
        __syncthreads();
    }
    ntt_fwd_smem<LOGN>(sm, tw, q, 1);
    for (int c = threadIdx.x; c < N; c += blockDim.x) c0[c] = add_mod(c0[c], barrett(sm[PAD(c)], q, md.mu), q);
    __syncthreads();
    // + e1
    fill_small<LOGN>(sm, seed, id, domain | 2u, q);
    __syncthreads();
    ntt_fwd_smem<LOGN>(sm, tw, q, 1);
    for (int c = threadIdx.x; c < N; c += blockDim.x) c1[c] = add_mod(c1[c], barrett(sm[PAD(c)], q, md.mu), q);
}

// test hook: the small polynomial fill_small draws for (id, stream), as signed coefficients
template <int LOGN>
__global__ void __launch_bounds__(1024) k_dump_small(const ChaChaKey seed, uint32_t id, uint32_t stream, u64 q, long long *__restrict__ out) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    fill_small<LOGN>(sm, seed, id, stream, q);
    __syncthreads();
    for (int c = threadIdx.x; c < N; c += blockDim.x) {
        const u64 v = sm[PAD(c)];
        out[c] = v > q / 2 ? -(long long)(q - v) : (long long)v;
    }
}

template <int LOGN>
__global__ void __launch_bounds__(1024) k_decrypt_limb(const ModDev *__restrict__ mods, const ulonglong2 *__restrict__ tw_inv, const CrtConsts *__restrict__ C,
                                                       int nl, const u64 *__restrict__ cts, uint64_t ct_words, const u64 *__restrict__ sk,
                                                       u64 *__restrict__ buf) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int i = blockIdx.x;
    const ModDev md = mods[i];
    const u64 q = md.q;
    const u64 *c0 = cts + (size_t)blockIdx.y * ct_words + ((size_t)i << LOGN), *c1 = cts + (size_t)blockIdx.y * ct_words + ((size_t)(nl + i) << LOGN);
    const u64 *s = sk + ((size_t)i << LOGN);
    for (int c = threadIdx.x; c < N; c += blockDim.x) sm[PAD(c)] = add_mod(c0[c], mont_mul(c1[c], s[c], q, md.qinv), q);
    __syncthreads();
    ntt_inv_smem<LOGN>(sm, tw_inv + ((size_t)i << LOGN), q);
    const u64 w = C->t_w[i], ws = C->t_ws[i];
    u64 *o = buf + (((size_t)blockIdx.y * nl + i) << LOGN);
    for (int c = threadIdx.x; c < N; c += blockDim.x) o[c] = mul_shoup(sm[PAD(c)], w, ws, q);
}

template <int LOGN>
__global__ void __launch_bounds__(1024) k_crt_decode(const PlainRing R, const CrtConsts *__restrict__ C, const ModDev *__restrict__ mods,
                                                     const u64 *__restrict__ buf, uint32_t n_values, u64 *__restrict__ values) {
    extern __shared__ u64 sm[];
    constexpr int N = 1 << LOGN;
    const int nl = C->nl;
    const u64 t = R.mod.q;
    const u64 *b = buf + (((size_t)blockIdx.x * nl) << LOGN);
    for (int c = threadIdx.x; c < N; c += blockDim.x) {
        u64 v[CRT_MAX];
        // mixed-radix digits of x + (Q - 1)/2 (the residues of (Q-1)/2 are (q_i - 1)/2)
        for (int i = 0; i < nl; i++) {
            const ModDev md = mods[i];
            u64 cur = add_mod(b[((size_t)i << LOGN) + c], C->half[i], md.q);
            for (int jj = 0; jj < i; jj++) cur = mul_shoup(sub_mod(cur, barrett(v[jj], md.q, md.mu), md.q), C->inv_w[i][jj], C->inv_ws[i][jj], md.q);
            v[i] = cur;
        }
        u64 acc = 0;
        for (int i = nl - 1; i >= 0; i--) acc = add_mod(mont_mul(acc, C->q_mod_t[i], t, R.mod.qinv), barrett(v[i], t, R.mod.mu), t);
        acc = sub_mod(acc, C->half_q_mod_t, t);                                   // centred representative of x mod Q, reduced mod t
        sm[PAD(c)] = mul_shoup(acc, C->sinv_w, C->sinv_ws, t);
    }
    __syncthreads();
    ntt_fwd_smem<LOGN>(sm, R.tw_fwd, t, 1);
    u64 *o = values + (size_t)blockIdx.x * n_values;
    for (uint32_t i = threadIdx.x; i < n_values; i += blockDim.x) o[i] = barrett(sm[PAD((int)R.perm[i])], t, R.mod.mu);
}

// ---------------------------------------------------------------------------------------------------
void launch_witness(const Launch &L, const ChaChaKey &key, u64 t, uint32_t rows, uint32_t cols, uint32_t first_col, uint32_t n_cols, u64 *W) {
    LaunchScope scope(L, KID_OTHER);
    const uint64_t total = (uint64_t)rows * n_cols;
    k_witness<<<(unsigned)((total + 255) / 256), 256, 0, L.stream>>>(key, t, (u64)((((unsigned __int128)1) << 64) / t), rows, cols, first_col, n_cols, W);
}
void launch_pt_encode(const Launch &L, int logn, const PlainRing &R, const u64 *values, uint32_t n_values, uint64_t value_stride, uint32_t n_cols, u64 *m) {
    LaunchScope scope(L, KID_OTHER);
    const int n = 1 << logn;
    LMN_CLIENT_DISPATCH(logn, k_pt_encode<LG><<<n_cols, client_threads(n), (size_t)padded_words(n) * 8, L.stream>>>(R, values, n_values, value_stride, m));
}
void launch_pt_lift(const Launch &L, const DeviceTables &T, int nl, const u64 *m, uint32_t n_cols, const ulonglong2 *tinv, u64 *pt) {
    LaunchScope scope(L, KID_OTHER);
    LMN_CLIENT_DISPATCH(T.logn, k_pt_lift<LG><<<dim3(nl, n_cols), client_threads(T.n), (size_t)padded_words(T.n) * 8, L.stream>>>(T.mods, T.tw_fwd, nl, m, tinv, pt));
}
void launch_encrypt(const Launch &L, const DeviceTables &T, const ChaChaKey &seed, uint32_t first_id, uint32_t domain, uint32_t n_cols, const u64 *m,
                    const ulonglong2 *tinv, const u64 *pk, u64 *cts, uint64_t ct_words) {
    LaunchScope scope(L, KID_OTHER);
    LMN_CLIENT_DISPATCH(T.logn, k_encrypt<LG><<<dim3(T.k, n_cols), client_threads(T.n), (size_t)padded_words(T.n) * 8, L.stream>>>(
                                    T.mods, T.tw_fwd, T.k, seed, first_id, domain, m, tinv, pk, cts, ct_words));
}
void launch_dump_small(const Launch &L, const DeviceTables &T, const ChaChaKey &seed, uint32_t id, uint32_t domain, u64 q, long long *out3) {
    LaunchScope scope(L, KID_OTHER);
    for (uint32_t s = 0; s < 3; s++)
        LMN_CLIENT_DISPATCH(T.logn, k_dump_small<LG><<<1, client_threads(T.n), (size_t)padded_words(T.n) * 8, L.stream>>>(seed, id, domain | s, q, out3 + (size_t)s * T.n));
}
void launch_decrypt_limbs(const Launch &L, const DeviceTables &T, const CrtConsts *Cd, int nl, const u64 *cts, uint64_t ct_words, uint32_t n_ct,
                          const u64 *sk, u64 *buf) {
    LaunchScope scope(L, KID_OTHER);
    LMN_CLIENT_DISPATCH(T.logn, k_decrypt_limb<LG><<<dim3(nl, n_ct), client_threads(T.n), (size_t)padded_words(T.n) * 8, L.stream>>>(
                                    T.mods, T.tw_inv, Cd, nl, cts, ct_words, sk, buf));
}
void launch_crt_decode(const Launch &L, int logn, const PlainRing &R, const CrtConsts *Cd, const ModDev *mods, const u64 *buf, uint32_t n_ct,
                       uint32_t n_values, u64 *values) {
    LaunchScope scope(L, KID_OTHER);
    const int n = 1 << logn;
    LMN_CLIENT_DISPATCH(logn, k_crt_decode<LG><<<n_ct, client_threads(n), (size_t)padded_words(n) * 8, L.stream>>>(R, Cd, mods, buf, n_values, values));
}

const char *kernels_client_init() {
    cudaError_t e = cudaSuccess;
    const int big = (int)padded_words(1 << 14) * 8;
#define SETSM(K) do { if (e == cudaSuccess) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, big); } while (0)
#define SETS(LG) SETSM(k_pt_encode<LG>); SETSM(k_pt_lift<LG>); SETSM(k_encrypt<LG>); SETSM(k_decrypt_limb<LG>); SETSM(k_crt_decode<LG>); SETSM(k_dump_small<LG>);
    SETS(13) SETS(14)
#undef SETS
#undef SETSM
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace lmn
