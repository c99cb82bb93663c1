// C ABI of the stages either side of the server hot path (include/lumenos_b200.h, "demo server / client side"):
// witness generation + batch encoding + public-key encryption of the matrix on the device, plaintext encoding for Prove, and the
// client's decrypt + decode.  Kernels in kernels_client.cu.
#include <algorithm>
#include <cstring>

#include "api_internal.hpp"
#include "client.cuh"

struct ClientState {
    u64 psi_t = 0;
    ulonglong2 *d_twf = nullptr, *d_twi = nullptr;     // plaintext ring tables
    uint32_t *d_perm = nullptr;
    ulonglong2 *d_tinv = nullptr;                      // [k] t^-1 mod q_i (Shoup pair)
    PlainRing R{};
    u64 *d_pk = nullptr, *d_sk = nullptr;
    u64 *d_vals = nullptr, *d_m = nullptr, *d_buf = nullptr;   // batch workspaces (values, ring-T polynomials, decrypt limbs)
    size_t vals_cap = 0, m_cap = 0, buf_cap = 0;
    CrtConsts *d_crt = nullptr;
};
void client_state_free(ClientState *s) {
    if (!s) return;
    cudaFree(s->d_twf); cudaFree(s->d_twi); cudaFree(s->d_perm); cudaFree(s->d_tinv); cudaFree(s->d_pk); cudaFree(s->d_sk);
    cudaFree(s->d_vals); cudaFree(s->d_m); cudaFree(s->d_buf); cudaFree(s->d_crt);
    delete s;
}
size_t client_state_trim(ClientState *s) {
    if (!s) return 0;
    const size_t bytes = (s->vals_cap + s->m_cap + s->buf_cap) * 8;
    cudaFree(s->d_vals); cudaFree(s->d_m); cudaFree(s->d_buf);
    s->d_vals = s->d_m = s->d_buf = nullptr; s->vals_cap = s->m_cap = s->buf_cap = 0;
    return bytes;
}
static int grow(u64 **p, size_t *cap, size_t words) {
    if (*cap >= words) return 0;
    cudaFree(*p); *p = nullptr; *cap = 0;
    cudaError_t e = cudaMalloc(p, words * 8);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(LMN_ERR_NOMEM, std::string("client workspace: ") + cudaGetErrorString(e)); }
    *cap = words;
    return 0;
}
static ChaChaKey key_words(const uint8_t *k32) {
    ChaChaKey k;
    for (int i = 0; i < 8; i++) k.w[i] = (uint32_t)k32[4 * i] | ((uint32_t)k32[4 * i + 1] << 8) | ((uint32_t)k32[4 * i + 2] << 16) | ((uint32_t)k32[4 * i + 3] << 24);
    return k;
}
static int need_ring(lmn_ctx *c) {
    if (!c->client || !c->client->d_twf) return fail(LMN_ERR_NOKEY, "plaintext ring not loaded (lmn_load_plaintext_ring)");
    return 0;
}
static const uint32_t BATCH = 256;   // columns per launch: bounds the workspaces (256 x N words each); fixed, independent of lmn_ctx_set_batch (header)

extern "C" {

int lmn_load_plaintext_ring(lmn_ctx *c, uint64_t psi_t) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    const u64 t = c->t;
    const int N = c->n, logn = c->logn;
    if (t >= (1ull << 59) || t > ~0ull / (3 * (u64)logn + 8)) return fail(LMN_ERR_UNSUPPORTED, "plaintext modulus too large for the lazy transforms");
    if (h_powmod(psi_t, (u64)N, t) != t - 1) return fail(LMN_ERR_INVALID, "psi_t is not a primitive 2N-th root of unity modulo t");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (const char *e = kernels_client_init()) return fail(LMN_ERR_CUDA, std::string("kernel attribute setup: ") + e);
    if (!c->client) c->client = new ClientState();
    ClientState *S = c->client;
    std::vector<ulonglong2> f, iv;
    build_twiddles(t, psi_t, logn, f, iv);
    // lattigo bgv: slot i of the first row sits at bitrev((5^i mod 2N) >> 1), slot i + N/2 at N - 1 - that
    std::vector<uint32_t> perm(N);
    u64 pw = 1, mask = 2 * (u64)N - 1;
    for (int i = 0, j = N / 2; i < N / 2; i++, j++) {
        const uint32_t pos = (uint32_t)h_bitrev(pw >> 1, logn);
        perm[i] = pos; perm[j] = (uint32_t)N - pos - 1;
        pw = (pw * 5) & mask;
    }
    std::vector<ulonglong2> tinv(c->k);
    for (int i = 0; i < c->k; i++) { const u64 q = c->q[i], w = h_invmod(t % q, q); tinv[i] = make_ulonglong2(w, h_shoup(w, q)); }
    if (!S->d_twf) {
        CK(cudaMalloc(&S->d_twf, sizeof(ulonglong2) * N)); CK(cudaMalloc(&S->d_twi, sizeof(ulonglong2) * N));
        CK(cudaMalloc(&S->d_perm, sizeof(uint32_t) * N)); CK(cudaMalloc(&S->d_tinv, sizeof(ulonglong2) * c->k));
    }
    CK(cudaMemcpy(S->d_twf, f.data(), sizeof(ulonglong2) * N, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(S->d_twi, iv.data(), sizeof(ulonglong2) * N, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(S->d_perm, perm.data(), sizeof(uint32_t) * N, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(S->d_tinv, tinv.data(), sizeof(ulonglong2) * c->k, cudaMemcpyHostToDevice));
    S->psi_t = psi_t;
    build_mod(S->R.mod, t, N);
    S->R.tw_fwd = S->d_twf; S->R.tw_inv = S->d_twi; S->R.perm = S->d_perm;
    return 0;
}

static int load_key(lmn_ctx *c, const uint64_t *host, size_t limbs, bool secret) {
    if (!c || !host) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);           // c->client is created and read under this lock only
    cudaSetDevice(c->device);
    if (!c->client) c->client = new ClientState();
    u64 **dst = secret ? &c->client->d_sk : &c->client->d_pk;
    const size_t bytes = limbs * (size_t)c->n * 8;
    if (!*dst) CK(cudaMalloc(dst, bytes));
    CK(cudaMemcpy(*dst, host, bytes, cudaMemcpyHostToDevice));
    return 0;
}
int lmn_load_public_key(lmn_ctx *c, const uint64_t *pk) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    return load_key(c, pk, 2 * (size_t)c->k, false);
}
int lmn_load_secret_key(lmn_ctx *c, const uint64_t *sk) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    return load_key(c, sk, (size_t)c->k, true);
}

int lmn_witness_columns(lmn_ctx *c, const uint8_t *chacha_key32, uint32_t rows, uint32_t cols, uint32_t first_col, uint32_t n_cols, uint64_t *host_out) {
    if (!c || !chacha_key32 || !host_out || rows == 0 || n_cols == 0) return fail(LMN_ERR_INVALID, "null argument");
    if ((uint64_t)first_col + n_cols > cols) return fail(LMN_ERR_INVALID, "column range out of bounds");
    if (((uint64_t)rows * cols + 7) / 8 > (1ull << 32)) return fail(LMN_ERR_UNSUPPORTED, "witness longer than one ChaCha20 (32-bit counter) stream");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (!c->client) c->client = new ClientState();
    ClientState *S = c->client;
    const ChaChaKey key = key_words(chacha_key32);
    for (uint32_t j0 = 0; j0 < n_cols; j0 += BATCH) {
        const uint32_t nj = std::min(BATCH, n_cols - j0);
        if (int rc = grow(&S->d_vals, &S->vals_cap, (size_t)nj * rows)) return rc;
        launch_witness(c->L(), key, c->t, rows, cols, first_col + j0, nj, S->d_vals);
        CK(cudaMemcpyAsync(host_out + (size_t)j0 * rows, S->d_vals, (size_t)nj * rows * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

int lmn_encrypt_columns(lmn_ctx *c, const uint64_t *values_colmajor_host, const uint8_t *chacha_key32, uint32_t rows, uint32_t cols_total,
                        uint32_t first_col, uint32_t count, const uint8_t *enc_seed32, lmn_store *matrix) {
    if (!c || !enc_seed32 || !matrix || (!values_colmajor_host && !chacha_key32)) return fail(LMN_ERR_INVALID, "null argument");
    if (matrix->ctx != c || (int)matrix->level != c->k - 1 || matrix->n != (uint32_t)c->n) return fail(LMN_ERR_INVALID, "the matrix store holds max-level ciphertexts of the context ring");
    if (count == 0 || matrix->count < count || (uint64_t)first_col + count > cols_total) return fail(LMN_ERR_INVALID, "column range out of bounds");
    if (rows == 0 || rows > (uint32_t)c->n) return fail(LMN_ERR_INVALID, "a column holds between 1 and N slot values");
    if (!values_colmajor_host && ((uint64_t)rows * cols_total + 7) / 8 > (1ull << 32)) return fail(LMN_ERR_UNSUPPORTED, "witness longer than one ChaCha20 (32-bit counter) stream");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_ring(c)) return rc;
    ClientState *S = c->client;
    if (!S->d_pk) return fail(LMN_ERR_NOKEY, "public key not loaded");
    const ChaChaKey seed = key_words(enc_seed32);
    ChaChaKey wkey{}; if (!values_colmajor_host) wkey = key_words(chacha_key32);
    const size_t N = c->n;
    Launch L = c->L();
    for (uint32_t j0 = 0; j0 < count; j0 += BATCH) {
        const uint32_t nj = std::min(BATCH, count - j0);
        if (int rc = grow(&S->d_vals, &S->vals_cap, (size_t)nj * rows)) return rc;
        if (int rc = grow(&S->d_m, &S->m_cap, (size_t)nj * N)) return rc;
        if (values_colmajor_host) {
            CK(cudaMemcpyAsync(S->d_vals, values_colmajor_host + (size_t)j0 * rows, (size_t)nj * rows * 8, cudaMemcpyHostToDevice, c->stream));
            CK(cudaStreamSynchronize(c->stream));     // the caller's buffer may be pageable and is free again when the call returns
        } else {
            launch_witness(L, wkey, c->t, rows, cols_total, first_col + j0, nj, S->d_vals);
        }
        launch_pt_encode(L, c->logn, S->R, S->d_vals, rows, rows, nj, S->d_m);
        launch_encrypt(L, c->T, seed, first_col + j0, ENC_DOMAIN_COLUMN, nj, S->d_m, S->d_tinv, S->d_pk, matrix->d + (size_t)j0 * matrix->words(), matrix->words());
    }
    CK(cudaGetLastError());
    return 0;
}

int lmn_encrypt_zero(lmn_ctx *c, const uint8_t *enc_seed32, uint32_t stream_id, lmn_store *out, uint32_t out_idx) {
    if (!c || !enc_seed32 || !out) return fail(LMN_ERR_INVALID, "null argument");
    if (out->ctx != c || (int)out->level != c->k - 1 || out->n != (uint32_t)c->n || out_idx >= out->count) return fail(LMN_ERR_INVALID, "output must be a max-level ciphertext slot of the context ring");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_ring(c)) return rc;
    ClientState *S = c->client;
    if (!S->d_pk) return fail(LMN_ERR_NOKEY, "public key not loaded");
    launch_encrypt(c->L(), c->T, key_words(enc_seed32), stream_id, ENC_DOMAIN_ZERO, 1, nullptr, S->d_tinv, S->d_pk, out->d + (size_t)out_idx * out->words(), out->words());
    CK(cudaGetLastError());
    return 0;
}

// test hook (tests/test_gpu_client.py): the sampler's output for one column, so that its distribution can be checked directly
int lmn_test_encryption_samples(lmn_ctx *c, const uint8_t *enc_seed32, uint32_t id, int zero_column, int64_t *out3_host) {
    if (!c || !enc_seed32 || !out3_host) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_ring(c)) return rc;
    DevScratch d;
    const size_t bytes = 3 * (size_t)c->n * sizeof(long long);
    if (d.alloc(bytes) != cudaSuccess) { cudaGetLastError(); return fail(LMN_ERR_NOMEM, "sample dump buffer"); }
    launch_dump_small(c->L(), c->T, key_words(enc_seed32), id, zero_column ? ENC_DOMAIN_ZERO : ENC_DOMAIN_COLUMN, c->q[0], d.as<long long>());
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out3_host, d.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int lmn_encode_plaintext(lmn_ctx *c, const uint64_t *values, uint32_t n_values, uint64_t *pt_host) {
    if (!c || !values || !pt_host) return fail(LMN_ERR_INVALID, "null argument");
    if (n_values == 0 || n_values > (uint32_t)c->n) return fail(LMN_ERR_INVALID, "a plaintext holds between 1 and N slot values");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_ring(c)) return rc;
    ClientState *S = c->client;
    const size_t N = c->n, k = c->k;
    if (int rc = grow(&S->d_vals, &S->vals_cap, n_values)) return rc;
    if (int rc = grow(&S->d_m, &S->m_cap, N)) return rc;
    if (int rc = grow(&S->d_buf, &S->buf_cap, k * N)) return rc;
    if (int rc = stage_begin(c, (size_t)n_values * 8 + 16)) return rc;
    if (int rc = stage_copy(c, S->d_vals, values, (size_t)n_values * 8)) return rc;
    if (int rc = stage_end(c)) return rc;
    Launch L = c->L();
    launch_pt_encode(L, c->logn, S->R, S->d_vals, n_values, n_values, 1, S->d_m);
    launch_pt_lift(L, c->T, (int)k, S->d_m, 1, S->d_tinv, S->d_buf);
    CK(cudaMemcpyAsync(pt_host, S->d_buf, k * N * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int lmn_decrypt_decode(lmn_ctx *c, const lmn_store *s, uint32_t first, uint32_t count, uint64_t scale, uint32_t n_values, uint64_t *values_host) {
    if (!c || !s || !values_host) return fail(LMN_ERR_INVALID, "null argument");
    if (s->ctx != c || s->n != (uint32_t)c->n) return fail(LMN_ERR_INVALID, "decryption needs ciphertexts of the context ring");
    if (count == 0 || (uint64_t)first + count > s->count) return fail(LMN_ERR_INVALID, "ciphertext range out of bounds");
    if (n_values == 0 || n_values > (uint32_t)c->n) return fail(LMN_ERR_INVALID, "between 1 and N slot values per ciphertext");
    if (scale % c->t == 0) return fail(LMN_ERR_INVALID, "Scale must be invertible modulo t");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_ring(c)) return rc;
    ClientState *S = c->client;
    if (!S->d_sk) return fail(LMN_ERR_NOKEY, "secret key not loaded");
    const int nl = (int)s->level + 1;
    if (nl > CRT_MAX) return fail(LMN_ERR_UNSUPPORTED, "level too high");
    const u64 t = c->t;
    const size_t N = c->n;
    CrtConsts K; memset(&K, 0, sizeof(K));
    K.nl = nl;
    u64 Qm = 1 % t;
    for (int i = 0; i < nl; i++) {
        const u64 qi = c->q[i];
        K.half[i] = (qi - 1) >> 1;
        for (int j = 0; j < i; j++) { const u64 w = h_invmod(c->q[j] % qi, qi); K.inv_w[i][j] = w; K.inv_ws[i][j] = h_shoup(w, qi); }
        K.q_mod_t[i] = h_mform(qi % t, t);
        Qm = h_mulmod(Qm, qi % t, t);
        const u64 w = h_mulmod(h_invmod((u64)N % qi, qi), t % qi, qi);
        K.t_w[i] = w; K.t_ws[i] = h_shoup(w, qi);
    }
    K.half_q_mod_t = h_mulmod(h_submod(Qm, 1 % t, t), h_invmod(2, t), t);
    K.sinv_w = h_invmod(scale % t, t); K.sinv_ws = h_shoup(K.sinv_w, t);
    if (!S->d_crt) CK(cudaMalloc(&S->d_crt, sizeof(CrtConsts)));
    if (int rc = stage_begin(c, sizeof(CrtConsts) + 16)) return rc;
    if (int rc = stage_copy(c, S->d_crt, &K, sizeof(CrtConsts))) return rc;
    if (int rc = stage_end(c)) return rc;
    Launch L = c->L();
    for (uint32_t j0 = 0; j0 < count; j0 += BATCH) {
        const uint32_t nj = std::min(BATCH, count - j0);
        if (int rc = grow(&S->d_buf, &S->buf_cap, (size_t)nj * nl * N)) return rc;
        if (int rc = grow(&S->d_vals, &S->vals_cap, (size_t)nj * n_values)) return rc;
        launch_decrypt_limbs(L, c->T, S->d_crt, nl, s->d + (size_t)(first + j0) * s->words(), s->words(), nj, S->d_sk, S->d_buf);
        launch_crt_decode(L, c->logn, S->R, S->d_crt, c->d_mods, S->d_buf, nj, n_values, S->d_vals);
        CK(cudaMemcpyAsync(values_host + (size_t)j0 * n_values, S->d_vals, (size_t)nj * n_values * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    CK(cudaGetLastError());
    return 0;
}

}  // extern "C"
