// C ABI of liblumenos_b200 (include/lumenos_b200.h): context, device-resident store, plan upload and the stage
// drivers that sequence the kernels of kernels.cu on the context's CUDA stream.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "api_internal.hpp"
#include "commit.cuh"
#include "plan.hpp"

static thread_local std::string g_err;
int lmn_fail(int code, const std::string &msg) { g_err = msg; return code; }

struct lmn_plan {
    lmn_ctx *ctx;
    EncodePlan pl;
    uint16_t *d_scale_after = nullptr;
    uint32_t *d_out_pos = nullptr;
    ulonglong2 *d_scal = nullptr;   // [k][nscal]
    uint32_t cols = 0, rho_inv = 0;
};

// ------------------------------------------------------------------------------------------------------------
static bool supported_logn(int l) { return l == 6 || l == 8 || l == 10 || l == 12 || l == 13 || l == 14; }

void build_mod(ModDev &m, u64 q, int n) {
    memset(&m, 0, sizeof(m));
    m.q = q; m.qinv = h_qinv64(q); m.mu = (u64)((((u128)1) << 64) / q);
    m.ninv = h_invmod((u64)n % q, q); m.ninvs = h_shoup(m.ninv, q); m.q3 = 3 * q;
}
void build_twiddles(u64 q, u64 psi, int logn, std::vector<ulonglong2> &fwd, std::vector<ulonglong2> &inv) {
    int n = 1 << logn;
    fwd.resize(n); inv.resize(n);
    u64 psiinv = h_invmod(psi, q), a = 1, b = 1;
    for (int j = 0; j < n; j++) {
        int r = (int)h_bitrev((u64)j, logn);
        fwd[r] = make_ulonglong2(a, h_shoup(a, q));
        inv[r] = make_ulonglong2(b, h_shoup(b, q));
        a = h_mulmod(a, psi, q); b = h_mulmod(b, psiinv, q);
    }
}

// s(d, gm) for every max-level digit d and every modulus gm outside it: the smallest e in [0, 16] for which all
// 2^e (Q_d/q_a) mod m are below 2^31 in signed form gives s = 64 - e (fast 64x32-bit base conversion); otherwise s = 64
// (generic 64x64-bit path).  Single-prime digits need no Montgomery step: s = 0.
static void build_key_scale(lmn_ctx *c) {
    const int k = c->k, np = c->np, KP = k + np, beta = (k + np - 1) / np;
    memset(c->key_scale, 0, sizeof(c->key_scale));
    for (int d = 0; d < beta; d++) {
        const int s0 = d * np, nd = std::min(np, k - s0);
        for (int gm = 0; gm < KP; gm++) {
            if (gm >= s0 && gm < s0 + nd) continue;
            if (nd < 2) { c->key_scale[d][gm] = 0; continue; }
            const u64 m = c->mod(gm);
            int best = 64;
            for (int e = 0; e <= 16 && best == 64; e++) {
                bool ok = true;
                for (int a = 0; a < nd && ok; a++) {
                    u64 pr = h_powmod(2, (u64)e, m);
                    for (int b = 0; b < nd; b++) if (b != a) pr = h_mulmod(pr, c->q[s0 + b] % m, m);
                    u64 mag = pr <= m / 2 ? pr : m - pr;
                    ok = mag < (1ull << 31);
                }
                if (ok) best = 64 - e;
            }
            c->key_scale[d][gm] = best;
        }
    }
}

static void build_ks_consts(const lmn_ctx *c, int level, KsConsts &K) {
    memset(&K, 0, sizeof(K));
    const int nq = level + 1, np = c->np, k = c->k;
    K.nq = nq; K.np = np; K.k = k; K.beta = (nq + np - 1) / np;
    for (int d = 0; d < K.beta; d++) { K.dig_start[d] = d * np; K.dig_size[d] = std::min(np, nq - d * np); }
    auto digit_prod = [&](int d, int skip, u64 m) {
        u64 r = 1 % m;
        for (int a = 0; a < K.dig_size[d]; a++) { int i = K.dig_start[d] + a; if (i != skip) r = h_mulmod(r, c->q[i] % m, m); }
        return r;
    };
    for (int d = 0; d < K.beta; d++) {
        for (int a = 0; a < K.dig_size[d]; a++) {
            int i = K.dig_start[d] + a; u64 qi = c->q[i];
            u64 ninv = h_invmod((u64)c->n % qi, qi);
            u64 w = K.dig_size[d] > 1 ? h_mulmod(ninv, h_invmod(digit_prod(d, i, qi), qi), qi) : ninv;
            K.intt_w[i] = w; K.intt_ws[i] = h_shoup(w, qi);
        }
        for (int tix = 0; tix < nq + np; tix++) {
            const int gm = ks_mod_index(tix, nq, k);
            if (gm >= K.dig_start[d] && gm < K.dig_start[d] + K.dig_size[d]) continue;   // own limb
            const u64 m = c->mod(gm);
            const int nd = K.dig_size[d], s = c->key_scale[d][gm];
            // r must equal value * 2^-s (mod m) because the key rows of (d, gm) carry 2^s
            const u64 ginv = h_invmod(h_powmod(2, (u64)s, m), m);
            K.mu_g_w[d][tix] = ginv; K.mu_g_ws[d][tix] = h_shoup(ginv, m);
            if (nd < 2) continue;
            const u64 scale = h_powmod(2, (u64)(64 - s), m);                 // c_a = 2^(64-s) Q_d/q_a mod m
            const bool same_as_max = nd == std::min(np, k - d * np);
            u64 cfull[MAX_P]; bool small = same_as_max;
            u128 maxneg = 0;
            for (int a = 0; a < nd; a++) {
                cfull[a] = h_mulmod(digit_prod(d, K.dig_start[d] + a, m), scale, m);
                u64 mag = cfull[a] <= m / 2 ? cfull[a] : m - cfull[a];
                if (mag >= (1ull << 31)) small = false;
            }
            const u64 kconst = [&] { u64 v = h_mulmod(h_mulmod((u64)(nd - 1) % m, digit_prod(d, -1, m), m), scale, m); return v ? m - v : 0; }();
            if (small) {
                K.mu_fast[d][tix] = 1;
                for (int a = 0; a < nd; a++) {
                    bool neg = cfull[a] > m / 2;
                    u64 mag = neg ? m - cfull[a] : cfull[a];
                    K.mu_mag[d][tix][a] = (uint32_t)mag;
                    if (neg) { K.mu_sgn[d][tix] |= 1u << a; maxneg += (u128)((u64)nd * c->q[K.dig_start[d] + a]) * mag; }
                }
                u128 kk = (u128)kconst + (u128)(maxneg / m + 1) * m;
                K.mu_k_lo[d][tix] = (u64)kk; K.mu_k_hi[d][tix] = (u64)(kk >> 64);
            } else {
                K.mu_fast[d][tix] = 0;
                for (int a = 0; a < nd; a++) K.mu_c[d][tix][a] = cfull[a];
                K.mu_k_lo[d][tix] = kconst; K.mu_k_hi[d][tix] = 0;
            }
        }
    }
    auto p_prod = [&](int skip, u64 m) {
        u64 r = 1 % m;
        for (int a = 0; a < np; a++) if (a != skip) r = h_mulmod(r, c->p[a] % m, m);
        return r;
    };
    for (int j = 0; j < np; j++) {
        u64 pj = c->p[j], ninv = h_invmod((u64)c->n % pj, pj), ci = h_invmod(p_prod(j, pj), pj);
        u64 w = h_mulmod(ninv, ci, pj);
        K.pin_w[j] = w; K.pin_ws[j] = h_shoup(w, pj);
        K.pin_add[j] = c->round_nearest ? h_mulmod((pj - 1) >> 1, ci, pj) : 0;
    }
    for (int i = 0; i < nq; i++) {
        u64 qi = c->q[i], Pm = p_prod(-1, qi);
        for (int a = 0; a < np; a++) K.md_c[i][a] = h_mform(p_prod(a, qi), qi);
        u64 half = c->round_nearest ? h_mulmod(h_submod(Pm, 1, qi), h_invmod(2, qi), qi) : 0;
        u64 v = h_addmod(h_mulmod((u64)(np - 1) % qi, Pm, qi), half, qi);
        K.md_neg[i] = v ? qi - v : 0;
        u64 pinv = h_invmod(Pm, qi);
        K.pinv_w[i] = pinv; K.pinv_ws[i] = h_shoup(pinv, qi);
        K.pmod_w[i] = Pm; K.pmod_ws[i] = h_shoup(Pm, qi);
    }
}

// staged host -> device copies on the context's stream: stage_begin (waits until the previous user of the staging buffer has been
// copied out, grows it), any number of stage_copy, stage_end.  Caller holds the context lock.
int stage_begin(lmn_ctx *c, size_t total) {
    if (c->stage_pending) { CK(cudaEventSynchronize(c->stage_ev)); c->stage_pending = false; }
    if (!c->stage_ev) CK(cudaEventCreateWithFlags(&c->stage_ev, cudaEventDisableTiming));
    if (c->h_stage_cap < total) {
        if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->limb_stage && c->limb_stage_free) c->limb_stage_free(c->limb_stage);
        c->h_stage = nullptr; c->h_stage_cap = 0;
        CK(cudaHostAlloc((void **)&c->h_stage, total, cudaHostAllocDefault));
        c->h_stage_cap = total;
    }
    c->h_stage_cur = 0;
    return 0;
}
int stage_copy(lmn_ctx *c, void *dst, const void *src, size_t bytes) {
    if (c->h_stage_cur + bytes > c->h_stage_cap) return fail(LMN_ERR_INVALID, "staging buffer overflow");
    uint8_t *s = c->h_stage + c->h_stage_cur;
    memcpy(s, src, bytes);
    c->h_stage_cur += (bytes + 15) & ~(size_t)15;
    CK(cudaMemcpyAsync(dst, s, bytes, cudaMemcpyHostToDevice, c->stream));
    return 0;
}
int stage_end(lmn_ctx *c) {
    CK(cudaEventRecord(c->stage_ev, c->stream));
    c->stage_pending = true;
    return 0;
}

static int get_ks(lmn_ctx *c, int level, const KsConsts **host, const KsConsts **dev) {
    auto it = c->ks.find(level);
    if (it == c->ks.end()) {
        KsConsts K; build_ks_consts(c, level, K);
        KsConsts *d = nullptr;
        CK(cudaMalloc(&d, sizeof(KsConsts)));
        CK(cudaMemcpy(d, &K, sizeof(KsConsts), cudaMemcpyHostToDevice));
        it = c->ks.emplace(level, std::make_pair(K, d)).first;
    }
    *host = &it->second.first; *dev = it->second.second;
    return 0;
}
static int get_rs(lmn_ctx *c, int level, const RescaleConsts **dev) {
    auto it = c->rs.find(level);
    if (it == c->rs.end()) {
        RescaleConsts R; memset(&R, 0, sizeof(R));
        u64 ql = c->q[level];
        R.half = (ql - 1) >> 1;
        for (int i = 0; i < level; i++) {
            u64 qi = c->q[i];
            R.half_mod[i] = R.half % qi;
            u64 inv = h_invmod(ql % qi, qi);
            R.qlinv_w[i] = inv; R.qlinv_ws[i] = h_shoup(inv, qi);
        }
        RescaleConsts *d = nullptr;
        CK(cudaMalloc(&d, sizeof(R)));
        CK(cudaMemcpy(d, &R, sizeof(R), cudaMemcpyHostToDevice));
        it = c->rs.emplace(level, d).first;
    }
    *dev = it->second;
    return 0;
}

static int ensure_workspace(lmn_ctx *c, uint32_t jobs) {
    if (c->ws_jobs >= jobs) return 0;
    auto drop = [&]() {
        cudaFree(c->ws_col); cudaFree(c->ws_col2); cudaFree(c->ws_y); cudaFree(c->ws_acc); cudaFree(c->ws_mu);
        c->ws_col = c->ws_col2 = c->ws_y = c->ws_acc = c->ws_mu = nullptr; c->ws_jobs = 0;
    };
    drop();
    size_t N = c->n, k = c->k, KP = c->k + c->np;
    cudaError_t e = cudaMalloc(&c->ws_col, jobs * 2 * k * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&c->ws_col2, jobs * 2 * k * N * 8);   // the block pipeline rotates out of place
    if (e == cudaSuccess) e = cudaMalloc(&c->ws_y, jobs * k * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&c->ws_acc, jobs * 2 * KP * N * 8);
    if (e == cudaSuccess) e = cudaMalloc(&c->ws_mu, jobs * (size_t)((c->k + c->np - 1) / c->np) * KP * N * 8);
    if (e != cudaSuccess) {
        drop();
        cudaGetLastError();   // an out-of-memory cudaMalloc leaves a sticky-looking last error behind; the context stays usable
        return fail(e == cudaErrorMemoryAllocation ? LMN_ERR_NOMEM : LMN_ERR_CUDA, std::string("workspace for ") + std::to_string(jobs) + " jobs: " + cudaGetErrorString(e));
    }
    c->ws_jobs = jobs;
    return 0;
}
static KsJobs jobs_view(lmn_ctx *c, uint32_t first, uint32_t jobs) {
    size_t N = c->n, k = c->k, KP = c->k + c->np;
    KsJobs J;
    J.col_words = 2 * k * N; J.y_words = k * N; J.acc_words = 2 * KP * N; J.mu_words = (size_t)((c->k + c->np - 1) / c->np) * KP * N;
    J.col = c->ws_col + first * J.col_words; J.col2 = c->ws_col2 + first * J.col_words; J.y = c->ws_y + first * J.y_words; J.acc = c->ws_acc + first * J.acc_words;
    J.mu = c->ws_mu + first * J.mu_words;
    J.jobs = (int)jobs;
    return J;
}
static u64 galois_element(const lmn_ctx *c, uint32_t rot) {  // 5^rot mod 2N
    u64 mask = 2 * (u64)c->n - 1, g = 1, b = 5;
    uint32_t e = rot & (uint32_t)(c->n / 2 - 1);
    while (e) { if (e & 1) g = (g * b) & mask; b = (b * b) & mask; e >>= 1; }
    return g;
}

// ------------------------------------------------------------------------------------------------------------
extern "C" {

const char *lmn_last_error(void) { return g_err.c_str(); }

int lmn_ctx_create(const lmn_params *P, int device, lmn_ctx **out) {
    if (!P || !out) return fail(LMN_ERR_INVALID, "null argument");
    if (!supported_logn((int)P->log_n)) return fail(LMN_ERR_UNSUPPORTED, "log_n must be one of 6, 8, 10, 12, 13, 14");
    if (P->num_q < 1 || P->num_p < 1 || P->num_p > MAX_P || P->num_q + P->num_p > MAX_LIMBS ||
        (P->num_q + P->num_p - 1) / P->num_p > MAX_DIGITS)
        return fail(LMN_ERR_UNSUPPORTED, "unsupported modulus chain shape");
    const u64 two_n = 2ull << P->log_n;
    for (uint32_t i = 0; i < P->num_q + P->num_p; i++) {
        u64 m = i < P->num_q ? P->q[i] : P->p[i - P->num_q];
        u64 psi = i < P->num_q ? P->psi_q[i] : P->psi_p[i - P->num_q];
        // lazy transforms: values grow by 3q per forward level on top of an 8q entry bound (ntt_r16.cuh) and must stay below 2^64
        const u64 lazy_cap = ~0ull / (3 * (u64)P->log_n + 8);
        if (m >= (1ull << 59) || m > lazy_cap || m < (1ull << 20) || m % two_n != 1)
            return fail(LMN_ERR_INVALID, "modulus must be an NTT-friendly prime below min(2^59, 2^64 / (3 log N + 8))");
        if (h_powmod(psi, two_n / 2, m) != m - 1) return fail(LMN_ERR_INVALID, "psi is not a primitive 2N-th root of unity");
    }
    if (P->t % two_n != 1) return fail(LMN_ERR_INVALID, "plaintext modulus must be 1 mod 2N");  // ref: fhe/bfv.go:136-140
    if (device >= 0) CK(cudaSetDevice(device)); else CK(cudaGetDevice(&device));
    int major = 0;
    CK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major < 10) return fail(LMN_ERR_UNSUPPORTED, "liblumenos_b200 requires an sm_100a (B200) device");
    if (const char *e = kernels_init()) return fail(LMN_ERR_CUDA, std::string("kernel attribute setup: ") + e);

    lmn_ctx *c = new lmn_ctx();
    c->device = device; c->logn = (int)P->log_n; c->n = 1 << c->logn; c->k = (int)P->num_q; c->np = (int)P->num_p; c->t = P->t;
    c->q.assign(P->q, P->q + P->num_q); c->p.assign(P->p, P->p + P->num_p);
    if (P->scalar_centered != 0xffffffffu) c->scalar_centered = P->scalar_centered != 0;
    if (P->moddown_round_nearest != 0xffffffffu) c->round_nearest = P->moddown_round_nearest != 0;
    if (P->innersum_rowswap_last != 0xffffffffu) c->rowswap_last = P->innersum_rowswap_last != 0;
    const int KP = c->k + c->np, N = c->n;
    std::vector<ModDev> mods(KP);
    std::vector<ulonglong2> twf((size_t)KP * N), twi((size_t)KP * N), f, iv;
    for (int i = 0; i < KP; i++) {
        u64 m = c->mod(i), psi = i < c->k ? P->psi_q[i] : P->psi_p[i - c->k];
        build_mod(mods[i], m, N);
        build_twiddles(m, psi, c->logn, f, iv);
        std::copy(f.begin(), f.end(), twf.begin() + (size_t)i * N);
        std::copy(iv.begin(), iv.end(), twi.begin() + (size_t)i * N);
    }
    cudaError_t e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_mods, sizeof(ModDev) * KP);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_twf, sizeof(ulonglong2) * KP * N);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_twi, sizeof(ulonglong2) * KP * N);
    if (e == cudaSuccess) e = cudaMalloc(&c->ws_pt, sizeof(u64) * 4 * c->k * N);
    if (e == cudaSuccess) e = cudaMalloc(&c->ws_scalar, sizeof(ulonglong2) * MAX_LIMBS);
    if (e == cudaSuccess) e = cudaMemcpy(c->d_mods, mods.data(), sizeof(ModDev) * KP, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(c->d_twf, twf.data(), sizeof(ulonglong2) * KP * N, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(c->d_twi, twi.data(), sizeof(ulonglong2) * KP * N, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { int rc = fail(LMN_ERR_CUDA, std::string("context setup: ") + cudaGetErrorString(e)); lmn_ctx_destroy(c); return rc; }
    c->stream = c->own_stream;
    c->T = DeviceTables{c->logn, c->n, c->k, c->np, c->d_mods, c->d_twf, c->d_twi, nullptr, nullptr};
    if (blk_blocks(c->logn)) {   // per-block twiddle tables of the 2048-point block pipeline (same words as d_twf / d_twi, block order)
        e = cudaMalloc(&c->d_twbf, sizeof(ulonglong2) * KP * N);
        if (e == cudaSuccess) e = cudaMalloc(&c->d_twbi, sizeof(ulonglong2) * KP * N);
        if (e == cudaSuccess) { launch_build_blk_tables(c->stream, c->T, c->d_twbf, c->d_twbi); e = cudaStreamSynchronize(c->stream); }
        if (e != cudaSuccess) { int rc = fail(LMN_ERR_CUDA, std::string("context setup: ") + cudaGetErrorString(e)); lmn_ctx_destroy(c); return rc; }
        c->T.twb_fwd = c->d_twbf; c->T.twb_inv = c->d_twbi;
    }
    build_key_scale(c);
    {
        const int beta = (c->k + c->np - 1) / c->np;
        std::vector<u64> fm((size_t)beta * KP, 0);
        for (int d = 0; d < beta; d++)
            for (int gm = 0; gm < KP; gm++) {
                const int s0 = d * c->np, nd = std::min(c->np, c->k - s0);
                if (gm >= s0 && gm < s0 + nd) continue;
                if (c->key_scale[d][gm] == 0) continue;
                const u64 m = c->mod(gm);
                fm[(size_t)d * KP + gm] = h_mform(h_powmod(2, (u64)c->key_scale[d][gm], m), m);
            }
        e = cudaMalloc(&c->d_key_fm, fm.size() * 8);
        if (e == cudaSuccess) e = cudaMemcpy(c->d_key_fm, fm.data(), fm.size() * 8, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { int rc = fail(LMN_ERR_CUDA, std::string("context setup: ") + cudaGetErrorString(e)); lmn_ctx_destroy(c); return rc; }
    }
    *out = c;
    return 0;
}

// Everything the context pools between calls and can rebuild on demand goes back to the driver: the key-switch / rescale batch
// workspaces, the serialised-leaf buffer of the Merkle build (n leaves x 512 KiB at log N 14: 4 GiB at config D on one GPU), the
// index list, the exchange staging buffers of the communicator and the client-side batch buffers.  Keys, tables, plans and stores stay.
int lmn_ctx_trim(lmn_ctx *c, size_t *released_bytes) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    std::lock_guard<std::mutex> g(c->mu);
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaStreamSynchronize(c->copy_stream));
    const size_t N = c->n, k = c->k, KP = c->k + c->np, beta = (size_t)((c->k + c->np - 1) / c->np);
    size_t bytes = (size_t)c->ws_jobs * (2 * 2 * k + k + 2 * KP + beta * KP) * N * 8 + c->ws_ser_cap + c->ws_idx_cap * sizeof(uint32_t);
    cudaFree(c->ws_col); cudaFree(c->ws_col2); cudaFree(c->ws_y); cudaFree(c->ws_acc); cudaFree(c->ws_mu); cudaFree(c->ws_ser); cudaFree(c->ws_idx);
    c->ws_col = c->ws_col2 = c->ws_y = c->ws_acc = c->ws_mu = c->ws_ser = nullptr; c->ws_idx = nullptr;
    c->ws_jobs = 0; c->ws_ser_cap = 0; c->ws_idx_cap = 0;
    bytes += comm_state_trim(c->comm);
    bytes += client_state_trim(c->client);
    if (released_bytes) *released_bytes = bytes;
    return 0;
}
void lmn_ctx_destroy(lmn_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    client_state_free(c->client);
    comm_state_free(c->comm);
    for (auto &kv : c->ks) cudaFree(kv.second.second);
    for (auto &kv : c->rs) cudaFree(kv.second);
    for (auto &kv : c->rs_chain) cudaFree(kv.second.second);
    for (auto &kv : c->galois) cudaFree(kv.second);
    cudaFree(c->d_mods); cudaFree(c->d_twf); cudaFree(c->d_twi); cudaFree(c->d_twbf); cudaFree(c->d_twbi); cudaFree(c->ws_pt); cudaFree(c->ws_scalar);
    cudaFree(c->ws_col); cudaFree(c->ws_col2); cudaFree(c->ws_y); cudaFree(c->ws_acc); cudaFree(c->ws_idx); cudaFree(c->ws_ser);
    cudaFree(c->d_rs_evk); cudaFree(c->d_small_mod); cudaFree(c->d_small_tw); cudaFree(c->d_key_fm); cudaFree(c->ws_mu);
    for (int i = 0; i < KID_COUNT; i++) for (auto &sp : c->prof.spans[i]) { cudaEventDestroy(sp.first); cudaEventDestroy(sp.second); }
    for (auto e : c->prof.pool) cudaEventDestroy(e);
    for (cudaEvent_t e : c->copy_events) cudaEventDestroy(e);
    if (c->stage_ev) cudaEventDestroy(c->stage_ev);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}
int lmn_ctx_set_stream(lmn_ctx *c, void *s) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    std::lock_guard<std::mutex> g(c->mu);
    c->stream = s ? (cudaStream_t)s : c->own_stream;
    return 0;
}
int lmn_ctx_synchronize(lmn_ctx *c) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
uint64_t lmn_ctx_launch_count(const lmn_ctx *c) { return c ? c->launches : 0; }
int lmn_ctx_set_batch(lmn_ctx *c, uint32_t jobs) {
    if (!c || jobs < 2 || jobs > 65535) return fail(LMN_ERR_INVALID, "batch must be in [2, 65535]");
    std::lock_guard<std::mutex> g(c->mu);
    c->batch = jobs;
    return 0;
}

// ---------------- store ----------------
static int store_new(lmn_ctx *c, void *wrap, uint32_t n_ct, uint32_t level, uint32_t ring_n, lmn_store **out) {
    if (!c || !out || n_ct == 0) return fail(LMN_ERR_INVALID, "bad store arguments");
    if ((int)level >= c->k) return fail(LMN_ERR_INVALID, "store level exceeds the modulus chain");
    lmn_store *s = new lmn_store();
    s->ctx = c; s->count = n_ct; s->level = level; s->n = ring_n;
    if (wrap) { s->d = (u64 *)wrap; s->owned = false; }
    else {
        cudaSetDevice(c->device);
        cudaError_t e = cudaMalloc(&s->d, s->words() * 8 * n_ct);
        if (e != cudaSuccess) { delete s; return fail(LMN_ERR_NOMEM, std::string("store allocation: ") + cudaGetErrorString(e)); }
    }
    *out = s;
    return 0;
}
int lmn_store_create(lmn_ctx *c, uint32_t n_ct, uint32_t level, lmn_store **out) { return store_new(c, nullptr, n_ct, level, c ? c->n : 0, out); }
int lmn_store_wrap(lmn_ctx *c, void *p, uint32_t n_ct, uint32_t level, lmn_store **out) {
    if (!p) return fail(LMN_ERR_INVALID, "null device pointer");
    return store_new(c, p, n_ct, level, c ? c->n : 0, out);
}
int lmn_store_create_small(lmn_ctx *c, uint32_t n_ct, lmn_store **out) {
    if (!c || !c->logn_small) return fail(LMN_ERR_NOKEY, "ring-switch key not loaded");
    return store_new(c, nullptr, n_ct, 0, 1u << c->logn_small, out);
}
void lmn_store_destroy(lmn_store *s) {
    if (!s) return;
    if (s->owned && s->d) { cudaSetDevice(s->ctx->device); cudaStreamSynchronize(s->ctx->stream); cudaFree(s->d); }
    delete s;
}
uint32_t lmn_store_count(const lmn_store *s) { return s ? s->count : 0; }
uint32_t lmn_store_level(const lmn_store *s) { return s ? s->level : 0; }
void *lmn_store_device_ptr(const lmn_store *s) { return s ? s->d : nullptr; }
size_t lmn_store_bytes(const lmn_store *s) { return s ? s->words() * 8 * s->count : 0; }

int lmn_store_upload(lmn_store *s, uint32_t first, uint32_t count, const uint64_t *host) {
    if (!s || !host || (uint64_t)first + count > s->count) return fail(LMN_ERR_INVALID, "upload range out of bounds");
    lmn_ctx *c = s->ctx; std::lock_guard<std::mutex> g(c->mu);
    CK(cudaMemcpyAsync(s->d + first * s->words(), host, count * s->words() * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
int lmn_store_download(const lmn_store *s, uint32_t first, uint32_t count, uint64_t *host) {
    if (!s || !host || (uint64_t)first + count > s->count) return fail(LMN_ERR_INVALID, "download range out of bounds");
    lmn_ctx *c = s->ctx; std::lock_guard<std::mutex> g(c->mu);
    CK(cudaMemcpyAsync(host, s->d + first * s->words(), count * s->words() * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
int lmn_store_upload_limb(lmn_store *s, uint32_t ct, uint32_t poly, uint32_t limb, const uint64_t *host) {
    if (!s || !host || ct >= s->count || poly > 1 || limb > s->level) return fail(LMN_ERR_INVALID, "limb index out of bounds");
    lmn_ctx *c = s->ctx; std::lock_guard<std::mutex> g(c->mu);
    u64 *dst = s->d + ct * s->words() + ((size_t)poly * (s->level + 1) + limb) * s->n;
    CK(cudaMemcpyAsync(dst, host, (size_t)s->n * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
int lmn_store_download_limb(const lmn_store *s, uint32_t ct, uint32_t poly, uint32_t limb, uint64_t *host) {
    if (!s || !host || ct >= s->count || poly > 1 || limb > s->level) return fail(LMN_ERR_INVALID, "limb index out of bounds");
    lmn_ctx *c = s->ctx; std::lock_guard<std::mutex> g(c->mu);
    const u64 *src = s->d + ct * s->words() + ((size_t)poly * (s->level + 1) + limb) * s->n;
    CK(cudaMemcpyAsync(host, src, (size_t)s->n * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
int lmn_host_alloc(size_t bytes, void **out) {
    if (!out) return fail(LMN_ERR_INVALID, "null argument");
    CK(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
    return 0;
}
void lmn_host_free(void *p) { if (p) cudaFreeHost(p); }

// ---------------- plan ----------------
static u64 scalar_mod_q(const lmn_ctx *c, u64 w, u64 q) {
    // bgv.Evaluator.Mul(ct, uint64): scalar mod t, centred when > t/2, Euclidean residue mod q
    u64 t = c->t; w %= t;
    if (c->scalar_centered && w > (t >> 1)) { u64 neg = (t - w) % q; return neg ? q - neg : 0; }
    return w % q;
}
int lmn_plan_create(lmn_ctx *c, uint32_t cols, uint32_t rho_inv, const uint64_t *roots, uint32_t field_n, lmn_plan **out) {
    if (!roots || !out || cols == 0 || rho_inv == 0) return fail(LMN_ERR_INVALID, "bad plan arguments");
    if (field_n == 0 || (field_n & (field_n - 1))) return fail(LMN_ERR_INVALID, "field_n (PrimeField.N()) must be a power of two");
    uint64_t n64 = (uint64_t)cols * rho_inv;
    if (n64 > (1u << 16)) return fail(LMN_ERR_UNSUPPORTED, "codeword length above 65536");
    uint32_t n = (uint32_t)n64;
    lmn_plan *p = new lmn_plan();
    p->ctx = c; p->cols = cols; p->rho_inv = rho_inv;
    if (!c) { delete p; return fail(LMN_ERR_INVALID, "null context"); }
    uint64_t t = c->t;
    std::string err = compile_encode_plan(n, field_n, roots, t, ENC_MAX_TILE, p->pl);
    if (!err.empty()) { delete p; return fail(LMN_ERR_INVALID, err); }
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    const EncodePlan &pl = p->pl;
    size_t nscal = pl.scalars.size();
    std::vector<ulonglong2> scal((size_t)c->k * nscal);
    for (int i = 0; i < c->k; i++)
        for (size_t s = 1; s < nscal; s++) {
            u64 w = scalar_mod_q(c, pl.scalars[s], c->q[i]);
            scal[(size_t)i * nscal + s] = make_ulonglong2(w, h_shoup(w, c->q[i]));
        }
    cudaError_t e = cudaMalloc(&p->d_scale_after, pl.scale_after.size() * sizeof(uint16_t));
    if (e == cudaSuccess) e = cudaMalloc(&p->d_out_pos, pl.out_pos.size() * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&p->d_scal, scal.size() * sizeof(ulonglong2));
    if (e == cudaSuccess) e = cudaMemcpy(p->d_scale_after, pl.scale_after.data(), pl.scale_after.size() * sizeof(uint16_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_out_pos, pl.out_pos.data(), pl.out_pos.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->d_scal, scal.data(), scal.size() * sizeof(ulonglong2), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { int rc = fail(LMN_ERR_CUDA, std::string("plan upload: ") + cudaGetErrorString(e)); lmn_plan_destroy(p); return rc; }
    *out = p;
    return 0;
}
void lmn_plan_destroy(lmn_plan *p) {
    if (!p) return;
    if (p->ctx) cudaSetDevice(p->ctx->device);
    cudaFree(p->d_scale_after); cudaFree(p->d_out_pos); cudaFree(p->d_scal);
    delete p;
}
int lmn_plan_counts(const lmn_plan *p, uint64_t *a, uint64_t *s, uint64_t *m) {
    if (!p) return fail(LMN_ERR_INVALID, "null plan");
    if (a) *a = p->pl.adds; if (s) *s = p->pl.subs; if (m) *m = p->pl.muls;
    return 0;
}
uint32_t lmn_plan_size(const lmn_plan *p) { return p ? p->pl.n : 0; }

// Host-only entry points for CPU-side checks of the plan compiler (no device needed).  They evaluate the compiled
// plan on a plaintext vector over Z_t, i.e. they must reproduce core.NTT (core/ntt.go) -- which is how the
// reference itself validates fhe.NTT (fhe/code_test.go:100-116).  Not part of the ciphertext data path.
int lmn_plan_host_check(uint32_t n, const uint64_t *roots, uint64_t t, uint64_t *values, uint64_t *counts3) {
    if (!roots || !values) return fail(LMN_ERR_INVALID, "null argument");
    EncodePlan pl;
    std::string err = compile_encode_plan(n, n, roots, t, ENC_MAX_TILE, pl);
    if (!err.empty()) return fail(LMN_ERR_INVALID, err);
    std::vector<u64> v(values, values + n);
    for (uint32_t s = 0; s < pl.logn; s++) {
        const uint32_t bit = 1u << pl.stage_bit[s];
        for (uint32_t lo = 0; lo < n; lo++) {
            if (lo & bit) continue;
            u64 a = v[lo], b = v[lo | bit];
            v[lo] = h_addmod(a, b, t); v[lo | bit] = h_submod(a, b, t);
        }
        for (uint32_t x = 0; x < n; x++) {
            uint16_t id = pl.scale_after[(size_t)s * n + x];
            if (id) v[x] = h_mulmod(v[x], pl.scalars[id] % t, t);
        }
    }
    for (uint32_t x = 0; x < n; x++) values[pl.out_pos[x]] = v[x];
    if (counts3) { counts3[0] = pl.adds; counts3[1] = pl.subs; counts3[2] = pl.muls; }
    return 0;
}

// ---------------- Encode ----------------
static int run_plan(lmn_ctx *c, const lmn_plan *p, const u64 *in, uint32_t in_count, const u64 *zero, u64 *out, int lane_logn = -1,
                    uint64_t lane_off = 0, uint64_t lane_words = 0) {
    const EncodePlan &pl = p->pl;
    if (lane_logn < 0) lane_logn = c->logn;
    EncPassArgs A; memset(&A, 0, sizeof(A));
    A.mods = c->d_mods; A.scale_after = p->d_scale_after; A.out_pos = p->d_out_pos; A.scal = p->d_scal;
    A.nscal = (uint32_t)pl.scalars.size(); A.n = pl.n; A.k = c->k; A.logn = lane_logn; A.ct_words = 2 * (uint64_t)c->k << lane_logn;
    A.out = out; A.lane_off = lane_off; A.lane_words = lane_words;
    Launch L = c->L();
    if (pl.split == 0) {
        A.in = in; A.zero = zero; A.in_count = in_count; A.in_by_pos = 0;
        A.log_tile = (int)pl.logn; A.s0 = 0; A.s1 = (int)pl.logn; A.tile_mul = 0; A.stride = 1;
        A.lbits = 0; for (uint32_t s = 0; s < pl.logn; s++) A.lbits |= (uint64_t)pl.stage_bit[s] << (4 * s);
        launch_encode_pass(L, A);
    } else {
        const int lg2 = h_log2(pl.n2);
        // pass 1: size-n1 column transforms + twiddles (fhe/ntt.go:254-274); tile = fixed column c, slots c + r*n2
        A.in = in; A.zero = zero; A.in_count = in_count; A.in_by_pos = 0;
        A.log_tile = (int)pl.split; A.s0 = 0; A.s1 = (int)pl.split; A.tile_mul = 1; A.stride = pl.n2;
        A.lbits = 0; for (uint32_t s = 0; s < pl.split; s++) A.lbits |= (uint64_t)(pl.stage_bit[s] - lg2) << (4 * s);
        launch_encode_pass(L, A);
        // pass 2: size-n2 row transforms (fhe/ntt.go:276), in place on the positions pass 1 wrote
        A.in = out; A.zero = nullptr; A.in_count = pl.n; A.in_by_pos = 1;
        A.log_tile = lg2; A.s0 = (int)pl.split; A.s1 = (int)pl.logn; A.tile_mul = pl.n2; A.stride = 1;
        A.lbits = 0; for (uint32_t s = pl.split; s < pl.logn; s++) A.lbits |= (uint64_t)pl.stage_bit[s] << (4 * (s - pl.split));
        launch_encode_pass(L, A);
    }
    CK(cudaGetLastError());
    return 0;
}
int lmn_encode(lmn_ctx *c, const lmn_plan *p, const lmn_store *matrix, const lmn_store *zero, uint32_t zero_idx, lmn_store *enc) {
    if (!c || !p || !matrix || !zero || !enc) return fail(LMN_ERR_INVALID, "null argument");
    if (p->ctx != c || matrix->ctx != c || zero->ctx != c || enc->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    const uint32_t L = (uint32_t)c->k - 1;
    if (matrix->level != L || zero->level != L || enc->level != L) return fail(LMN_ERR_INVALID, "Encode operates on max-level ciphertexts");
    if (matrix->count != p->cols || enc->count != p->pl.n || zero_idx >= zero->count) return fail(LMN_ERR_INVALID, "store sizes do not match the plan");
    if (enc->d == matrix->d) return fail(LMN_ERR_INVALID, "encoded store must be distinct from the matrix");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    return run_plan(c, p, matrix->d, matrix->count, zero->d + zero_idx * zero->words(), enc->d);
}
static cudaEvent_t copy_event(lmn_ctx *c, size_t i) {
    while (c->copy_events.size() <= i) { cudaEvent_t e; cudaEventCreateWithFlags(&e, cudaEventDisableTiming); c->copy_events.push_back(e); }
    return c->copy_events[i];
}
// Encode with the matrix still on the host: the upload runs limb by limb (one strided copy of that limb of every column) on the
// context's copy stream while the kernels encode the limbs that have landed (Encode is lane-wise, fhe/ntt.go touches one
// (poly, limb, coefficient) lane at a time), so the span costs max(upload, kernels) instead of their sum.
int lmn_encode_host(lmn_ctx *c, const lmn_plan *p, const uint64_t *host_matrix, lmn_store *matrix, const lmn_store *zero, uint32_t zero_idx, lmn_store *enc) {
    if (!c || !p || !host_matrix || !matrix || !zero || !enc) return fail(LMN_ERR_INVALID, "null argument");
    if (p->ctx != c || matrix->ctx != c || zero->ctx != c || enc->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    const uint32_t Lm = (uint32_t)c->k - 1;
    if (matrix->level != Lm || zero->level != Lm || enc->level != Lm) return fail(LMN_ERR_INVALID, "Encode operates on max-level ciphertexts");
    if (matrix->count != p->cols || enc->count != p->pl.n || zero_idx >= zero->count) return fail(LMN_ERR_INVALID, "store sizes do not match the plan");
    if (enc->d == matrix->d) return fail(LMN_ERR_INVALID, "encoded store must be distinct from the matrix");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    const size_t N = c->n, ctb = matrix->words() * 8, limbs = 2 * (size_t)c->k;
    cudaEvent_t start = copy_event(c, limbs);
    CK(cudaEventRecord(start, c->stream));
    CK(cudaStreamWaitEvent(c->copy_stream, start, 0));      // earlier work on the context's stream may still read the matrix store
    for (size_t l = 0; l < limbs; l++) {
        CK(cudaMemcpy2DAsync(matrix->d + l * N, ctb, host_matrix + l * N, ctb, N * 8, matrix->count, cudaMemcpyHostToDevice, c->copy_stream));
        cudaEvent_t ev = copy_event(c, l);
        CK(cudaEventRecord(ev, c->copy_stream));
        CK(cudaStreamWaitEvent(c->stream, ev, 0));
        if (int rc = run_plan(c, p, matrix->d, matrix->count, zero->d + zero_idx * zero->words(), enc->d, -1, l * N, N)) return rc;
    }
    return 0;
}
int lmn_ntt(lmn_ctx *c, const lmn_plan *p, const lmn_store *in, lmn_store *out) {
    if (!c || !p || !in || !out) return fail(LMN_ERR_INVALID, "null argument");
    const uint32_t L = (uint32_t)c->k - 1;
    if (in->level != L || out->level != L || in->count != p->pl.n || out->count != p->pl.n) return fail(LMN_ERR_INVALID, "NTT needs plan-size stores of max-level ciphertexts");
    if (in->d == out->d) return fail(LMN_ERR_INVALID, "in and out must be distinct stores");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    return run_plan(c, p, in->d, in->count, nullptr, out->d);
}

int lmn_encode_lanes(lmn_ctx *c, const lmn_plan *p, const void *matrix_dev, const void *zero_dev, void *encoded_dev, uint32_t log_coeffs) {
    if (!c || !p || !matrix_dev || !zero_dev || !encoded_dev) return fail(LMN_ERR_INVALID, "null argument");
    if (p->ctx != c) return fail(LMN_ERR_INVALID, "plan belongs to a different context");
    if (log_coeffs < 5 || (int)log_coeffs > c->logn) return fail(LMN_ERR_INVALID, "a coefficient slice holds between 32 and N coefficients per limb");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    return run_plan(c, p, (const u64 *)matrix_dev, p->cols, (const u64 *)zero_dev, (u64 *)encoded_dev, (int)log_coeffs);
}

// ---------------- profiling ----------------
int lmn_profile_enable(lmn_ctx *c, uint32_t mask) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    std::lock_guard<std::mutex> g(c->mu);
    c->prof.mask = mask;
    return 0;
}
int lmn_profile_read(lmn_ctx *c, int kernel_id, double *ms, uint64_t *launches, uint64_t *jobs, int reset) {
    if (!c || kernel_id < 0 || kernel_id >= KID_COUNT) return fail(LMN_ERR_INVALID, "bad kernel id");
    std::lock_guard<std::mutex> g(c->mu);
    CK(cudaStreamSynchronize(c->stream));
    double total = 0;
    for (auto &sp : c->prof.spans[kernel_id]) {
        float t = 0; CK(cudaEventElapsedTime(&t, sp.first, sp.second));
        total += t;
    }
    if (ms) *ms = total;
    if (launches) *launches = c->prof.launches[kernel_id];
    if (jobs) *jobs = c->prof.jobs[kernel_id];
    if (reset) {
        for (auto &sp : c->prof.spans[kernel_id]) { c->prof.pool.push_back(sp.first); c->prof.pool.push_back(sp.second); }
        c->prof.spans[kernel_id].clear();
        c->prof.launches[kernel_id] = 0;
        c->prof.jobs[kernel_id] = 0;
    }
    return 0;
}
int lmn_kernel_count(void) { return KID_COUNT; }
const char *lmn_kernel_name(int id) { return kernel_name(id); }
int lmn_bench_int_peak(lmn_ctx *c, double *butterflies_per_s) {
    if (!c || !butterflies_per_s) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    *butterflies_per_s = measure_peak_butterflies(c->stream, c->d_mods, sms);
    CK(cudaGetLastError());
    return 0;
}

int lmn_bench_product_butterfly(lmn_ctx *c, double *butterflies_per_s) {
    if (!c || !butterflies_per_s) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    *butterflies_per_s = measure_peak_butterflies(c->stream, c->d_mods, sms, 1);
    CK(cudaGetLastError());
    return 0;
}

// ---------------- keys ----------------
int lmn_load_galois_key(lmn_ctx *c, uint64_t gal_el, const uint64_t *evk) {
    if (!c || !evk) return fail(LMN_ERR_INVALID, "null argument");
    if (!(gal_el & 1) || gal_el >= 2 * (u64)c->n) return fail(LMN_ERR_INVALID, "Galois element must be odd and below 2N");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (gal_el == 1) return 0;  // identity: Lattigo's key list may contain it; never used
    const int beta = (c->k + c->np - 1) / c->np;
    size_t bytes = (size_t)beta * 2 * (c->k + c->np) * c->n * 8;
    auto it = c->galois.find(gal_el);
    DevScratch fresh, stage;     // a key that is not yet in the map is owned here until the upload has succeeded
    if (it == c->galois.end()) CK(fresh.alloc(bytes));
    u64 *d = it != c->galois.end() ? it->second : fresh.as<u64>();
    CK(stage.alloc(bytes));
    CK(cudaMemcpyAsync(stage.p, evk, bytes, cudaMemcpyHostToDevice, c->stream));
    launch_scale_rows(c->L(), c->T, stage.as<u64>(), d, c->d_key_fm, beta);   // fold 2^s into the rows the base-converted digits multiply
    CK(cudaStreamSynchronize(c->stream));
    fresh.release();
    c->galois[gal_el] = d;
    return 0;
}
int lmn_load_ring_switch_key(lmn_ctx *c, uint32_t logns, const uint64_t *evk, uint32_t rns_digits, uint32_t base2_digits, uint64_t psi_small) {
    if (!c || !evk) return fail(LMN_ERR_INVALID, "null argument");
    if (rns_digits != (uint32_t)((c->k + c->np - 1) / c->np)) return fail(LMN_ERR_INVALID, "ring-switch key must have ceil(k / np) RNS digits");
    if (base2_digits != 1) return fail(LMN_ERR_UNSUPPORTED, "ring-switch keys with base-2 sub-digits (BaseTwoDecomposition) are not supported: evk.Value[d] must have one entry");
    if ((int)logns >= c->logn) return fail(LMN_ERR_INVALID, "small ring must be smaller than the ciphertext ring");
    bool ok = (logns == 10 && c->logn >= 12) || (logns == 6 && (c->logn == 8 || c->logn == 10)) || (logns == 8 && (c->logn == 10 || c->logn == 12));
    if (!ok) return fail(LMN_ERR_UNSUPPORTED, "ring switch is instantiated for N' = 1024 (N >= 4096) and the small test rings");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    const int beta = (c->k + c->np - 1) / c->np;
    size_t bytes = (size_t)beta * 2 * (c->k + c->np) * c->n * 8;
    if (!c->d_rs_evk) CK(cudaMalloc(&c->d_rs_evk, bytes));
    DevScratch stage;
    CK(stage.alloc(bytes));
    CK(cudaMemcpy(stage.p, evk, bytes, cudaMemcpyHostToDevice));
    launch_scale_rows(c->L(), c->T, stage.as<u64>(), c->d_rs_evk, c->d_key_fm, beta);
    CK(cudaStreamSynchronize(c->stream));
    u64 q0 = c->q[0];
    if (!psi_small) {
        // psi_small = psi^(N/N'): read psi back from the forward table (entry bitrev(1) = N/2 holds psi)
        ulonglong2 w; CK(cudaMemcpy(&w, c->d_twf + c->n / 2, sizeof(w), cudaMemcpyDeviceToHost));
        psi_small = h_powmod(w.x, (u64)c->n >> logns, q0);
    }
    if (h_powmod(psi_small, 1ull << logns, q0) != q0 - 1) return fail(LMN_ERR_INVALID, "psi_small is not a primitive 2N'-th root");
    ModDev m; build_mod(m, q0, 1 << logns);
    std::vector<ulonglong2> f, iv; build_twiddles(q0, psi_small, (int)logns, f, iv);
    if (c->d_small_mod) { cudaFree(c->d_small_mod); cudaFree(c->d_small_tw); }
    CK(cudaMalloc(&c->d_small_mod, sizeof(ModDev)));
    CK(cudaMalloc(&c->d_small_tw, sizeof(ulonglong2) << logns));
    CK(cudaMemcpy(c->d_small_mod, &m, sizeof(m), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->d_small_tw, f.data(), sizeof(ulonglong2) << logns, cudaMemcpyHostToDevice));
    c->logn_small = (int)logns;
    return 0;
}

// ---------------- InnerSum on the job workspace ----------------
static int innersum_jobs(lmn_ctx *c, int level, KsJobs &J, uint32_t n) {
    const uint32_t N = (uint32_t)c->n;
    if (n == 0 || (n & (n - 1)) || n > N) return fail(LMN_ERR_UNSUPPORTED, "InnerSum is implemented for power-of-two n <= N");
    const KsConsts *Kh, *Kd;
    if (int rc = get_ks(c, level, &Kh, &Kd)) return rc;
    const bool full = n == N;
    const uint32_t m = full ? N / 2 : n;
    std::vector<u64> gals;
    if (full && !c->rowswap_last) gals.push_back(2 * (u64)N - 1);
    for (uint32_t s = 1; s < m; s <<= 1) gals.push_back(galois_element(c, s));
    if (full && c->rowswap_last) gals.push_back(2 * (u64)N - 1);
    for (u64 g : gals) if (!c->galois.count(g)) return fail(LMN_ERR_NOKEY, "missing Galois key for element " + std::to_string(g));
    Launch L = c->L();
    const ptrdiff_t off = J.col - c->ws_col;
    for (u64 g : gals) launch_rotate_accumulate(L, c->T, Kd, *Kh, J, c->galois[g], (uint32_t)g);
    if (J.col != c->ws_col + off) std::swap(c->ws_col, c->ws_col2);   // the view ended in the second buffer: make it the current one
    CK(cudaGetLastError());
    return 0;
}
static int get_rs_chain(lmn_ctx *c, int from, int to, const RescaleChain **host, const RescaleChain **dev) {
    const int key = from * 64 + to;
    auto it = c->rs_chain.find(key);
    if (it == c->rs_chain.end()) {
        RescaleChain R; memset(&R, 0, sizeof(R));
        R.from = from; R.to = to;
        for (int l = from; l > to; l--) {
            const u64 ql = c->q[l];
            R.half[l] = (ql - 1) >> 1;
            for (int i = 0; i < l; i++) {
                const u64 qi = c->q[i];
                R.half_mod[l][i] = R.half[l] % qi;
                const u64 inv = h_invmod(ql % qi, qi);
                R.qlinv_w[l][i] = inv; R.qlinv_ws[l][i] = h_shoup(inv, qi);
                R.cadd[l][i] = R.half_mod[l][i] + ((ql + qi - 1) / qi) * qi;      // < 2^61: moduli are below 2^59 and within a factor 4 of each other
            }
        }
        RescaleChain *d = nullptr;
        CK(cudaMalloc(&d, sizeof(R)));
        CK(cudaMemcpy(d, &R, sizeof(R), cudaMemcpyHostToDevice));
        it = c->rs_chain.emplace(key, std::make_pair(R, d)).first;
    }
    *host = &it->second.first; *dev = it->second.second;
    return 0;
}
// src != nullptr: the jobs' input ciphertexts are the consecutive entries of a store starting at src (not yet in J.col)
static int rescale_jobs_to(lmn_ctx *c, int from_level, int to_level, const KsJobs &J, const u64 *src = nullptr) {
    Launch L = c->L();
    static const bool stepwise = getenv("LMN_RESCALE_STEPWISE") != nullptr;   // the one-limb-at-a-time evaluation, for A/B tests
    // log N >= 12, two or more steps: the whole chain in the coefficient domain (3 launches instead of 3 per step)
    if (!stepwise && blk_blocks(c->logn) > 0 && from_level - to_level >= 2 &&
        (size_t)(from_level - to_level) * blk_blocks(c->logn) * 128 * 8 <= 96 * 1024 && 2 * (from_level + 1) <= MAX_LIMBS) {
        const RescaleChain *Rh, *Rd;
        if (int rc = get_rs_chain(c, from_level, to_level, &Rh, &Rd)) return rc;
        launch_rescale_chain(L, c->T, Rd, *Rh, J, src, J.col_words);
        CK(cudaGetLastError());
        return 0;
    }
    if (src) CK(cudaMemcpyAsync(J.col, src, (size_t)J.jobs * J.col_words * 8, cudaMemcpyDeviceToDevice, c->stream));   // the stepwise path works in place
    for (int l = from_level; l > to_level; l--) {
        const RescaleConsts *Rd;
        if (int rc = get_rs(c, l, &Rd)) return rc;
        launch_rescale_step(L, c->T, Rd, l, J);
    }
    CK(cudaGetLastError());
    return 0;
}

// ---------------- Inner product eval ----------------
static int inner_product2_impl(lmn_ctx *c, const lmn_store *matrix, const uint64_t *pt_r, const uint64_t *pt_b, uint32_t rows,
                               lmn_store *out_r, lmn_store *out_b, uint64_t *host_r, uint64_t *host_b) {
    if (!c || !matrix || !pt_r || !out_r || (pt_b && !out_b)) return fail(LMN_ERR_INVALID, "null argument");
    const int Lmax = c->k - 1, lout = std::min(1, Lmax);
    if ((int)matrix->level != Lmax) return fail(LMN_ERR_INVALID, "matrix must hold max-level ciphertexts");
    if ((int)out_r->level != lout || out_r->count != matrix->count || (pt_b && ((int)out_b->level != lout || out_b->count != matrix->count)))
        return fail(LMN_ERR_INVALID, "output stores must hold one level-1 ciphertext per column");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    const size_t N = c->n, k = c->k;
    const int nvec = pt_b ? 2 : 1;
    uint32_t bc = std::max(1u, c->batch / nvec);
    // results that also go home are copied batch by batch behind the next batch's kernels: keep at least four batches (of at least 64
    // columns) so that a rank whose whole block fits one launch still hides three quarters of its device-to-host copy
    if (host_r) bc = std::min(bc, std::max(64u, (matrix->count + 3) / 4));
    if (int rc = ensure_workspace(c, std::min(bc, matrix->count) * nvec)) return rc;
    Launch L = c->L();
    // plaintexts: upload, fold T and the Montgomery factor (tensorStandard's T * pt)
    u64 *d_pt = c->ws_pt, *d_ptm = c->ws_pt + 2 * k * N;
    if (int rc = stage_begin(c, 2 * k * N * 8 + 32)) return rc;
    if (int rc = stage_copy(c, d_pt, pt_r, k * N * 8)) return rc;
    if (pt_b) if (int rc = stage_copy(c, d_pt + k * N, pt_b, k * N * 8)) return rc;
    if (int rc = stage_end(c)) return rc;
    launch_prep_pt(L, c->T, Lmax, d_pt, c->t, d_ptm);
    if (pt_b) launch_prep_pt(L, c->T, Lmax, d_pt + k * N, c->t, d_ptm + k * N);
    const size_t out_words = 2 * (size_t)(lout + 1) * N;
    for (uint32_t c0 = 0; c0 < matrix->count; c0 += bc) {
        const uint32_t nc = std::min(bc, matrix->count - c0);
        KsJobs J = jobs_view(c, 0, nc * nvec);
        launch_mul_pt(L, c->T, Lmax, matrix->d + c0 * matrix->words(), matrix->words(), (int)nc, d_ptm, pt_b ? d_ptm + k * N : nullptr,
                      J.col, J.col + nc * J.col_words, J.col_words);
        if (int rc = innersum_jobs(c, Lmax, J, rows)) return rc;
        if (int rc = rescale_jobs_to(c, Lmax, lout, J)) return rc;
        KsJobs Jr = jobs_view(c, 0, nc);
        launch_pack_out(L, c->T, Jr, lout + 1, out_r->d + c0 * out_words);
        if (pt_b) { KsJobs Jb = jobs_view(c, nc, nc); launch_pack_out(L, c->T, Jb, lout + 1, out_b->d + c0 * out_words); }
        if (host_r) {
            // this batch's results go home on the copy stream while the next batch computes
            cudaEvent_t ev = copy_event(c, c0 / bc);
            CK(cudaEventRecord(ev, c->stream));
            CK(cudaStreamWaitEvent(c->copy_stream, ev, 0));
            CK(cudaMemcpyAsync(host_r + c0 * out_words, out_r->d + c0 * out_words, nc * out_words * 8, cudaMemcpyDeviceToHost, c->copy_stream));
            if (pt_b && host_b) CK(cudaMemcpyAsync(host_b + c0 * out_words, out_b->d + c0 * out_words, nc * out_words * 8, cudaMemcpyDeviceToHost, c->copy_stream));
        }
    }
    if (host_r) {
        cudaEvent_t done = copy_event(c, (matrix->count + bc - 1) / bc);
        CK(cudaEventRecord(done, c->copy_stream));
        CK(cudaStreamWaitEvent(c->stream, done, 0));    // whoever synchronises the context's stream also has the host copies
    }
    CK(cudaGetLastError());
    return 0;
}
int lmn_inner_product2(lmn_ctx *c, const lmn_store *matrix, const uint64_t *pt_r, const uint64_t *pt_b, uint32_t rows,
                       lmn_store *out_r, lmn_store *out_b) {
    return inner_product2_impl(c, matrix, pt_r, pt_b, rows, out_r, out_b, nullptr, nullptr);
}
// the same, with the level-1 results also copied to (pinned) host buffers batch by batch behind the kernels
int lmn_inner_product2_host(lmn_ctx *c, const lmn_store *matrix, const uint64_t *pt_r, const uint64_t *pt_b, uint32_t rows,
                            lmn_store *out_r, lmn_store *out_b, uint64_t *host_r, uint64_t *host_b) {
    if (!host_r || (pt_b && !host_b)) return fail(LMN_ERR_INVALID, "null host buffer");
    return inner_product2_impl(c, matrix, pt_r, pt_b, rows, out_r, out_b, host_r, host_b);
}
int lmn_inner_product(lmn_ctx *c, const lmn_store *matrix, const uint64_t *pt, uint32_t rows, lmn_store *out) {
    return lmn_inner_product2(c, matrix, pt, nullptr, rows, out, nullptr);
}

uint64_t lmn_rescale_scale(const lmn_ctx *c, uint32_t from_level, uint32_t to_level) {
    if (!c || (int)from_level >= c->k || to_level > from_level) return 0;
    u64 s = 1;
    for (uint32_t l = from_level; l > to_level; l--) s = h_mulmod(s, h_invmod(c->q[l] % c->t, c->t), c->t);
    return s;
}

// ---------------- Query columns / leaf rescale ----------------
// out[j] = enc[idx[j]] rescaled to level 1, for host indices idx (nullptr: idx[j] = j).  Caller holds the context lock.
static int rescale_select(lmn_ctx *c, const lmn_store *enc, const uint32_t *idx, uint32_t n_idx, lmn_store *out) {
    const int Lmax = c->k - 1, lout = std::min(1, Lmax);
    if ((int)enc->level != Lmax || (int)out->level != lout || out->count < n_idx) return fail(LMN_ERR_INVALID, "rescale needs max-level input and a level-1 output store with one entry per index");
    const bool all = idx == nullptr;   // every entry in order (Commit's leaves): the first transform reads the store itself, no gather
    if (all && n_idx > enc->count) return fail(LMN_ERR_INVALID, "column index out of range");
    for (uint32_t i = 0; !all && i < n_idx; i++) if (idx[i] >= enc->count) return fail(LMN_ERR_INVALID, "column index out of range");
    cudaSetDevice(c->device);
    if (!all) {
        if (c->ws_idx_cap < n_idx) { cudaFree(c->ws_idx); c->ws_idx = nullptr; c->ws_idx_cap = 0; CK(cudaMalloc(&c->ws_idx, n_idx * sizeof(uint32_t))); c->ws_idx_cap = n_idx; }
        if (int rc = stage_begin(c, n_idx * sizeof(uint32_t) + 16)) return rc;
        if (int rc = stage_copy(c, c->ws_idx, idx, n_idx * sizeof(uint32_t))) return rc;
        if (int rc = stage_end(c)) return rc;
    }
    const uint32_t bc = c->batch;
    if (int rc = ensure_workspace(c, std::min(bc, n_idx))) return rc;
    Launch L = c->L();
    const size_t out_words = 2 * (size_t)(lout + 1) * c->n;
    for (uint32_t j0 = 0; j0 < n_idx; j0 += bc) {
        const uint32_t nj = std::min(bc, n_idx - j0);
        KsJobs J = jobs_view(c, 0, nj);
        if (!all) launch_gather_in(L, c->T, enc->d, enc->words(), c->ws_idx + j0, J);
        if (int rc = rescale_jobs_to(c, Lmax, lout, J, all ? enc->d + (size_t)j0 * enc->words() : nullptr)) return rc;
        launch_pack_out(L, c->T, J, lout + 1, out->d + j0 * out_words);
    }
    CK(cudaGetLastError());
    return 0;
}
}  // extern "C"
// out[j] (level 1, packed) = src[j] (max level, consecutive raw device ciphertexts) rescaled, j < count.  Caller holds the lock.
int rescale_consecutive(lmn_ctx *c, const u64 *src, uint32_t count, u64 *out) {
    const int Lmax = c->k - 1, lout = std::min(1, Lmax);
    const uint32_t bc = c->batch;
    if (int rc = ensure_workspace(c, std::min(bc, count))) return rc;
    Launch L = c->L();
    const size_t out_words = 2 * (size_t)(lout + 1) * c->n, in_words = 2 * (size_t)c->k * c->n;
    for (uint32_t j0 = 0; j0 < count; j0 += bc) {
        const uint32_t nj = std::min(bc, count - j0);
        KsJobs J = jobs_view(c, 0, nj);
        if (int rc = rescale_jobs_to(c, Lmax, lout, J, src + (size_t)j0 * in_words)) return rc;
        launch_pack_out(L, c->T, J, lout + 1, out + j0 * out_words);
    }
    CK(cudaGetLastError());
    return 0;
}
extern "C" {
int lmn_query_columns(lmn_ctx *c, const lmn_store *enc, const uint32_t *idx, uint32_t n_idx, lmn_store *out) {
    if (!c || !enc || !idx || !out || n_idx == 0) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);
    return rescale_select(c, enc, idx, n_idx, out);
}

// ---------------- ring switch ----------------
int lmn_ring_switch(lmn_ctx *c, const lmn_store *in, lmn_store *out) {
    if (!c || !in || !out) return fail(LMN_ERR_INVALID, "null argument");
    if (!c->logn_small) return fail(LMN_ERR_NOKEY, "ring-switch key not loaded");
    if (in->level != 1 || out->level != 0 || out->n != (1u << c->logn_small) || out->count != in->count)
        return fail(LMN_ERR_INVALID, "ring switch maps level-1 ciphertexts to level-0 ciphertexts of the small ring");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    const KsConsts *Kh, *Kd;
    if (int rc = get_ks(c, 0, &Kh, &Kd)) return rc;
    const uint32_t bc = c->batch;
    if (int rc = ensure_workspace(c, bc)) return rc;
    Launch L = c->L();
    const size_t N = c->n, k = c->k;
    for (uint32_t j0 = 0; j0 < in->count; j0 += bc) {
        const uint32_t nj = std::min(bc, in->count - j0);
        KsJobs J = jobs_view(c, 0, nj);
        // limb 0 of both polys of each level-1 ct -> col[c][0]   (the key switch runs at level min(1, 0) = 0)
        CK(cudaMemcpy2DAsync(J.col, k * N * 8, in->d + j0 * in->words(), 2 * N * 8, N * 8, (size_t)nj * 2, cudaMemcpyDeviceToDevice, c->stream));
        launch_keyswitch_level0(L, c->T, Kd, *Kh, J, c->d_rs_evk);
        launch_ring_decimate(L, c->T, J, c->logn_small, c->d_small_mod, c->d_small_tw, out->d + j0 * out->words());
    }
    CK(cudaGetLastError());
    return 0;
}

// ---------------- operator level ----------------
static int check_pair(lmn_ctx *c, const lmn_store *a, uint32_t ai, const lmn_store *o, uint32_t oi) {
    if (!c || !a || !o) return fail(LMN_ERR_INVALID, "null argument");
    if (ai >= a->count || oi >= o->count) return fail(LMN_ERR_INVALID, "ciphertext index out of range");
    if (a->n != (uint32_t)c->n || o->n != (uint32_t)c->n) return fail(LMN_ERR_INVALID, "operator needs ciphertexts of the context ring");
    return 0;
}
static int addsub(lmn_ctx *c, const lmn_store *a, uint32_t ai, const lmn_store *b, uint32_t bi, lmn_store *o, uint32_t oi, int sub) {
    if (int rc = check_pair(c, a, ai, o, oi)) return rc;
    if (!b || bi >= b->count || a->level != b->level || a->level != o->level) return fail(LMN_ERR_INVALID, "operands must share one level");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    launch_addsub(c->L(), c->T, (int)a->level, a->d + ai * a->words(), b->d + bi * b->words(), o->d + oi * o->words(), sub);
    CK(cudaGetLastError());
    return 0;
}
int lmn_ct_add(lmn_ctx *c, const lmn_store *a, uint32_t ai, const lmn_store *b, uint32_t bi, lmn_store *o, uint32_t oi) { return addsub(c, a, ai, b, bi, o, oi, 0); }
int lmn_ct_sub(lmn_ctx *c, const lmn_store *a, uint32_t ai, const lmn_store *b, uint32_t bi, lmn_store *o, uint32_t oi) { return addsub(c, a, ai, b, bi, o, oi, 1); }
int lmn_ct_mul_scalar(lmn_ctx *c, const lmn_store *a, uint32_t ai, uint64_t w, lmn_store *o, uint32_t oi) {
    if (int rc = check_pair(c, a, ai, o, oi)) return rc;
    if (a->level != o->level) return fail(LMN_ERR_INVALID, "operands must share one level");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    ulonglong2 ws[MAX_LIMBS];
    for (int i = 0; i <= (int)a->level; i++) { u64 wi = scalar_mod_q(c, w, c->q[i]); ws[i] = make_ulonglong2(wi, h_shoup(wi, c->q[i])); }
    CK(cudaMemcpyAsync(c->ws_scalar, ws, sizeof(ulonglong2) * (a->level + 1), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // ws is a stack buffer
    launch_mul_scalar(c->L(), c->T, (int)a->level, a->d + ai * a->words(), c->ws_scalar, o->d + oi * o->words());
    CK(cudaGetLastError());
    return 0;
}
int lmn_ct_mul_pt(lmn_ctx *c, const lmn_store *a, uint32_t ai, const uint64_t *pt, lmn_store *o, uint32_t oi) {
    if (int rc = check_pair(c, a, ai, o, oi)) return rc;
    if (!pt || a->level != o->level) return fail(LMN_ERR_INVALID, "operands must share one level");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    const size_t N = c->n, nl = a->level + 1;
    u64 *d_pt = c->ws_pt, *d_ptm = c->ws_pt + 2 * (size_t)c->k * N;
    if (int rc = stage_begin(c, nl * N * 8 + 16)) return rc;
    if (int rc = stage_copy(c, d_pt, pt, nl * N * 8)) return rc;
    if (int rc = stage_end(c)) return rc;
    Launch L = c->L();
    launch_prep_pt(L, c->T, (int)a->level, d_pt, c->t, d_ptm);
    launch_mul_pt(L, c->T, (int)a->level, a->d + ai * a->words(), a->words(), 1, d_ptm, nullptr, o->d + oi * o->words(), nullptr, o->words());
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
// copy a store entry at `level` into / out of job 0's col (limb stride k)
static int col_load(lmn_ctx *c, const u64 *src, int level) {
    const size_t N = c->n;
    CK(cudaMemcpy2DAsync(c->ws_col, c->k * N * 8, src, (size_t)(level + 1) * N * 8, (size_t)(level + 1) * N * 8, 2, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}
static int col_store(lmn_ctx *c, u64 *dst, int level) {
    const size_t N = c->n;
    CK(cudaMemcpy2DAsync(dst, (size_t)(level + 1) * N * 8, c->ws_col, c->k * N * 8, (size_t)(level + 1) * N * 8, 2, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}
int lmn_ct_inner_sum(lmn_ctx *c, lmn_store *a, uint32_t ai, uint32_t n) {
    if (int rc = check_pair(c, a, ai, a, ai)) return rc;
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = ensure_workspace(c, 2)) return rc;
    if (int rc = col_load(c, a->d + ai * a->words(), (int)a->level)) return rc;
    KsJobs J = jobs_view(c, 0, 1);
    if (int rc = innersum_jobs(c, (int)a->level, J, n)) return rc;
    return col_store(c, a->d + ai * a->words(), (int)a->level);
}
int lmn_ct_rescale(lmn_ctx *c, const lmn_store *a, uint32_t ai, lmn_store *o, uint32_t oi) {
    if (int rc = check_pair(c, a, ai, o, oi)) return rc;
    if (a->level == 0 || o->level + 1 != a->level) return fail(LMN_ERR_INVALID, "Rescale output must be one level below a non-zero input level");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = ensure_workspace(c, 2)) return rc;
    if (int rc = col_load(c, a->d + ai * a->words(), (int)a->level)) return rc;
    KsJobs J = jobs_view(c, 0, 1);
    if (int rc = rescale_jobs_to(c, (int)a->level, (int)a->level - 1, J)) return rc;
    return col_store(c, o->d + oi * o->words(), (int)o->level);
}
int lmn_ct_copy(lmn_ctx *c, const lmn_store *a, uint32_t ai, lmn_store *o, uint32_t oi) {
    if (int rc = check_pair(c, a, ai, o, oi)) return rc;
    if (a->level != o->level) return fail(LMN_ERR_INVALID, "operands must share one level");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    CK(cudaMemcpyAsync(o->d + oi * o->words(), a->d + ai * a->words(), a->words() * 8, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}
int lmn_limb_ntt(lmn_ctx *c, lmn_store *a, uint32_t ai, uint32_t poly, uint32_t limb, int inverse) {
    if (int rc = check_pair(c, a, ai, a, ai)) return rc;
    if (poly > 1 || limb > a->level) return fail(LMN_ERR_INVALID, "limb index out of bounds");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    launch_limb_ntt(c->L(), c->T, (int)limb, a->d + ai * a->words() + ((size_t)poly * (a->level + 1) + limb) * c->n, inverse);
    CK(cudaGetLastError());
    return 0;
}

// ---------------- Commit eval: leaves and Merkle tree (fhe/ligero.go:105-116, 126-183; core/tree.go) ----------------
struct lmn_merkle {
    uint32_t n_leaves = 0;
    std::vector<uint32_t> level_off;     // node offset of every level (level 0 = leaf hashes)
    std::vector<uint32_t> level_cnt;
    std::vector<uint8_t> nodes;          // 32 bytes per node, host copy of the whole tree
};

int lmn_commit_leaves(lmn_ctx *c, const lmn_store *enc, lmn_store *leaves) {
    if (!c || !enc || !leaves) return fail(LMN_ERR_INVALID, "null argument");
    if (leaves->count != enc->count) return fail(LMN_ERR_INVALID, "one leaf per encoded column");
    std::lock_guard<std::mutex> g(c->mu);
    return rescale_select(c, enc, nullptr, enc->count, leaves);
}

int lmn_store_gather(lmn_ctx *c, const lmn_store *src, const uint32_t *idx, uint32_t n_idx, lmn_store *out) {
    if (!c || !src || !idx || !out || n_idx == 0) return fail(LMN_ERR_INVALID, "null argument");
    if (src->level != out->level || src->n != out->n || out->count < n_idx) return fail(LMN_ERR_INVALID, "gather needs stores of one level and n_idx output entries");
    for (uint32_t i = 0; i < n_idx; i++) if (idx[i] >= src->count) return fail(LMN_ERR_INVALID, "index out of range");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    // copies queued in index order on the context's stream (<= a few hundred level-1 ciphertexts on the query path)
    for (uint32_t i = 0; i < n_idx; i++)
        CK(cudaMemcpyAsync(out->d + i * out->words(), src->d + (size_t)idx[i] * src->words(), src->words() * 8, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}

static int make_leaf_format(lmn_ctx *c, const lmn_store *leaves, const lmn_leaf_format *fmt, LeafFormat &F, uint8_t **d_hdr) {
    if (fmt->poly_header_len > LEAF_HDR_MAX || fmt->limb_header_len > LEAF_HDR_MAX) return fail(LMN_ERR_INVALID, "poly / limb headers are at most 16 bytes");
    if ((fmt->header_len && !fmt->header) || (fmt->poly_header_len && !fmt->poly_header) || (fmt->limb_header_len && !fmt->limb_header))
        return fail(LMN_ERR_INVALID, "null header with non-zero length");
    memset(&F, 0, sizeof(F));
    F.hdr_len = fmt->header_len; F.poly_hdr_len = fmt->poly_header_len; F.limb_hdr_len = fmt->limb_header_len;
    if (F.poly_hdr_len) memcpy(F.poly_hdr, fmt->poly_header, F.poly_hdr_len);
    if (F.limb_hdr_len) memcpy(F.limb_hdr, fmt->limb_header, F.limb_hdr_len);
    F.nl = leaves->level + 1;
    int logn = 0; while ((1u << logn) < leaves->n) logn++;
    F.logn = (uint32_t)logn;
    const uint64_t limb_len = F.limb_hdr_len + 8ull * leaves->n, poly_len = F.poly_hdr_len + F.nl * limb_len, total = F.hdr_len + 2 * poly_len;
    if (total + 72 >= (1ull << 31)) return fail(LMN_ERR_UNSUPPORTED, "leaf stream longer than 2 GiB");
    F.limb_len = (uint32_t)limb_len; F.poly_len = (uint32_t)poly_len; F.total_len = (uint32_t)total;
    F.padded_len = (uint32_t)((total + 9 + 63) / 64 * 64);
    *d_hdr = nullptr;
    if (F.hdr_len) {
        CK(cudaMalloc(d_hdr, F.hdr_len));
        const cudaError_t e = cudaMemcpy(*d_hdr, fmt->header, F.hdr_len, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { cudaFree(*d_hdr); *d_hdr = nullptr; return fail(LMN_ERR_CUDA, cudaGetErrorString(e)); }
    }
    F.hdr = *d_hdr;
    return 0;
}

int lmn_leaf_bytes(lmn_ctx *c, const lmn_store *leaves, uint32_t index, const lmn_leaf_format *fmt, uint8_t *out, size_t cap, size_t *len) {
    if (!c || !leaves || !fmt || !len) return fail(LMN_ERR_INVALID, "null argument");
    if (index >= leaves->count) return fail(LMN_ERR_INVALID, "leaf index out of range");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    LeafFormat F; uint8_t *d_hdr = nullptr;
    if (int rc = make_leaf_format(c, leaves, fmt, F, &d_hdr)) return rc;
    *len = F.total_len;
    int rc = 0;
    if (out) {
        if (cap < F.total_len) rc = fail(LMN_ERR_INVALID, "output buffer too small for the leaf");
        else {
            u64 *ser = nullptr;
            cudaError_t e = cudaMalloc(&ser, F.padded_len);
            if (e == cudaSuccess) {
                launch_leaf_serialize(c->L(), F, leaves->d + (size_t)index * leaves->words(), leaves->words(), 1, ser);
                e = cudaMemcpyAsync(out, ser, F.total_len, cudaMemcpyDeviceToHost, c->stream);
                if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
                cudaFree(ser);
            }
            if (e != cudaSuccess) rc = fail(LMN_ERR_CUDA, cudaGetErrorString(e));
        }
    }
    cudaFree(d_hdr);
    return rc;
}

// ---- wire format: EncryptedProof.WriteTo, fhe/ligero.go:660-704 ----
// ct.WriteTo(cts[first]) || ... || ct.WriteTo(cts[first + count - 1]) into host memory: the byte streams are produced on the GPU in
// chunks of two alternating staging buffers, each chunk copied home on the copy stream while the next one is serialised, so the
// marshalled bytes land in the caller's buffer without an intermediate host copy of the limbs.  Caller holds the lock.
static int store_marshal_locked(lmn_ctx *c, const lmn_store *s, uint32_t first, uint32_t count, const lmn_leaf_format *fmt, uint8_t *out,
                                size_t cap, size_t *len) {
    LeafFormat F; uint8_t *d_hdr = nullptr;
    if (int rc = make_leaf_format(c, s, fmt, F, &d_hdr)) return rc;
    const uint64_t total = (uint64_t)count * F.total_len;
    *len = (size_t)total;
    int rc = 0;
    if (out && total) {
        if (cap < total) rc = fail(LMN_ERR_INVALID, "output buffer too small for the marshalled ciphertexts");
        else {
            const uint64_t CH = std::min<uint64_t>((total + 7) / 8 * 8, 64ull << 20);    // bytes per staging buffer (multiple of 8)
            u64 *stg[2] = {nullptr, nullptr};
            cudaError_t e = cudaMalloc(&stg[0], CH);
            if (e == cudaSuccess && total > CH) e = cudaMalloc(&stg[1], CH);
            const u64 *cts = s->d + (size_t)first * s->words();
            Launch L = c->L();
            uint64_t nchunk = 0;
            for (uint64_t b0 = 0; e == cudaSuccess && b0 < total; b0 += CH, nchunk++) {
                const int bi = (int)(nchunk & 1);
                const uint64_t b1 = std::min(total, b0 + CH);
                cudaEvent_t ser = copy_event(c, bi), copied = copy_event(c, 2 + bi);
                if (nchunk >= 2) e = cudaStreamWaitEvent(c->stream, copied, 0);          // the buffer's previous chunk has gone home
                if (e != cudaSuccess) break;
                launch_ct_stream(L, F, cts, s->words(), b0, b1, stg[bi]);
                e = cudaGetLastError();
                if (e == cudaSuccess) e = cudaEventRecord(ser, c->stream);
                if (e == cudaSuccess) e = cudaStreamWaitEvent(c->copy_stream, ser, 0);
                if (e == cudaSuccess) e = cudaMemcpyAsync(out + b0, stg[bi], b1 - b0, cudaMemcpyDeviceToHost, c->copy_stream);
                if (e == cudaSuccess) e = cudaEventRecord(copied, c->copy_stream);
            }
            cudaError_t e2 = cudaStreamSynchronize(c->copy_stream);
            cudaError_t e3 = cudaStreamSynchronize(c->stream);
            if (e == cudaSuccess) e = e2 != cudaSuccess ? e2 : e3;
            cudaFree(stg[0]); cudaFree(stg[1]);
            if (e != cudaSuccess) rc = fail(LMN_ERR_CUDA, cudaGetErrorString(e));
        }
    }
    cudaFree(d_hdr);
    return rc;
}
int lmn_store_marshal(lmn_ctx *c, const lmn_store *s, uint32_t first, uint32_t count, const lmn_leaf_format *fmt, uint8_t *out, size_t cap, size_t *len) {
    if (!c || !s || !fmt || !len) return fail(LMN_ERR_INVALID, "null argument");
    if (s->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    if ((uint64_t)first + count > s->count) return fail(LMN_ERR_INVALID, "ciphertext range out of bounds");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    return store_marshal_locked(c, s, first, count, fmt, out, cap, len);
}
int lmn_proof_marshal(lmn_ctx *c, const lmn_ligero_metadata *md, const lmn_store *mat_r, const lmn_store *mat_z, const lmn_store *queried,
                      const lmn_leaf_format *fmt_mat, const lmn_leaf_format *fmt_query, const uint8_t *paths, size_t paths_len,
                      const uint8_t *root, size_t root_len, uint8_t *out, size_t cap, size_t *len) {
    if (!c || !md || !mat_r || !mat_z || !queried || !fmt_mat || !len) return fail(LMN_ERR_INVALID, "null argument");
    if ((paths_len && !paths) || (root_len && !root)) return fail(LMN_ERR_INVALID, "null paths / root with non-zero length");
    if (mat_r->ctx != c || mat_z->ctx != c || queried->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    if (mat_r->level != mat_z->level || mat_r->n != mat_z->n) return fail(LMN_ERR_INVALID, "MatR and MatZ must have one shape");
    if (!fmt_query) fmt_query = fmt_mat;
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    // LigeroMetadata.WriteTo (ligero.go:757-763): u32 rows, u32 cols, u8 rhoInv, u16 queries, little-endian (Go's conversions truncate)
    constexpr size_t MD = 11;
    size_t off = MD, n = 0;
    const lmn_store *parts[3] = {mat_r, mat_z, queried};
    const lmn_leaf_format *fm[3] = {fmt_mat, fmt_mat, fmt_query};
    for (int i = 0; i < 3; i++) {
        uint8_t *dst = out ? out + std::min(off, cap) : nullptr;
        if (int rc = store_marshal_locked(c, parts[i], 0, parts[i]->count, fm[i], dst, out && cap > off ? cap - off : 0, &n)) return rc;
        off += n;
    }
    const size_t total = off + paths_len + root_len;
    *len = total;
    if (!out) return 0;
    if (cap < total) return fail(LMN_ERR_INVALID, "output buffer too small for the proof");
    const uint32_t rows = md->rows, cols = md->cols;
    const uint16_t queries = (uint16_t)md->queries;
    for (int i = 0; i < 4; i++) { out[i] = (uint8_t)(rows >> (8 * i)); out[4 + i] = (uint8_t)(cols >> (8 * i)); }
    out[8] = (uint8_t)md->rho_inv;
    out[9] = (uint8_t)queries; out[10] = (uint8_t)(queries >> 8);
    if (paths_len) memcpy(out + off, paths, paths_len);                    // MerklePath.WriteTo, core/tree.go:270-277
    if (root_len) memcpy(out + off + paths_len, root, root_len);           // buf.Write(p.Root), ligero.go:698
    return 0;
}

// leaf hashes (level 0) computed on the device into d_nodes[0 .. n); caller holds the lock
static cudaError_t hash_leaves_dev(lmn_ctx *c, const lmn_store *leaves, const LeafFormat &F, uint32_t *d_nodes) {
    const uint32_t n = leaves->count;
    const size_t bytes = (size_t)n * F.padded_len;
    if (c->ws_ser_cap < bytes) {
        cudaFree(c->ws_ser); c->ws_ser = nullptr; c->ws_ser_cap = 0;
        cudaError_t e = cudaMalloc(&c->ws_ser, bytes);
        if (e != cudaSuccess) return e;
        c->ws_ser_cap = bytes;
    }
    Launch L = c->L();
    launch_leaf_serialize(L, F, leaves->d, leaves->words(), n, c->ws_ser);
    launch_sha256_leaves(L, c->ws_ser, F.padded_len, n, d_nodes);
    return cudaGetLastError();
}
// parent levels on the device from the n leaf hashes in d_nodes, the whole tree copied to the host object
static cudaError_t finish_tree(lmn_ctx *c, lmn_merkle *m, uint32_t *d_nodes, uint32_t total_nodes) {
    Launch L = c->L();
    for (size_t l = 0; l + 1 < m->level_cnt.size(); l++)
        launch_sha256_nodes(L, d_nodes + (size_t)m->level_off[l] * 8, m->level_cnt[l], d_nodes + (size_t)m->level_off[l + 1] * 8);
    cudaError_t e = cudaGetLastError();
    m->nodes.resize((size_t)total_nodes * 32);
    if (e == cudaSuccess) e = cudaMemcpyAsync(m->nodes.data(), d_nodes, m->nodes.size(), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    return e;
}
static lmn_merkle *new_tree_shape(uint32_t n, uint32_t *total_nodes) {
    auto *m = new lmn_merkle();
    m->n_leaves = n;
    uint32_t total = 0;
    for (uint32_t cnt = n;; cnt = (cnt + 1) / 2) { m->level_off.push_back(total); m->level_cnt.push_back(cnt); total += cnt; if (cnt == 1) break; }
    *total_nodes = total;
    return m;
}

int lmn_merkle_build(lmn_ctx *c, const lmn_store *leaves, const lmn_leaf_format *fmt, lmn_merkle **out) {
    if (!c || !leaves || !fmt || !out) return fail(LMN_ERR_INVALID, "null argument");
    if (leaves->count == 0) return fail(LMN_ERR_INVALID, "no leaves");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    LeafFormat F; uint8_t *d_hdr = nullptr;
    if (int rc = make_leaf_format(c, leaves, fmt, F, &d_hdr)) return rc;
    uint32_t total_nodes = 0;
    lmn_merkle *m = new_tree_shape(leaves->count, &total_nodes);
    uint32_t *d_nodes = nullptr;
    cudaError_t e = cudaMalloc(&d_nodes, (size_t)total_nodes * 32);
    if (e == cudaSuccess) e = hash_leaves_dev(c, leaves, F, d_nodes);
    if (e == cudaSuccess) e = finish_tree(c, m, d_nodes, total_nodes);
    cudaFree(d_nodes); cudaFree(d_hdr);
    if (e != cudaSuccess) { delete m; return fail(LMN_ERR_CUDA, std::string("merkle build: ") + cudaGetErrorString(e)); }
    *out = m;
    return 0;
}
// multi-GPU Commit: every rank hashes the leaves of its own columns ...
int lmn_leaf_hashes(lmn_ctx *c, const lmn_store *leaves, const lmn_leaf_format *fmt, uint8_t *out) {
    if (!c || !leaves || !fmt || !out) return fail(LMN_ERR_INVALID, "null argument");
    if (leaves->count == 0) return fail(LMN_ERR_INVALID, "no leaves");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    LeafFormat F; uint8_t *d_hdr = nullptr;
    if (int rc = make_leaf_format(c, leaves, fmt, F, &d_hdr)) return rc;
    uint32_t *d_nodes = nullptr;
    cudaError_t e = cudaMalloc(&d_nodes, (size_t)leaves->count * 32);
    if (e == cudaSuccess) e = hash_leaves_dev(c, leaves, F, d_nodes);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_nodes, (size_t)leaves->count * 32, cudaMemcpyDeviceToHost, c->stream);
    cudaError_t e2 = cudaStreamSynchronize(c->stream);        // the hashes are home; d_nodes / d_hdr may go
    if (e == cudaSuccess) e = e2;
    cudaFree(d_nodes); cudaFree(d_hdr);
    if (e != cudaSuccess) return fail(LMN_ERR_CUDA, std::string("leaf hashes: ") + cudaGetErrorString(e));
    return 0;
}
// ... and the tree is built from the gathered leaf hashes (in column order)
int lmn_merkle_from_leaf_hashes(lmn_ctx *c, const uint8_t *hashes, uint32_t n, lmn_merkle **out) {
    if (!c || !hashes || !out || n == 0) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    uint32_t total_nodes = 0;
    lmn_merkle *m = new_tree_shape(n, &total_nodes);
    uint32_t *d_nodes = nullptr;
    cudaError_t e = cudaMalloc(&d_nodes, (size_t)total_nodes * 32);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_nodes, hashes, (size_t)n * 32, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = finish_tree(c, m, d_nodes, total_nodes);
    cudaFree(d_nodes);
    if (e != cudaSuccess) { delete m; return fail(LMN_ERR_CUDA, std::string("merkle from hashes: ") + cudaGetErrorString(e)); }
    *out = m;
    return 0;
}
void lmn_merkle_destroy(lmn_merkle *m) { delete m; }
uint32_t lmn_merkle_leaf_count(const lmn_merkle *m) { return m ? m->n_leaves : 0; }
uint32_t lmn_merkle_depth(const lmn_merkle *m) { return m ? (uint32_t)m->level_cnt.size() - 1 : 0; }
int lmn_merkle_root(const lmn_merkle *m, uint8_t *out32) {
    if (!m || !out32) return fail(LMN_ERR_INVALID, "null argument");
    memcpy(out32, m->nodes.data() + (size_t)m->level_off.back() * 32, 32);
    return 0;
}
int lmn_merkle_leaf_hash(const lmn_merkle *m, uint32_t index, uint8_t *out32) {
    if (!m || !out32) return fail(LMN_ERR_INVALID, "null argument");
    if (index >= m->n_leaves) return fail(LMN_ERR_INVALID, "index out of bounds");
    memcpy(out32, m->nodes.data() + (size_t)index * 32, 32);
    return 0;
}
// sibling hashes from the leaf level up (core/tree.go:176-219); the unpaired last node of a level is its own sibling
int lmn_merkle_path(const lmn_merkle *m, uint32_t index, uint8_t *out) {
    if (!m || !out) return fail(LMN_ERR_INVALID, "null argument");
    if (index >= m->n_leaves) return fail(LMN_ERR_INVALID, "index out of bounds");
    uint32_t i = index;
    for (size_t l = 0; l + 1 < m->level_cnt.size(); l++, i >>= 1) {
        uint32_t sib = i ^ 1u;
        if (sib >= m->level_cnt[l]) sib = i;
        memcpy(out + 32 * l, m->nodes.data() + ((size_t)m->level_off[l] + sib) * 32, 32);
    }
    return 0;
}

}  // extern "C"
