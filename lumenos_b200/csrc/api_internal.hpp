// Private definitions shared by the translation units that implement the C ABI (api.cu, api_client.cu, api_comm.cu).
#pragma once
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/lumenos_b200.h"
#include "hostmath.hpp"
#include "kernels.cuh"

using namespace lmn;

int lmn_fail(int code, const std::string &msg);     // records the calling thread's error text, returns code
#define fail lmn_fail
#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) return fail(LMN_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// device scratch that is released on every exit path of the function that owns it
struct DevScratch {
    void *p = nullptr;
    ~DevScratch() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes); }
    template <class T> T *as() const { return static_cast<T *>(p); }
    void *release() { void *r = p; p = nullptr; return r; }
};

struct CommState;
struct ClientState;     // plaintext ring, public / secret key material of the client-side entry points (api_client.cu)
void client_state_free(ClientState *);
// pooled batch workspaces released (the state itself, keys and tables stay); return the bytes given back.  Caller holds the context lock
// and has drained the streams.
size_t client_state_trim(ClientState *);
size_t comm_state_trim(CommState *);

struct lmn_ctx {
    int device = 0;
    ClientState *client = nullptr;
    CommState *comm = nullptr;
    int logn = 0, n = 0, k = 0, np = 0;
    u64 t = 0;
    std::vector<u64> q, p;                  // moduli: mod(i) = i < k ? q[i] : p[i-k]
    bool scalar_centered = true, round_nearest = true, rowswap_last = true;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaStream_t copy_stream = nullptr;     // host <-> device copies that overlap the kernels (lmn_encode_host, lmn_inner_product2_host)
    std::vector<cudaEvent_t> copy_events;   // pool, reused call after call
    // small host inputs (plaintexts, index lists) go through this pinned buffer, so the caller's memory is free to be reused or
    // released the moment a call returns even when it is pinned (an async copy from pinned memory runs behind the queued kernels)
    uint8_t *h_stage = nullptr; size_t h_stage_cap = 0, h_stage_cur = 0; cudaEvent_t stage_ev = nullptr; bool stage_pending = false;
    ModDev *d_mods = nullptr;
    ulonglong2 *d_twf = nullptr, *d_twi = nullptr, *d_twbf = nullptr, *d_twbi = nullptr;
    DeviceTables T{};
    std::map<int, std::pair<KsConsts, KsConsts *>> ks;      // per level
    std::map<int, RescaleConsts *> rs;                      // per level
    std::map<int, std::pair<RescaleChain, RescaleChain *>> rs_chain;   // per (from, to)
    std::map<u64, u64 *> galois;                            // galEl -> device evk (rows pre-scaled, see key_scale)
    int key_scale[MAX_DIGITS][MAX_LIMBS];                   // s(d, gm): evk rows of non-own (d, gm) carry the factor 2^s
    u64 *d_key_fm = nullptr;                                // [beta][k+np] 2^s in Montgomery form (0 = own row)
    // ring switch
    int logn_small = 0;
    u64 *d_rs_evk = nullptr;
    ModDev *d_small_mod = nullptr;
    ulonglong2 *d_small_tw = nullptr;
    // workspace
    uint32_t batch = 64;
    u64 *ws_col = nullptr, *ws_col2 = nullptr, *ws_y = nullptr, *ws_acc = nullptr, *ws_mu = nullptr, *ws_pt = nullptr;
    uint32_t *ws_idx = nullptr; size_t ws_idx_cap = 0;
    u64 *ws_ser = nullptr; size_t ws_ser_cap = 0;           // serialised leaves of the Merkle build (pooled: no allocation per Commit)
    ulonglong2 *ws_scalar = nullptr;
    uint32_t ws_jobs = 0;
    void *limb_stage = nullptr; void (*limb_stage_free)(void *) = nullptr;   // pinned staging of lmn_store_upload_limbs / download_limbs (api_limbs.cu)
    uint64_t launches = 0;
    Profiler prof;
    std::mutex mu;

    u64 mod(int i) const { return i < k ? q[i] : p[i - k]; }
    size_t ct_words(int level) const { return 2 * (size_t)(level + 1) * n; }
    Launch L() { return Launch{stream, &launches, &prof}; }
};

struct lmn_store {
    lmn_ctx *ctx;
    u64 *d = nullptr;
    bool owned = true;
    uint32_t count = 0, level = 0;
    uint32_t n = 0;   // ring degree of the entries (ctx->n, or the small ring for ring-switch outputs)
    size_t words() const { return 2 * (size_t)(level + 1) * n; }
};

struct CommState;       // NCCL communicator and exchange buffers (api_comm.cu)
void comm_state_free(CommState *);
int rescale_consecutive(lmn_ctx *c, const u64 *src, uint32_t count, u64 *out);   // api.cu; caller holds the context lock

// staged host -> device copies on the context's stream (api.cu); caller holds the context lock
int stage_begin(lmn_ctx *c, size_t total);
int stage_copy(lmn_ctx *c, void *dst, const void *src, size_t bytes);
int stage_end(lmn_ctx *c);
void build_mod(ModDev &m, u64 q, int n);
void build_twiddles(u64 q, u64 psi, int logn, std::vector<ulonglong2> &fwd, std::vector<ulonglong2> &inv);
