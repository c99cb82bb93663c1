// Multi-GPU data plane of the path behind the C ABI (SURVEY 8e, 8b "-lnccl"; one process per GPU): the layout changes of the encoded
// matrix between Encode's coefficient slices and the column blocks every other stage works on, as grouped ncclSend / ncclRecv
// all-to-alls over NVLink with the layout kernels of kernels_comm.cu, and the all-gather of the Merkle leaf hashes.
// There is no compute to fuse the transfers with except the leaf rescale, which lmn_commit_leaves_from_slices pipelines chunk by
// chunk behind the exchange.  NCCL is bound at lmn_comm_init with dlopen("libnccl.so.2"), so the library uses the very NCCL the
// host process already has (a Go host links the system one; under PyTorch it is torch's bundled copy) and single-GPU hosts need none.
#include <dlfcn.h>

#include <algorithm>
#include <cstring>

#include <nccl.h>

#include "api_internal.hpp"
#include "comm.cuh"

namespace {
struct NcclApi {
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl() {
    std::lock_guard<std::mutex> g(g_nccl_mu);
    if (g_nccl.h) return 0;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);        // the copy the process already uses, if any
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(LMN_ERR_UNSUPPORTED, std::string("NCCL not found: ") + dlerror());
#define SYM(field, name) do { *(void **)(&g_nccl.field) = dlsym(h, name); if (!g_nccl.field) return fail(LMN_ERR_UNSUPPORTED, "NCCL symbol missing: " name); } while (0)
    SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommDestroy, "ncclCommDestroy");
    SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd"); SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv");
    SYM(AllGather, "ncclAllGather"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    g_nccl.h = h;
    return 0;
}
}  // namespace
#define NK(call)                                                                                              \
    do {                                                                                                      \
        ncclResult_t r_ = (call);                                                                             \
        if (r_ != ncclSuccess) return fail(LMN_ERR_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r_)); \
    } while (0)

// One grouped set of sends / receives.  The first failure is remembered and the group is ALWAYS closed: a return between
// ncclGroupStart and ncclGroupEnd would leave the calling thread inside an open group, and every later NCCL call of that thread
// (the communicator's destruction included) would be queued into it and never run.
struct NcclGroup {
    ncclResult_t first = ncclSuccess;
    const char *what = "";
    bool open = false;
    NcclGroup() { note(g_nccl.GroupStart(), "ncclGroupStart"); open = first == ncclSuccess; }
    ~NcclGroup() { if (open) g_nccl.GroupEnd(); }
    void note(ncclResult_t r, const char *w) { if (r != ncclSuccess && first == ncclSuccess) { first = r; what = w; } }
    void send(const void *p, size_t words, int peer, ncclComm_t comm, cudaStream_t st) { if (first == ncclSuccess) note(g_nccl.Send(p, words, ncclUint64, peer, comm, st), "ncclSend"); }
    void recv(void *p, size_t words, int peer, ncclComm_t comm, cudaStream_t st) { if (first == ncclSuccess) note(g_nccl.Recv(p, words, ncclUint64, peer, comm, st), "ncclRecv"); }
    int end() {
        if (open) { open = false; note(g_nccl.GroupEnd(), "ncclGroupEnd"); }
        return first == ncclSuccess ? 0 : fail(LMN_ERR_CUDA, std::string(what) + ": " + g_nccl.GetErrorString(first));
    }
};

struct CommState {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    cudaStream_t stream = nullptr;                 // the exchanges run here, beside the kernels of the context's stream
    cudaEvent_t ev_ready = nullptr;                // context stream -> comm stream (inputs produced)
    cudaEvent_t ev_recv[2] = {nullptr, nullptr};   // comm stream -> context stream (staging buffer b received)
    cudaEvent_t ev_free[2] = {nullptr, nullptr};   // context stream -> comm stream (staging buffer b consumed)
    u64 *stage[2] = {nullptr, nullptr}; size_t stage_cap[2] = {0, 0};
    u64 *whole = nullptr; size_t whole_cap = 0;
    u64 *sendbuf = nullptr; size_t send_cap = 0;
    uint32_t *d_idx = nullptr; size_t idx_cap = 0;
    uint8_t *d_hash = nullptr; size_t hash_cap = 0;
};
void comm_state_free(CommState *s) {
    if (!s) return;
    if (s->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(s->comm);
    for (int b = 0; b < 2; b++) { cudaFree(s->stage[b]); if (s->ev_recv[b]) cudaEventDestroy(s->ev_recv[b]); if (s->ev_free[b]) cudaEventDestroy(s->ev_free[b]); }
    cudaFree(s->whole); cudaFree(s->sendbuf); cudaFree(s->d_idx); cudaFree(s->d_hash);
    if (s->ev_ready) cudaEventDestroy(s->ev_ready);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}
size_t comm_state_trim(CommState *s) {
    if (!s) return 0;
    if (s->stream) cudaStreamSynchronize(s->stream);
    const size_t bytes = (s->stage_cap[0] + s->stage_cap[1] + s->whole_cap + s->send_cap) * 8 + s->idx_cap * 4 + s->hash_cap;
    for (int b = 0; b < 2; b++) { cudaFree(s->stage[b]); s->stage[b] = nullptr; s->stage_cap[b] = 0; }
    cudaFree(s->whole); cudaFree(s->sendbuf); cudaFree(s->d_idx); cudaFree(s->d_hash);
    s->whole = s->sendbuf = nullptr; s->d_idx = nullptr; s->d_hash = nullptr;
    s->whole_cap = s->send_cap = s->idx_cap = s->hash_cap = 0;
    return bytes;
}
template <class T> static int grow_dev(T **p, size_t *cap, size_t count) {
    if (*cap >= count) return 0;
    cudaFree(*p); *p = nullptr; *cap = 0;
    cudaError_t e = cudaMalloc(p, count * sizeof(T));
    if (e != cudaSuccess) { cudaGetLastError(); return fail(LMN_ERR_NOMEM, std::string("exchange buffer: ") + cudaGetErrorString(e)); }
    *cap = count;
    return 0;
}
static int need_comm(lmn_ctx *c) {
    if (!c->comm || !c->comm->comm) return fail(LMN_ERR_INVALID, "no communicator: call lmn_comm_init first");
    return 0;
}
// queries handled by rank d: positions d, d + G, ... of the query list (round-robin keeps the 309 queries balanced)
static uint32_t queries_of(uint32_t nq, uint32_t d, uint32_t G) { return nq > d ? (nq - d + G - 1) / G : 0; }

extern "C" {

int lmn_comm_unique_id(uint8_t *out128) {
    if (!out128) return fail(LMN_ERR_INVALID, "null argument");
    if (int rc = load_nccl()) return rc;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NK(g_nccl.GetUniqueId(&id));
    memcpy(out128, &id, 128);
    return 0;
}

int lmn_comm_init(lmn_ctx *c, const uint8_t *id128, int rank, int world) {
    if (!c || !id128 || world < 1 || rank < 0 || rank >= world) return fail(LMN_ERR_INVALID, "bad communicator arguments");
    if (world & (world - 1)) return fail(LMN_ERR_UNSUPPORTED, "the number of ranks must be a power of two");
    if (int rc = load_nccl()) return rc;
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (c->comm) { comm_state_free(c->comm); c->comm = nullptr; }
    CommState *S = new CommState();
    S->rank = rank; S->world = world;
    ncclUniqueId id; memcpy(&id, id128, 128);
    ncclResult_t r = g_nccl.CommInitRank(&S->comm, world, id, rank);
    cudaError_t e = cudaSuccess;
    if (r == ncclSuccess) {
        e = cudaStreamCreateWithFlags(&S->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->ev_ready, cudaEventDisableTiming);
        for (int b = 0; b < 2 && e == cudaSuccess; b++) {
            e = cudaEventCreateWithFlags(&S->ev_recv[b], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->ev_free[b], cudaEventDisableTiming);
        }
    }
    if (r != ncclSuccess || e != cudaSuccess) {
        std::string msg = r != ncclSuccess ? std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r) : std::string("communicator setup: ") + cudaGetErrorString(e);
        S->comm = r == ncclSuccess ? S->comm : nullptr;
        comm_state_free(S);
        return fail(LMN_ERR_CUDA, msg);
    }
    c->comm = S;
    return 0;
}

int lmn_comm_destroy(lmn_ctx *c) {
    if (!c) return fail(LMN_ERR_INVALID, "null context");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (c->comm) { cudaStreamSynchronize(c->comm->stream); comm_state_free(c->comm); c->comm = nullptr; }
    return 0;
}
int lmn_comm_rank(const lmn_ctx *c) { return c && c->comm ? c->comm->rank : 0; }
int lmn_comm_world(const lmn_ctx *c) { return c && c->comm ? c->comm->world : 1; }

// ---- coefficient slices of the queried columns -> whole ciphertexts on the rank that rescales them (Prove without a leaves store) ----
int lmn_exchange_queried(lmn_ctx *c, const void *enc_slice_dev, uint32_t n, const uint32_t *query_idx, uint32_t n_queries, lmn_store *out) {
    if (!c || !enc_slice_dev || !query_idx || !out || n_queries == 0) return fail(LMN_ERR_INVALID, "null argument");
    if (out->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_comm(c)) return rc;
    CommState *S = c->comm;
    const uint32_t G = (uint32_t)S->world, me = (uint32_t)S->rank;
    const uint32_t N = (uint32_t)c->n, ns = N / G, PL = 2 * (uint32_t)c->k;
    if (N % G || ns < 32) return fail(LMN_ERR_INVALID, "a coefficient slice holds at least 32 coefficients");
    const uint32_t mine = queries_of(n_queries, me, G);
    if ((int)out->level != c->k - 1 || out->n != N || out->count < std::max(1u, mine)) return fail(LMN_ERR_INVALID, "output store must hold this rank's queries as max-level ciphertexts");
    for (uint32_t i = 0; i < n_queries; i++) if (query_idx[i] >= n) return fail(LMN_ERR_INVALID, "query index out of range");
    const size_t slice_words = (size_t)PL * ns;
    // send order: destination-major, then the destination's queries in list order
    std::vector<uint32_t> order; order.reserve(n_queries);
    for (uint32_t d = 0; d < G; d++) for (uint32_t p = d; p < n_queries; p += G) order.push_back(query_idx[p]);
    if (int rc = grow_dev(&S->d_idx, &S->idx_cap, (size_t)n_queries)) return rc;
    if (int rc = grow_dev(&S->sendbuf, &S->send_cap, (size_t)n_queries * slice_words)) return rc;
    if (int rc = grow_dev(&S->stage[0], &S->stage_cap[0], (size_t)G * std::max(1u, mine) * slice_words)) return rc;
    if (int rc = stage_begin(c, (size_t)n_queries * 4 + 16)) return rc;
    if (int rc = stage_copy(c, S->d_idx, order.data(), (size_t)n_queries * 4)) return rc;
    if (int rc = stage_end(c)) return rc;
    Launch L = c->L();
    launch_copy_rows(L, (const u64 *)enc_slice_dev, S->d_idx, S->sendbuf, n_queries, slice_words);
    CK(cudaEventRecord(S->ev_ready, c->stream));
    CK(cudaStreamWaitEvent(S->stream, S->ev_ready, 0));
    {
        NcclGroup grp;
        size_t off = 0;
        for (uint32_t d = 0; d < G; d++) {
            const uint32_t cnt = queries_of(n_queries, d, G);
            if (cnt) grp.send(S->sendbuf + off * slice_words, cnt * slice_words, (int)d, S->comm, S->stream);
            off += cnt;
            if (mine) grp.recv(S->stage[0] + (size_t)d * mine * slice_words, mine * slice_words, (int)d, S->comm, S->stream);
        }
        if (int rc = grp.end()) return rc;
    }
    CK(cudaEventRecord(S->ev_recv[0], S->stream));
    CK(cudaStreamWaitEvent(c->stream, S->ev_recv[0], 0));
    if (mine) launch_slices_to_whole(L, S->stage[0], out->d, G, mine, PL, ns);
    CK(cudaGetLastError());
    return 0;
}

}  // extern "C"

// one chunk of the slices -> column-block all-to-all: columns [c0, c0 + cnt) of every rank's block, received into stage[b]
static int exchange_chunk(lmn_ctx *c, const u64 *enc_slice, uint32_t per, uint32_t c0, uint32_t cnt, int b) {
    CommState *S = c->comm;
    const uint32_t G = (uint32_t)S->world;
    const size_t slice_words = (size_t)2 * c->k * (c->n / G);
    NcclGroup grp;
    for (uint32_t d = 0; d < G; d++) {
        grp.send(enc_slice + ((size_t)d * per + c0) * slice_words, cnt * slice_words, (int)d, S->comm, S->stream);
        grp.recv(S->stage[b] + (size_t)d * cnt * slice_words, cnt * slice_words, (int)d, S->comm, S->stream);
    }
    return grp.end();
}
// shared driver: the whole encoded matrix from coefficient slices to this rank's column block, chunk by chunk; consume(chunk) runs on
// the context's stream on the whole ciphertexts of the chunk while the next chunk is in flight on the communicator's stream
template <class F>
static int slices_to_blocks(lmn_ctx *c, const u64 *enc_slice, uint32_t n, uint32_t chunk, u64 *whole_dst, F consume) {
    CommState *S = c->comm;
    const uint32_t G = (uint32_t)S->world, N = (uint32_t)c->n, ns = N / G, PL = 2 * (uint32_t)c->k;
    const uint32_t per = n / G;
    const size_t slice_words = (size_t)PL * ns, ct_words = (size_t)PL * N;
    chunk = std::max(1u, std::min(chunk, per));
    for (int b = 0; b < 2; b++) if (int rc = grow_dev(&S->stage[b], &S->stage_cap[b], (size_t)G * chunk * slice_words)) return rc;
    if (!whole_dst) if (int rc = grow_dev(&S->whole, &S->whole_cap, (size_t)chunk * ct_words)) return rc;
    Launch L = c->L();
    CK(cudaEventRecord(S->ev_ready, c->stream));              // the slices are produced on the context's stream (Encode)
    CK(cudaStreamWaitEvent(S->stream, S->ev_ready, 0));
    uint32_t it = 0;
    for (uint32_t c0 = 0; c0 < per; c0 += chunk, it++) {
        const uint32_t cnt = std::min(chunk, per - c0);
        const int b = (int)(it & 1);
        if (it >= 2) CK(cudaStreamWaitEvent(S->stream, S->ev_free[b], 0));   // the unpack of chunk it - 2 has read this buffer
        if (int rc = exchange_chunk(c, enc_slice, per, c0, cnt, b)) return rc;
        CK(cudaEventRecord(S->ev_recv[b], S->stream));
        CK(cudaStreamWaitEvent(c->stream, S->ev_recv[b], 0));
        u64 *w = whole_dst ? whole_dst + (size_t)c0 * ct_words : S->whole;
        launch_slices_to_whole(L, S->stage[b], w, G, cnt, PL, ns);
        CK(cudaEventRecord(S->ev_free[b], c->stream));
        if (int rc = consume(w, c0, cnt)) return rc;
    }
    CK(cudaGetLastError());
    return 0;
}
static int check_slices(lmn_ctx *c, const void *p, uint32_t n) {
    if (int rc = need_comm(c)) return rc;
    const uint32_t G = (uint32_t)c->comm->world;
    if ((uint32_t)c->n % G || (uint32_t)c->n / G < 32) return fail(LMN_ERR_INVALID, "a coefficient slice holds at least 32 coefficients");
    if (!p || n == 0 || n % G) return fail(LMN_ERR_INVALID, "the number of ciphertexts must divide evenly over the ranks");
    return 0;
}

extern "C" {

int lmn_exchange_encoded(lmn_ctx *c, const void *enc_slice_dev, uint32_t n, lmn_store *block) {
    if (!c || !block) return fail(LMN_ERR_INVALID, "null argument");
    if (block->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = check_slices(c, enc_slice_dev, n)) return rc;
    const uint32_t per = n / (uint32_t)c->comm->world;
    if ((int)block->level != c->k - 1 || block->n != (uint32_t)c->n || block->count < per) return fail(LMN_ERR_INVALID, "the block store holds n / world max-level ciphertexts");
    return slices_to_blocks(c, (const u64 *)enc_slice_dev, n, c->batch, block->d, [](u64 *, uint32_t, uint32_t) { return 0; });
}

// Commit's leaves on column-sharded ranks straight from Encode's coefficient slices: the all-to-all of chunk i + 1 runs on the
// communicator's stream while chunk i is rescaled to level 1 (processLeafParallel, fhe/ligero.go:146-154) on the context's stream
int lmn_commit_leaves_from_slices(lmn_ctx *c, const void *enc_slice_dev, uint32_t n, lmn_store *leaves) {
    if (!c || !leaves) return fail(LMN_ERR_INVALID, "null argument");
    if (leaves->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = check_slices(c, enc_slice_dev, n)) return rc;
    const uint32_t per = n / (uint32_t)c->comm->world;
    const int lout = std::min(1, c->k - 1);
    if ((int)leaves->level != lout || leaves->n != (uint32_t)c->n || leaves->count < per) return fail(LMN_ERR_INVALID, "the leaves store holds n / world level-1 ciphertexts");
    const size_t out_words = 2 * (size_t)(lout + 1) * c->n;
    // at least four chunks, so that the rescale of one overlaps the all-to-all of the next even when the whole block fits one batch
    const uint32_t chunk = std::max(1u, std::min(c->batch, (per + 3) / 4));
    return slices_to_blocks(c, (const u64 *)enc_slice_dev, n, chunk, nullptr,
                            [&](u64 *w, uint32_t c0, uint32_t cnt) { return rescale_consecutive(c, w, cnt, leaves->d + (size_t)c0 * out_words); });
}

// the inverse layout change: this rank's block of whole ciphertexts -> its coefficient slice of every ciphertext (a matrix uploaded
// once as column blocks feeds the slice-wise Encode)
int lmn_exchange_to_slices(lmn_ctx *c, const lmn_store *block, uint32_t n, void *slice_dev) {
    if (!c || !block) return fail(LMN_ERR_INVALID, "null argument");
    if (block->ctx != c) return fail(LMN_ERR_INVALID, "objects belong to different contexts");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = check_slices(c, slice_dev, n)) return rc;
    CommState *S = c->comm;
    const uint32_t G = (uint32_t)S->world, N = (uint32_t)c->n, ns = N / G, PL = 2 * ((uint32_t)block->level + 1);
    const uint32_t per = n / G;
    if (block->n != N || block->count < per) return fail(LMN_ERR_INVALID, "the block store holds n / world ciphertexts of the context ring");
    const size_t slice_words = (size_t)PL * ns, ct_words = (size_t)PL * N;
    const uint32_t chunk = std::max(1u, std::min(c->batch, per));
    for (int b = 0; b < 2; b++) if (int rc = grow_dev(&S->stage[b], &S->stage_cap[b], (size_t)G * chunk * slice_words)) return rc;
    Launch L = c->L();
    u64 *out = (u64 *)slice_dev;
    uint32_t it = 0;
    for (uint32_t c0 = 0; c0 < per; c0 += chunk, it++) {
        const uint32_t cnt = std::min(chunk, per - c0);
        const int b = (int)(it & 1);
        if (it >= 2) CK(cudaStreamWaitEvent(c->stream, S->ev_free[b], 0));    // the sends of chunk it - 2 have left this buffer
        launch_whole_to_slices(L, block->d + (size_t)c0 * ct_words, S->stage[b], G, cnt, PL, ns);
        CK(cudaEventRecord(S->ev_recv[b], c->stream));
        CK(cudaStreamWaitEvent(S->stream, S->ev_recv[b], 0));
        {
            NcclGroup grp;
            for (uint32_t d = 0; d < G; d++) {
                grp.send(S->stage[b] + (size_t)d * cnt * slice_words, cnt * slice_words, (int)d, S->comm, S->stream);
                grp.recv(out + ((size_t)d * per + c0) * slice_words, cnt * slice_words, (int)d, S->comm, S->stream);
            }
            if (int rc = grp.end()) return rc;
        }
        CK(cudaEventRecord(S->ev_free[b], S->stream));
    }
    CK(cudaEventRecord(S->ev_ready, S->stream));
    CK(cudaStreamWaitEvent(c->stream, S->ev_ready, 0));        // the slices are complete for whatever the context's stream runs next
    CK(cudaGetLastError());
    return 0;
}

// Merkle leaf hashes of all ranks in column order (32 bytes per leaf; every rank then builds the same tree)
int lmn_allgather_leaf_hashes(lmn_ctx *c, const uint8_t *mine, uint32_t n_local, uint8_t *all) {
    if (!c || !mine || !all || n_local == 0) return fail(LMN_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    if (int rc = need_comm(c)) return rc;
    CommState *S = c->comm;
    const size_t bytes = (size_t)n_local * 32, total = bytes * S->world;
    if (int rc = grow_dev(&S->d_hash, &S->hash_cap, bytes + total)) return rc;
    CK(cudaMemcpyAsync(S->d_hash, mine, bytes, cudaMemcpyHostToDevice, S->stream));
    NK(g_nccl.AllGather(S->d_hash, S->d_hash + bytes, bytes, ncclUint8, S->comm, S->stream));
    CK(cudaMemcpyAsync(all, S->d_hash + bytes, total, cudaMemcpyDeviceToHost, S->stream));
    CK(cudaStreamSynchronize(S->stream));
    return 0;
}

}  // extern "C"
