// Batched host <-> device copies for the layout the Go host really holds: rlwe.Ciphertext.Value[p].Coeffs[limb] is one Go slice per
// limb (core/utils.go:12-33, vdec/prover.go:110-112), pageable and scattered over the heap.  lmn_store_upload_limb / download_limb
// move one limb per call with a synchronisation each (98 304 calls for config D's matrix); the two entry points here take the whole
// pointer table at once and pipeline it: limbs are packed into (unpacked from) two alternating pinned staging buffers by a few host
// threads while the previous buffer's single cudaMemcpyAsync is in flight on the context's copy stream.  No synchronisation per limb.
#include <algorithm>
#include <cstring>
#include <thread>

#include "api_internal.hpp"

namespace {

constexpr size_t STAGE_BYTES = 32u << 20;     // per staging buffer
constexpr int COPY_THREADS = 8;               // host threads packing / unpacking one chunk (a single core's memcpy is ~10 GB/s)

struct LimbStage {                            // per context, created on first use, lives until the context is destroyed
    uint8_t *buf[2] = {nullptr, nullptr};
    cudaEvent_t done[2] = {nullptr, nullptr}; // the DMA out of / into buf[b] has finished
    cudaEvent_t order = nullptr;              // context stream <-> copy stream
};

int stage_of(lmn_ctx *c, LimbStage **out) {
    if (!c->limb_stage) {
        auto *S = new LimbStage();
        cudaError_t e = cudaSuccess;
        for (int b = 0; b < 2 && e == cudaSuccess; b++) {
            e = cudaHostAlloc((void **)&S->buf[b], STAGE_BYTES, cudaHostAllocDefault);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->done[b], cudaEventDisableTiming);
        }
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&S->order, cudaEventDisableTiming);
        c->limb_stage = S;
        c->limb_stage_free = [](void *p) {
            auto *S = static_cast<LimbStage *>(p);
            for (int b = 0; b < 2; b++) { if (S->buf[b]) cudaFreeHost(S->buf[b]); if (S->done[b]) cudaEventDestroy(S->done[b]); }
            if (S->order) cudaEventDestroy(S->order);
            delete S;
        };
        if (e != cudaSuccess) { cudaGetLastError(); return fail(LMN_ERR_NOMEM, std::string("pinned staging for limb copies: ") + cudaGetErrorString(e)); }
    }
    *out = static_cast<LimbStage *>(c->limb_stage);
    if (!(*out)->buf[1] || !(*out)->order) return fail(LMN_ERR_NOMEM, "pinned staging for limb copies is not available");
    return 0;
}

// limbs [l0, l1) of the table <-> consecutive slots of the staging buffer, split over a few host threads
template <bool TO_STAGE, class P>
void pack(uint8_t *stage, P limb_ptrs, size_t l0, size_t l1, size_t limb_bytes) {
    const size_t n = l1 - l0;
    auto work = [&](size_t a, size_t b) {
        for (size_t i = a; i < b; i++) {
            if (TO_STAGE) memcpy(stage + i * limb_bytes, (const void *)limb_ptrs[l0 + i], limb_bytes);
            else memcpy((void *)limb_ptrs[l0 + i], stage + i * limb_bytes, limb_bytes);
        }
    };
    const int nt = (int)std::min<size_t>(COPY_THREADS, n * limb_bytes >> 21 ? n : 1);   // small chunks: the calling thread alone
    if (nt <= 1) { work(0, n); return; }
    std::thread th[COPY_THREADS];
    for (int t = 1; t < nt; t++) th[t] = std::thread(work, n * t / nt, n * (t + 1) / nt);
    work(0, n / nt);
    for (int t = 1; t < nt; t++) th[t].join();
}

int check_args(const lmn_store *s, uint32_t first, uint32_t count, const void *const *ptrs, size_t *n_limbs) {
    if (!s || !ptrs) return fail(LMN_ERR_INVALID, "null argument");
    if (count == 0 || (uint64_t)first + count > s->count) return fail(LMN_ERR_INVALID, "ciphertext range out of bounds");
    *n_limbs = (size_t)count * 2 * (s->level + 1);
    for (size_t i = 0; i < *n_limbs; i++) if (!ptrs[i]) return fail(LMN_ERR_INVALID, "null limb pointer in the table (entry " + std::to_string(i) + ")");
    if ((size_t)s->n * 8 > STAGE_BYTES) return fail(LMN_ERR_UNSUPPORTED, "limb larger than the staging buffer");
    return 0;
}

}  // namespace

extern "C" {

int lmn_store_upload_limbs(lmn_store *s, uint32_t first, uint32_t count, const uint64_t *const *limb_ptrs) {
    size_t n_limbs = 0;
    if (int rc = check_args(s, first, count, (const void *const *)limb_ptrs, &n_limbs)) return rc;
    lmn_ctx *c = s->ctx;
    std::lock_guard<std::mutex> g(c->mu);
    CK(cudaSetDevice(c->device));
    LimbStage *S = nullptr;
    if (int rc = stage_of(c, &S)) return rc;
    const size_t limb_bytes = (size_t)s->n * 8, per = STAGE_BYTES / limb_bytes;
    uint8_t *dst = (uint8_t *)(s->d + (size_t)first * s->words());
    CK(cudaEventRecord(S->order, c->stream));                    // kernels already queued may still read the store
    CK(cudaStreamWaitEvent(c->copy_stream, S->order, 0));
    cudaError_t e = cudaSuccess;
    int b = 0;
    for (size_t l0 = 0; l0 < n_limbs && e == cudaSuccess; l0 += per, b ^= 1) {
        const size_t l1 = std::min(n_limbs, l0 + per);
        e = cudaEventSynchronize(S->done[b]);                    // the DMA that last read this buffer (a never-recorded event is complete)
        if (e != cudaSuccess) break;
        pack<true>(S->buf[b], limb_ptrs, l0, l1, limb_bytes);    // overlaps the DMA of the other buffer
        e = cudaMemcpyAsync(dst + l0 * limb_bytes, S->buf[b], (l1 - l0) * limb_bytes, cudaMemcpyHostToDevice, c->copy_stream);
        if (e == cudaSuccess) e = cudaEventRecord(S->done[b], c->copy_stream);
    }
    // the caller's slices are free again (everything was copied into the staging buffers); work queued on the context's stream after
    // this call sees the uploaded data
    if (e == cudaSuccess) e = cudaEventRecord(S->order, c->copy_stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, S->order, 0);
    if (e != cudaSuccess) { cudaStreamSynchronize(c->copy_stream); return fail(LMN_ERR_CUDA, std::string("upload of limbs: ") + cudaGetErrorString(e)); }
    return 0;
}

int lmn_store_download_limbs(const lmn_store *s, uint32_t first, uint32_t count, uint64_t *const *limb_ptrs) {
    size_t n_limbs = 0;
    if (int rc = check_args(s, first, count, (const void *const *)limb_ptrs, &n_limbs)) return rc;
    lmn_ctx *c = s->ctx;
    std::lock_guard<std::mutex> g(c->mu);
    CK(cudaSetDevice(c->device));
    LimbStage *S = nullptr;
    if (int rc = stage_of(c, &S)) return rc;
    const size_t limb_bytes = (size_t)s->n * 8, per = STAGE_BYTES / limb_bytes;
    const uint8_t *src = (const uint8_t *)(s->d + (size_t)first * s->words());
    CK(cudaEventRecord(S->order, c->stream));                    // the kernels that produce the store
    CK(cudaStreamWaitEvent(c->copy_stream, S->order, 0));
    cudaError_t e = cudaSuccess;
    for (int b = 0; b < 2 && e == cudaSuccess; b++) e = cudaEventSynchronize(S->done[b]);   // an earlier upload may still read the buffers
    size_t prev0 = 0, prev1 = 0; int pb = 0, b = 0;
    for (size_t l0 = 0; l0 < n_limbs && e == cudaSuccess; l0 += per, b ^= 1) {
        const size_t l1 = std::min(n_limbs, l0 + per);
        e = cudaMemcpyAsync(S->buf[b], src + l0 * limb_bytes, (l1 - l0) * limb_bytes, cudaMemcpyDeviceToHost, c->copy_stream);
        if (e == cudaSuccess) e = cudaEventRecord(S->done[b], c->copy_stream);
        if (e == cudaSuccess && prev1 > prev0) {                 // unpack the previous chunk while this one is in flight
            e = cudaEventSynchronize(S->done[pb]);
            if (e == cudaSuccess) pack<false>(S->buf[pb], limb_ptrs, prev0, prev1, limb_bytes);
        }
        prev0 = l0; prev1 = l1; pb = b;
    }
    if (e == cudaSuccess && prev1 > prev0) {
        e = cudaEventSynchronize(S->done[pb]);
        if (e == cudaSuccess) pack<false>(S->buf[pb], limb_ptrs, prev0, prev1, limb_bytes);
    }
    if (e != cudaSuccess) { cudaStreamSynchronize(c->copy_stream); return fail(LMN_ERR_CUDA, std::string("download of limbs: ") + cudaGetErrorString(e)); }
    return 0;
}

}  // extern "C"
