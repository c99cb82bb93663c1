// lmn_merkle_from_leaf_hashes_noalloc: the Merkle tree over gathered leaf hashes (core.NewTree, core/tree.go:76-163) for hosts that
// must not allocate.  It is the twin of lmn_merkle_from_leaf_hashes (api.cu), which cudaMallocs the node array of the tree for every
// call and news an lmn_merkle object; here the node array lives in the file-scope device array below and the results go straight
// into buffers the caller owns.
//
// Who needs this.  A prover service sizes its HBM once: the matrix, the encoded matrix, the leaves and the key-switch workspaces
// of one context already take tens of GiB at the benchmark shape, and what is left is usually handed to another pool (a second
// context, torch's caching allocator, another tenant of an MPS / MIG partition).  In such a process a cudaMalloc in the middle of
// Commit can fail although the proof itself would fit, and even when it succeeds cudaMalloc / cudaFree stall the device until every
// stream has drained - which stops the kernels another host thread has in flight on the same GPU (the reference runs its R and Z
// inner products from two goroutines, fhe/ligero.go:231-242).  Some hosts are simply not allowed to call the allocator at all once
// they have entered their steady-state loop.  This entry point gives them the tree without touching cudaMalloc, cudaFree, new or any
// other allocating routine on its success path (a failing call formats its message into the thread's std::string, host heap only).
//
// Cost and limits.
//   - The scratch is LMN_MERKLE_NOALLOC_MAX_LEAVES leaves' worth of tree nodes (2.0 MiB).  It is part of the library image: every
//     GPU on which the process loads this library's kernels holds the 2 MiB until the process exits, whether or not the entry point
//     is ever called.  More leaves than the maximum are refused with LMN_ERR_INVALID; use lmn_merkle_from_leaf_hashes for those.
//   - The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device current) and keeps it;
//     later calls on that device reuse it.  Make that first call during set-up, before any stream capture is begun.
//   - One array per device is shared by every call, every context and every host thread of the process.  Two uses that overlapped
//     on the GPU would race on it (one tree's parents overwriting the other's leaves).  The entry point therefore takes a
//     process-wide lock and waits for its own work to finish before it returns: calls from several host threads are safe and simply
//     run one after the other.  Do not expect it to overlap with itself, and do not read g_tree_nodes from code of your own.
//   - The array keeps the previous call's nodes.  Nothing in it has to start at zero: level 0 is overwritten by the copy of the
//     caller's hashes and every parent is written by k_sha256_nodes before the level above reads it, and only the first
//     `total nodes of this call` entries are ever read.  tests/test_zz_no_alloc.py checks a small tree built right after a large one.
#include <cstring>
#include <mutex>

#include "api_internal.hpp"
#include "commit.cuh"

namespace {

constexpr uint32_t MAX_LEAVES = LMN_MERKLE_NOALLOC_MAX_LEAVES;
// levels hold n, ceil(n/2), ceil(n/4), ..., 1 nodes: fewer than 2n + 32 in total for any n < 2^32
constexpr uint32_t MAX_NODES = 2 * MAX_LEAVES + 32;
constexpr int MAX_DEVICES = 64;

// Tree nodes, 8 words (32 bytes) each, levels back to back from the leaf hashes up.  In HOST code this name is only a handle for
// cudaGetSymbolAddress; the pointer a kernel may be given is the one scratch_on_device() returns.
static __device__ __align__(256) uint32_t g_tree_nodes[(size_t)MAX_NODES * 8];

std::mutex g_mu;                                  // guards g_addr and the array's contents (one user at a time)
uint32_t *g_addr[MAX_DEVICES] = {};               // device address of g_tree_nodes on each device, looked up once

// caller holds g_mu and has made `device` current
int scratch_on_device(int device, uint32_t **out) {
    if (device < 0 || device >= MAX_DEVICES) return fail(LMN_ERR_UNSUPPORTED, "no-alloc Merkle scratch: device ordinal beyond 63");
    if (!g_addr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_tree_nodes);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(LMN_ERR_CUDA, std::string("cudaGetSymbolAddress(no-alloc Merkle scratch): ") + cudaGetErrorString(e));
        }
        g_addr[device] = static_cast<uint32_t *>(p);
    }
    *out = g_addr[device];
    return 0;
}

}  // namespace

extern "C" int lmn_merkle_from_leaf_hashes_noalloc(lmn_ctx *c, const uint8_t *hashes, uint32_t n_leaves, uint8_t *root_out32,
                                                   uint8_t *nodes_out, size_t nodes_cap) {
    if (!c || !hashes || !root_out32) return fail(LMN_ERR_INVALID, "null argument");
    if (n_leaves == 0) return fail(LMN_ERR_INVALID, "no leaves");
    if (n_leaves > MAX_LEAVES)
        return fail(LMN_ERR_INVALID, "lmn_merkle_from_leaf_hashes_noalloc: " + std::to_string(n_leaves) + " leaves exceed the static scratch of " +
                                         std::to_string(MAX_LEAVES) + " (LMN_MERKLE_NOALLOC_MAX_LEAVES); use lmn_merkle_from_leaf_hashes");
    uint32_t total = 0;
    for (uint32_t cnt = n_leaves;; cnt = (cnt + 1) / 2) { total += cnt; if (cnt == 1) break; }
    if (nodes_out && nodes_cap < (size_t)total * 32)
        return fail(LMN_ERR_INVALID, "nodes_out holds " + std::to_string(nodes_cap) + " bytes, the tree has " + std::to_string((size_t)total * 32));

    std::lock_guard<std::mutex> gc(c->mu);        // the context's stream and launch counter
    std::lock_guard<std::mutex> gs(g_mu);         // the shared array, until the results are on the host
    CK(cudaSetDevice(c->device));
    uint32_t *d_nodes = nullptr;
    if (int rc = scratch_on_device(c->device, &d_nodes)) return rc;

    cudaError_t e = cudaMemcpyAsync(d_nodes, hashes, (size_t)n_leaves * 32, cudaMemcpyHostToDevice, c->stream);
    const Launch L{c->stream, &c->launches, nullptr};      // no profiler: its event pool grows on demand
    uint32_t off = 0;                                      // first node of the level being read; the root's index at the end
    for (uint32_t cnt = n_leaves; e == cudaSuccess && cnt > 1; cnt = (cnt + 1) / 2) {
        launch_sha256_nodes(L, d_nodes + (size_t)off * 8, cnt, d_nodes + (size_t)(off + cnt) * 8);
        e = cudaGetLastError();
        off += cnt;
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(root_out32, d_nodes + (size_t)off * 8, 32, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && nodes_out) e = cudaMemcpyAsync(nodes_out, d_nodes, (size_t)total * 32, cudaMemcpyDeviceToHost, c->stream);
    // also on a failure: nothing queued above may still be using the array or the caller's buffers once the locks are dropped
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) return fail(LMN_ERR_CUDA, std::string("merkle from hashes (no-alloc): ") + cudaGetErrorString(e));
    return 0;
}
