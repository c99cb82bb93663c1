"""Multi-GPU partitioning of the hot path (one process per GPU, torch.distributed for the plumbing).

SURVEY 8(e): the inner products, the query rescale and the leaf rescale are independent per column ciphertext
(fhe/ligero.go:318-345 has no cross-ciphertext reduction), so columns shard by rank with no data-path collective.
Encode is independent per (poly, limb, coefficient) lane (fhe/ntt.go applies only ct+ct, ct-ct, ct*scalar), so it shards
by coefficient slice.  The real exchanges of the path are therefore layout changes of the ENCODED matrix from coefficient
slices to whole ciphertexts: of the few hundred queried columns (exchange_queried, Prove without a leaves store) or of all
columns (exchange_encoded, Commit's leaf rescale + Merkle span, after which the query stage is a local gather from the
leaves), plus the gather of 32-byte leaf hashes for the tree (commit_tree).

The exchanges themselves live behind the C ABI (lmn_comm_* / lmn_exchange_* / lmn_commit_leaves_from_slices /
lmn_allgather_leaf_hashes in include/lumenos_b200.h: NCCL send/recv groups + layout kernels on the communicator's stream), so a Go
host can drive them; the functions here only bootstrap the communicator and call them.  The `*_torch` twins are the same layout
changes written with torch.distributed: they run on CPU tensors under gloo (tests/test_sharding_gloo.py pins the index arithmetic
without a GPU) and serve as an independent cross-check of the library path in the multi-GPU tests."""
import numpy as np
import torch
import torch.distributed as dist


def column_range(cols, rank, world):
    """contiguous block of columns owned by `rank` (inner products, fhe/ligero.go:299-370)"""
    if cols % world:
        raise ValueError("cols must be divisible by the number of ranks")
    per = cols // world
    return rank * per, (rank + 1) * per


def coefficient_range(n_ring, rank, world):
    """contiguous coefficient slice of every limb owned by `rank` during Encode"""
    if n_ring % world or n_ring // world < 32:
        raise ValueError("ring degree must split into slices of at least 32 coefficients")
    per = n_ring // world
    return rank * per, (rank + 1) * per


def queries_of(n_queries, rank, world):
    """positions in the query list handled by `rank` (round-robin keeps the 309 queries balanced)"""
    return list(range(rank, n_queries, world))


def init_comm(backend, rank, world, group=None):
    """lmn_comm_init on every rank: rank 0 draws the NCCL unique id, torch.distributed (any backend) carries it to the others."""
    import ctypes as C
    from . import _capi
    buf = C.create_string_buffer(128)
    if rank == 0:
        _capi.check(_capi.lib().lmn_comm_unique_id(buf))
    box = [buf.raw]
    if world > 1:
        dist.broadcast_object_list(box, src=0, group=group)
    _capi.check(_capi.lib().lmn_comm_init(backend.ctx, box[0], rank, world))


def exchange_queried(backend, enc_slice, query_idx, out=None):
    """enc_slice: device tensor [n, 2, k, N/world]; returns a store of the whole max-level ciphertexts of this rank's queries"""
    import ctypes as C
    from . import _capi, fhe
    lib = _capi.lib()
    rank, world = lib.lmn_comm_rank(backend.ctx), lib.lmn_comm_world(backend.ctx)
    mine = len(queries_of(len(query_idx), rank, world))
    out = fhe.CiphertextStore(backend, max(1, mine), backend.params.MaxLevel()) if out is None else out
    idx = np.ascontiguousarray(query_idx, dtype=np.uint32)
    _capi.check(lib.lmn_exchange_queried(backend.ctx, C.c_void_p(enc_slice.data_ptr()), enc_slice.shape[0], idx.ctypes.data_as(C.c_void_p), len(idx), out.h))
    return out


def exchange_encoded(backend, enc_slice, out=None):
    """slices [n, 2, k, N/world] -> store of this rank's column block (n / world whole ciphertexts)"""
    import ctypes as C
    from . import _capi, fhe
    world = _capi.lib().lmn_comm_world(backend.ctx)
    out = fhe.CiphertextStore(backend, enc_slice.shape[0] // world, backend.params.MaxLevel()) if out is None else out
    _capi.check(_capi.lib().lmn_exchange_encoded(backend.ctx, C.c_void_p(enc_slice.data_ptr()), enc_slice.shape[0], out.h))
    return out


def commit_leaves_from_slices(backend, enc_slice, leaves=None):
    """Commit's leaves of this rank's column block, pipelined behind the all-to-all (lmn_commit_leaves_from_slices)"""
    import ctypes as C
    from . import _capi, fhe
    world = _capi.lib().lmn_comm_world(backend.ctx)
    lout = min(1, backend.params.MaxLevel())
    leaves = fhe.CiphertextStore(backend, enc_slice.shape[0] // world, lout) if leaves is None else leaves
    _capi.check(_capi.lib().lmn_commit_leaves_from_slices(backend.ctx, C.c_void_p(enc_slice.data_ptr()), enc_slice.shape[0], leaves.h))
    leaves.scale = backend.RescaleScale(backend.params.MaxLevel(), lout)
    return leaves


def to_slices(backend, block, n, out):
    """column block store (n / world ciphertexts) -> this rank's coefficient slice of all n ciphertexts, written into the device tensor out"""
    import ctypes as C
    from . import _capi
    _capi.check(_capi.lib().lmn_exchange_to_slices(backend.ctx, block.h, n, C.c_void_p(out.data_ptr())))
    return out


def commit_tree(backend, leaves, leaf_format):
    """Commit's Merkle span on column-sharded leaves (fhe/ligero.go:105-116): every rank hashes its own leaves on its GPU, the
    32-byte hashes are all-gathered in column order (lmn_allgather_leaf_hashes) and every rank builds the same tree from them."""
    import ctypes as C
    from . import _capi, fhe
    lib = _capi.lib()
    world = lib.lmn_comm_world(backend.ctx)
    mine = fhe.LeafHashes(backend, leaves, leaf_format)
    if world == 1:
        return fhe.MerkleTree(backend, None, None, leafHashes=mine)
    allh = C.create_string_buffer(len(mine) * world)
    _capi.check(lib.lmn_allgather_leaf_hashes(backend.ctx, mine, leaves.count, allh))
    return fhe.MerkleTree(backend, None, None, leafHashes=allh.raw)


def exchange_queried_torch(enc_slice, query_idx, rank, world, group=None):
    """enc_slice: [n, 2, k, N/world] this rank's coefficient slice of the encoded matrix;
    query_idx: the full list of queried column indices (identical on all ranks).
    Returns [len(queries_of(rank)), 2, k, N]: whole ciphertexts of the queried columns this rank rescales."""
    nq = len(query_idx)
    mine = queries_of(nq, rank, world)
    idx = torch.as_tensor([int(query_idx[i]) for d in range(world) for i in queries_of(nq, d, world)],
                          dtype=torch.long, device=enc_slice.device)
    send = enc_slice.index_select(0, idx).contiguous()
    send_splits = [len(queries_of(nq, d, world)) for d in range(world)]
    recv = torch.empty((world * len(mine),) + tuple(enc_slice.shape[1:]), dtype=enc_slice.dtype, device=enc_slice.device)
    if world == 1:
        recv.copy_(send)
    else:
        dist.all_to_all_single(recv, send, [len(mine)] * world, send_splits, group=group)
    _, two, k, ns = enc_slice.shape
    # recv is [source rank][my query][poly][limb][slice]; coefficient index = source * ns + c
    return recv.view(world, len(mine), two, k, ns).permute(1, 2, 3, 0, 4).reshape(len(mine), two, k, world * ns).contiguous()


def exchange_encoded_torch(enc_slice, rank, world, group=None, chunks=8):
    """enc_slice: [n, 2, k, N/world], this rank's coefficient slice of the whole encoded matrix.
    Returns [n/world, 2, k, N]: whole ciphertexts of the encoded columns column_range(n, rank, world), which this rank
    rescales to leaves, hashes and later serves queries from.  Runs as `chunks` all-to-alls to bound the staging buffers."""
    n, two, k, ns = enc_slice.shape
    lo, hi = column_range(n, rank, world)
    per = hi - lo
    out = torch.empty((per, two, k, world * ns), dtype=enc_slice.dtype, device=enc_slice.device)
    if world == 1:
        out.copy_(enc_slice)
        return out
    chunks = max(1, min(chunks, per))
    while per % chunks:
        chunks -= 1
    sub = per // chunks
    ov = out.view(per, two, k, world, ns)
    for c in range(chunks):
        # to rank d: its columns d*per + [c*sub, (c+1)*sub) of my slice; from rank s: slice s of my columns [c*sub, (c+1)*sub)
        send = torch.stack([enc_slice[d * per + c * sub: d * per + (c + 1) * sub] for d in range(world)])
        recv = torch.empty_like(send)
        dist.all_to_all_single(recv, send, group=group)
        ov[c * sub:(c + 1) * sub].copy_(recv.permute(1, 2, 3, 0, 4))
    return out


def commit_tree_torch(backend, leaves, leaf_format, rank, world, group=None):
    """Commit's Merkle span on column-sharded leaves (fhe/ligero.go:105-116): every rank hashes its own leaves on its GPU, the
    32-byte hashes are all-gathered in column order and every rank builds the same tree (root, paths) from them."""
    from . import fhe
    mine = fhe.LeafHashes(backend, leaves, leaf_format)
    if world == 1:
        return fhe.MerkleTree(backend, None, None, leafHashes=mine)
    dev = torch.device("cuda", backend.device) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    local = torch.frombuffer(bytearray(mine), dtype=torch.uint8).to(dev)
    gathered = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(gathered, local, group=group)
    return fhe.MerkleTree(backend, None, None, leafHashes=b"".join(bytes(t.cpu().numpy().tobytes()) for t in gathered))


def marshal_columns(backend, store, leafFormat, first=0, count=None):
    """Wire format on a column shard (EncryptedProof.WriteTo, fhe/ligero.go:665-679): the ct.WriteTo streams of this rank's
    ciphertexts.  Every stream has the same length L, so in the proof buffer rank r's MatR block starts at byte
    11 + column_range(cols, r, world)[0] * L (MatZ: + cols * L); the host concatenates / writes the blocks at those offsets."""
    import ctypes as C
    from . import _capi
    count = store.count - first if count is None else count
    f = leafFormat.c_struct()
    n = C.c_size_t()
    _capi.check(_capi.lib().lmn_store_marshal(backend.ctx, store.h, first, count, C.byref(f), None, 0, C.byref(n)))
    buf = np.empty(n.value, dtype=np.uint8)
    if n.value:
        _capi.check(_capi.lib().lmn_store_marshal(backend.ctx, store.h, first, count, C.byref(f), C.c_void_p(buf.ctypes.data), buf.nbytes, C.byref(n)))
    return buf.tobytes()
