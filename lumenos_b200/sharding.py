"""Multi-GPU partitioning of the hot path (one process per GPU, torch.distributed for the plumbing).

SURVEY 8(e): the inner products, the query rescale and the leaf rescale are independent per column ciphertext
(fhe/ligero.go:318-345 has no cross-ciphertext reduction), so columns shard by rank with no data-path collective.
Encode is independent per (poly, limb, coefficient) lane (fhe/ntt.go applies only ct+ct, ct-ct, ct*scalar), so it shards
by coefficient slice.  The single real exchange is turning the coefficient slices of the few hundred QUERIED encoded
columns into whole ciphertexts on the rank that rescales them: one all-to-all (NCCL over NVLink on the GPU box,
gloo in the CPU tests)."""
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


def exchange_queried(enc_slice, query_idx, rank, world, group=None):
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
