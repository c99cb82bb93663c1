"""debug: one InnerSum with few rotations under the current LMN_MAC_VAR; dumps outputs to argv[1]"""
import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, "/root/repo")
from oracle import oracle as O
from lumenos_b200 import _capi, fhe
logN, k, cols, rows = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
P = O.Params(logN, k=k)
N = P.N
params = fhe.BGVParameters(P.logN, P.q, P.p, P.t, P.psi_q, P.psi_p)
be = fhe.ServerBFV(fhe.PrimeField(P.t, O.field_roots_forward(128)), params, stream=torch.cuda.current_stream().cuda_stream)
lib = _capi.lib()
rng = np.random.default_rng(11)
mods = list(P.q) + list(P.p)
for g in P.galois_elements_for_innersum(rows):
    e = np.empty((P.beta, 2, k + P.np, N), np.uint64)
    for i, m in enumerate(mods):
        e[:, :, i, :] = rng.integers(0, int(m), size=(P.beta, 2, N), dtype=np.uint64)
    be.LoadGaloisKey(g, e)
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(5)
M = torch.empty((cols, 2, k, N), dtype=torch.int64, device=dev)
for i, m in enumerate(P.q):
    M[:, :, i, :] = torch.randint(0, int(m), (cols, 2, N), generator=g, dtype=torch.int64, device=dev)
sM = fhe.CiphertextStore(be, cols, k - 1, wrap_ptr=M.data_ptr())
rpt = np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in P.q])
bpt = np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in P.q])
vp = C.c_void_p
res = []
for rep in range(3):
    mr, mz = be.NewStore(cols, min(1, k - 1)), be.NewStore(cols, min(1, k - 1))
    _capi.check(lib.lmn_inner_product2(be.ctx, sM.h, rpt.ctypes.data_as(vp), bpt.ctypes.data_as(vp), rows, mr.h, mz.h))
    res.append(np.stack([mr.download(), mz.download()]))
np.save(sys.argv[1], np.stack(res))
print("saved", res[0].shape, "rep0==rep1", np.array_equal(res[0], res[1]), "rep0==rep2", np.array_equal(res[0], res[2]))
