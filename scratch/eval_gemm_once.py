"""One filtered ranking call per mode of the eval_gemm bench section (DistMult, FB15k sizes): the command profiled for the
per-launch times of eval_mma_kernel (profiles/r2_eval_mma_pair.csv)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import customknowledgegraphembedding_b200 as kb

name = sys.argv[1] if len(sys.argv) > 1 else "DistMult"
de, dr = (True, True) if name == "ComplEx" else (False, False)
E, R, H, nq = 14951, 1345, 1000, 4096
torch.manual_seed(1)
m = kb.KGEModel(name, E, R, H, 12.0, de, dr).cuda()
rng = np.random.RandomState(11)
q = torch.from_numpy(np.stack([rng.randint(0, E, nq), rng.randint(0, R, nq), rng.randint(0, E, nq)], 1)).cuda()
lens = rng.randint(0, 6, nq)
indptr = np.zeros(nq + 1, dtype=np.int64)
np.cumsum(lens, out=indptr[1:])
idx = torch.from_numpy(rng.randint(0, E, int(indptr[-1])).astype(np.int64)).cuda()
indptr = torch.from_numpy(indptr).cuda()
for _ in range(2):
    for mode in ("head-batch", "tail-batch"):
        out = m.rank_queries(q, mode, indptr, idx)
torch.cuda.synchronize()
print("checksum", int(out[1].sum()))
