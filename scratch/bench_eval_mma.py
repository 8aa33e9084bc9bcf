"""Eval throughput of config 4 (ComplEx / DistMult, FB15k sizes, H=1000): tensor path vs fp32 CUDA-core path."""
import sys, os, json, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import customknowledgegraphembedding_b200 as kb
from customknowledgegraphembedding_b200 import _lib
lib = _lib.load()
E, R, H = 14951, 1345, 1000
Qn = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
out = {}
for name, de, dr in (("DistMult", False, False), ("ComplEx", True, True)):
    torch.manual_seed(0)
    m = kb.KGEModel(name, E, R, H, 12.0, de, dr).cuda()
    rng = np.random.RandomState(0)
    q = torch.from_numpy(np.stack([rng.randint(0, E, Qn), rng.randint(0, R, Qn), rng.randint(0, E, Qn)], 1)).cuda()
    lens = rng.randint(0, 6, Qn); indptr = np.zeros(Qn + 1, dtype=np.int64); np.cumsum(lens, out=indptr[1:])
    idx = torch.from_numpy(rng.randint(0, E, int(indptr[-1])).astype(np.int64)).cuda(); indptr = torch.from_numpy(indptr).cuda()
    for path in (1, 0):
        lib.kge_eval_set_tensor_path(path)
        for _ in range(2): m.rank_queries(q, "tail-batch", indptr, idx)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for i in range(reps): r = m.rank_queries(q, "head-batch" if i % 2 else "tail-batch", indptr, idx)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        K = m.entity_dim
        out["%s/%s" % (name, "tcgen05" if path else "fp32-simt")] = {"ms": ms, "queries_per_s": Qn / ms * 1e3,
            "useful_tflops": 2.0 * Qn * E * K / ms / 1e9, "mma_tflops_3x": 6.0 * Qn * E * K / ms / 1e9 if path else None}
    lib.kge_eval_set_tensor_path(1)
print(json.dumps(out, indent=1))
