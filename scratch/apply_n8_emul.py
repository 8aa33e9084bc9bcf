"""Stage times of the owner-parallel apply of ONE rank of an 8-GPU data-parallel step, emulated on one GPU: the sweep
runs on the whole global batch (8 x 1024 positives), the apply on the 1/8 of the entity rows the rank would own."""
import argparse, ctypes as C, sys
import torch
import customknowledgegraphembedding_b200 as kb
from customknowledgegraphembedding_b200 import _lib, dist as kd
import bench

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
w = bench.WORKLOAD
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = kb.KGEModel(w["model"], w["nentity"], w["nrelation"], w["hidden_dim"], w["gamma"], True, False).to(dev)
opt = torch.optim.Adam(filter(lambda p: p.requires_grad, model.parameters()), lr=w["lr"])
targs = argparse.Namespace(cuda=True, negative_adversarial_sampling=True, adversarial_temperature=1.0, uni_weight=False,
                           regularization=0.0)
dp = kd.DataParallelStep(model, opt)
dp.lo, dp.hi = kd.shard_range(w["nentity"], 0, G)
batches = []
for s in range(4):
    parts = [bench.synthetic_batch(w["nentity"], w["nrelation"], w["batch"], w["negatives"], seed=100 * s + r) for r in range(G)]
    mode = parts[0][3]
    batches.append((torch.cat([p[0] for p in parts]).to(dev), torch.cat([p[1] for p in parts]).to(dev),
                    torch.cat([p[2] for p in parts]).to(dev), mode))
lib = _lib.load()
for i in range(6):
    dp.train_step(iter([batches[i % 4]]), targs, sync=False)
torch.cuda.synchronize()
lib.kge_profile_enable(1)
K = 20
for i in range(K):
    dp.train_step(iter([batches[i % 4]]), targs, sync=False)
torch.cuda.synchronize()
nst = lib.kge_profile_num_stages()
ms = (C.c_float * nst)(); ln = (C.c_int64 * nst)()
lib.kge_profile_read(ms, ln, 1)
print("emulated rank 0 of %d:" % G, {lib.kge_profile_stage_name(i).decode(): round(ms[i] / K, 4) for i in range(nst) if ms[i] > 0})
