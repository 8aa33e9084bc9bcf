"""Where the host time of one e2e train step goes (pinned host batch, public API)."""
import argparse, time
import torch
import customknowledgegraphembedding_b200 as kb
from customknowledgegraphembedding_b200 import _lib
import bench

w = bench.WORKLOAD
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = kb.KGEModel(w["model"], w["nentity"], w["nrelation"], w["hidden_dim"], w["gamma"], True, False).to(dev)
opt = torch.optim.Adam(filter(lambda p: p.requires_grad, model.parameters()), lr=w["lr"])
targs = argparse.Namespace(cuda=True, negative_adversarial_sampling=True, adversarial_temperature=1.0, uni_weight=False,
                           regularization=0.0)
host = []
for s in range(8):
    pos, neg, sw, mode = bench.synthetic_batch(w["nentity"], w["nrelation"], w["batch"], w["negatives"], seed=s)
    host.append((pos.pin_memory(), neg.pin_memory(), sw.pin_memory(), mode))
lib = _lib.load()
real = lib.kge_train_step_host
acc = {"c": 0.0, "n": 0}
class Wrap:
    def __call__(self, *a):
        t0 = time.perf_counter(); r = real(*a); acc["c"] += time.perf_counter() - t0; acc["n"] += 1; return r
K = 300
it = iter([host[i % 8] for i in range(K + 20)])
for _ in range(20):
    kb.KGEModel.train_step(model, opt, it, targs)
lib.kge_train_step_host = Wrap()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(K):
    kb.KGEModel.train_step(model, opt, it, targs)
tot = time.perf_counter() - t0
print("per step: total %.1f us, inside the C call (copies + launches + sync) %.1f us, python around it %.1f us"
      % (tot / K * 1e6, acc["c"] / K * 1e6, (tot - acc["c"]) / K * 1e6))
import cProfile, pstats
lib.kge_train_step_host = real
it = iter([host[i % 8] for i in range(K)])
pr = cProfile.Profile(); pr.enable()
for _ in range(K):
    kb.KGEModel.train_step(model, opt, it, targs)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
