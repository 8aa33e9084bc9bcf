"""Host time of one e2e train step with the input lookahead: wall per step, and time from the end of fetch to the
enqueue of the next step's first kernel (GPU idle gap)."""
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
K = 300
it = iter([host[i % 8] for i in range(3 * K + 40)])
for _ in range(20):
    kb.KGEModel.train_step(model, opt, it, targs)
marks = {"fetch_end": 0.0, "gap": 0.0, "step_call": 0.0, "fetch": 0.0, "stage": 0.0}
real_step, real_fetch, real_stage = lib.kge_train_step, lib.kge_fetch_floats, lib.kge_stage_batch
def step(*a):
    t0 = time.perf_counter()
    if marks["fetch_end"]:
        marks["gap"] += t0 - marks["fetch_end"]
    r = real_step(*a); marks["step_call"] += time.perf_counter() - t0; return r
def fetch(*a):
    t0 = time.perf_counter(); r = real_fetch(*a); t1 = time.perf_counter(); marks["fetch"] += t1 - t0; marks["fetch_end"] = t1; return r
def stage(*a):
    t0 = time.perf_counter(); r = real_stage(*a); marks["stage"] += time.perf_counter() - t0; return r
lib.kge_train_step, lib.kge_fetch_floats, lib.kge_stage_batch = step, fetch, stage
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(K):
    kb.KGEModel.train_step(model, opt, it, targs)
tot = time.perf_counter() - t0
print("per step: wall %.1f us | python gap fetch-end -> next step call %.1f us | step C call %.1f us | stage C call %.1f us | fetch (waits for the GPU) %.1f us"
      % (tot / K * 1e6, marks["gap"] / K * 1e6, marks["step_call"] / K * 1e6, marks["stage"] / K * 1e6, marks["fetch"] / K * 1e6))
lib.kge_train_step, lib.kge_fetch_floats, lib.kge_stage_batch = real_step, real_fetch, real_stage
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(K):
    kb.KGEModel.train_step(model, opt, it, targs)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
