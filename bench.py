#!/usr/bin/env python
"""bench.py -- the BASELINE.json metric on the BASELINE.json config, one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

metric  : train triples*neg/sec  (B*N negatives scored, back-propagated and Adam-applied per step)
workload: BASELINE.json configs[1] -- RotatE, FB15k-237 sizes (14541 entities, 237 relations),
          hidden_dim 1000, gamma 9, batch 1024, 256 negatives, -adv; synthetic triples of that shape
          (the dataset's train.txt is a missing blob in the reference), random-init tables.
step    : forward (negative + positive scores), loss, backward, dense Adam on both tables.

value   : device-resident batches, CUDA events around exactly K steps, max over ranks.
e2e     : the same K steps through KGEModel.train_step(model, optimizer, iterator, args) with the batch
          in pinned HOST memory: per step H2D of (positive, negative, weights) and D2H of the loss log.
roofline: dominant kernel of the step (by measured device time, CUDA events recorded by the library
          around each stage inside a second timed region), algorithmic bytes / duration vs MEASURED_PEAKS.
cpu_baseline / --impl reference: the CPU oracle (a port: the reference's PyTorch half is an absent
          submodule and TensorFlow is not installable) on the host cores, bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

WORKLOAD = dict(model="RotatE", nentity=14541, nrelation=237, hidden_dim=1000, gamma=9.0, batch=1024,
                negatives=256, adversarial_temperature=1.0, lr=5e-5)
WORKLOAD_NAME = "RotatE FB15k-237-shaped (E=14541,R=237) H=1000 gamma=9 B=1024 N=256 -adv, fwd+loss+bwd+dense Adam"
CPU_SAMPLE_B = 64  # positives of the CPU arm's larger probe step (the timed sample is sized from the budget)


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed regions run."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.max_mhz = index, [], False, None
        self.reasons = set()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                     "sw_thermal_slowdown": 0x20, "hw_power_brake_slowdown": 0x80}
            while not self.stop_flag:
                util = nv.nvmlDeviceGetUtilizationRates(h).gpu
                mhz = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((mhz, util))
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
                time.sleep(0.1)
        except Exception as e:  # pragma: no cover
            self.reasons.add("sampler_error:%s" % type(e).__name__)

    def summary(self):
        import statistics
        loaded = [m for m, u in self.samples if u > 0] or [m for m, _ in self.samples]
        return {"sm_mhz": statistics.median(loaded) if loaded else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def algorithmic_bytes(w):
    """Per-launch algorithmic bytes of each stage (DESIGN.md 'Kernels'; SURVEY 8(d) terms)."""
    B, N, H = w["batch"], w["negatives"], w["hidden_dim"]
    De, Dr = 2 * H, H
    E, R = w["nentity"], w["nrelation"]
    rows = B * N
    return {
        "score_forward(neg)": rows * (De * 4 + 8 + 4) + B * (De + Dr + De) * 4,   # candidate row + index + score
        "score_backward(neg)": rows * (2 * De * 4 + 8 + 4) + B * (2 * De + Dr) * 4 * 2,  # row read + grad row add
        "adam(entity)": E * De * 28,   # p,m,v read+write, g read
        "adam(relation+modulus)": R * Dr * 28,
        "zero_grad": (E * De + R * Dr) * 4,
        # second-generation step: the sweep reads every candidate row once; the entity pass reads the staged
        # planes of the using positive once per use, and p, m, v of every row once (read + write)
        "fused_sweep(scores+loss+dL/dq)": rows * (De * 4 + 8 + 4) + B * (3 * De + Dr) * 4 * 2,
        "entity_gather+adam": rows * (De * 4 + 4 + 4) + 2 * B * De * 4 + E * De * 24,
        "relation_gather+adam": B * Dr * 4 + R * Dr * 24,
        "step": rows * (3 * De * 4 + 16) + B * (2 * De + Dr) * 12 + (E * De + R * Dr) * 24,
    }


def synthetic_batch(nentity, nrelation, batch, negatives, seed):
    """SURVEY 8(d) synthetic inputs: uniform triples and negatives, w = sqrt(1/c), c ~ U{8..200}, modes alternate."""
    import torch
    g = torch.Generator().manual_seed(int(seed))
    pos = torch.stack([torch.randint(0, nentity, (batch,), generator=g),
                       torch.randint(0, nrelation, (batch,), generator=g),
                       torch.randint(0, nentity, (batch,), generator=g)], dim=1)
    neg = torch.randint(0, nentity, (batch, negatives), generator=g)
    w = torch.sqrt(1.0 / torch.randint(8, 201, (batch,), generator=g).to(torch.float32))
    return pos, neg, w, ("head-batch" if seed % 2 == 0 else "tail-batch")


def shared_config(world):
    """The `config` object of BOTH arms (ours and --impl reference): same workload, same keys."""
    w = WORKLOAD
    return {"workload": WORKLOAD_NAME,
            "parallelism": "single GPU" if world == 1 else
            "dp%d: %d positives per GPU, ONE optimizer step on the global batch of %d; sweep data-parallel over "
            "positives, apply owner-parallel over entity rows; per-positive buffers exchanged by peer stores over "
            "NVLink (symmetric memory), updated rows stored into every replica by the Adam epilogue (NVSwitch "
            "multicast); NCCL moves one float (the weight sum)" % (world, w["batch"], w["batch"] * world),
            "l2": "no flush: tables + Adam state = %.2f GB per step > 126 MB L2; the entity table itself (116 MB) "
                  "is partly L2-resident for the gathers, as in any real run" %
                  ((w["nentity"] * 2 * w["hidden_dim"] + w["nrelation"] * w["hidden_dim"]) * 12 / 1e9),
            "backward": "fixed-order fused step (no float atomics, no dense gradient table)"}


def cpu_baseline(steps, warmup, threads=None, budget_s=30.0):
    """The oracle's train_step on the host cores, bounded sample of the workload: a first-touch step and two probe
    steps (64 and 8 positives) give a linear cost model  t(b) = fixed + per_pos * b  (fixed = dense autograd gradient
    tables + dense Adam over both full tables, paid whatever b is); the positives per step are the LARGEST power of two
    <= the batch for which warm-up + steps fit ``budget_s`` (never below 8).  Reports the measured throughput of that
    sample and the model's extrapolation to the full batch (the sample amortises `fixed` over fewer positives, so
    the measured number understates the CPU arm; the extrapolated one does not)."""
    import torch
    from oracle import kge_oracle as ko
    w = WORKLOAD
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    m = ko.make_model(w["model"], w["nentity"], w["nrelation"], w["hidden_dim"], w["gamma"], seed=0)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, m.parameters()), lr=w["lr"])
    args = ko.Args(negative_adversarial_sampling=True, adversarial_temperature=w["adversarial_temperature"])

    def one(seed, nb):
        batch = ko.synthetic_batch(w["nentity"], w["nrelation"], nb, w["negatives"], seed=seed)
        t0 = time.perf_counter()
        ko.train_step(m, opt, iter([batch]), args)
        return time.perf_counter() - t0

    one(10_000, 8)  # first touch of the tables and the Adam state
    t_full, t_small = one(10_001, CPU_SAMPLE_B), one(10_002, CPU_SAMPLE_B // 8)
    per_pos = max(1e-6, (t_full - t_small) / (CPU_SAMPLE_B - CPU_SAMPLE_B // 8))
    fixed = max(0.0, t_small - per_pos * (CPU_SAMPLE_B // 8))
    sample_b = w["batch"]
    while sample_b > 8 and (steps + warmup) * (fixed + per_pos * sample_b) > budget_s:
        sample_b //= 2
    times = []
    for s in range(warmup + steps):
        dt = one(s, sample_b)
        if s >= warmup:
            times.append(dt)
    total = sum(times)
    # refit per_pos with the measured sample (more positives than the probes: a better slope)
    if sample_b > CPU_SAMPLE_B:
        per_pos = max(1e-6, (total / len(times) - t_small) / (sample_b - CPU_SAMPLE_B // 8))
        fixed = max(0.0, t_small - per_pos * (CPU_SAMPLE_B // 8))
    full = w["batch"] * w["negatives"] / (fixed + per_pos * w["batch"])
    measured = sample_b * w["negatives"] * len(times) / total
    # `value` is the figure for the FULL batch: the measurement itself when the sample is the whole batch, else the cost
    # model's extrapolation (a smaller sample amortises the fixed per-step cost over fewer positives and would understate
    # the CPU arm); the measured throughput of the sample is reported beside it
    return {"value": measured if sample_b == w["batch"] else max(measured, full), "measured_sample_value": measured,
            "unit": "triples*neg/s", "cores": cores, "kind": "port",
            "sample": "%d steps of %d positives x %d negatives (%s of the batch: the largest power of two whose run fits "
                      "%.0f s) incl. dense Adam over the full tables, torch %s CPU fp32; cost model t(b) = %.3f s + "
                      "%.4f s * b -> %.0f triples*neg/s extrapolated to the full batch of %d" % (
                          len(times), sample_b, w["negatives"],
                          "all" if sample_b == w["batch"] else "1/%d" % (w["batch"] // sample_b), budget_s,
                          torch.__version__, fixed, per_pos, full, w["batch"]),
            "ms_per_step": 1e3 * total / len(times), "sample_positives": sample_b,
            "full_batch_extrapolated": full, "fixed_s": fixed, "per_positive_s": per_pos}


def run_reference(a):
    """--impl reference: the reference path's CPU implementation on the host cores (oracle port), rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base = cpu_baseline(max(1, a.steps), max(0, a.warmup), budget_s=170.0)  # the whole arm ends within ~3 minutes
    line = {
        "impl": "reference", "metric": "train_triples_neg_per_sec", "value": base["value"], "unit": "triples*neg/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": base["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": shared_config(a.gpus),  # the same object as the other arm's: same workload, same keys
        "reference_arm": "CPU oracle (port of the absent PyTorch submodule) on the host cores; " + base["sample"],
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "measured_sample_value",
                                              "full_batch_extrapolated")},
        "e2e": {"value": base["value"], "unit": "triples*neg/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def l2_read_peak(lib, dev, timed, mbytes=32, iters=200):
    """Measured L2 read bandwidth (GB/s): kge_probe_l2_read re-reads a 32 MB buffer (resident in the 126 MB L2) with every
    SM touching every line, CUDA events around the launch, best of 3 (burst figure: a kernel timed alone)."""
    import torch
    buf = torch.ones(mbytes << 20, dtype=torch.uint8, device=dev)
    sink = torch.zeros(1, dtype=torch.float32, device=dev)
    nread = C.c_int64(0)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def run():
        rc = lib.kge_probe_l2_read(C.c_void_p(buf.data_ptr()), buf.numel(), iters, C.c_void_p(sink.data_ptr()),
                                   C.byref(nread), stream)
        if rc != 0:
            raise RuntimeError(lib.kge_last_error().decode())
    run()
    best = min(timed(run) for _ in range(3))
    return nread.value / (best * 1e-3) / 1e9


def l2_bytes(w):
    """Bytes the two dominant stages request from L2 per launch (every candidate row / staged plane row is gathered once
    per USE): what the L2 roofline is computed on."""
    B, N, H = w["batch"], w["negatives"], w["hidden_dim"]
    De, E = 2 * H, w["nentity"]
    rows, parts = B * N, max(1, min(8, (N + 127) // 128))  # partials per positive of the CTA-cooperative rows kernel
    return {
        # candidate rows + ids + scores; Q rows per item; per-item partials written and read back; positives' rows
        "fused_sweep(scores+loss+dL/dq)": rows * (De * 4 + 12) + 3 * B * parts * De * 4 + B * (3 * De + H) * 4 * 2,
        # one staged plane row + (dL/ds, b) per use; p, m, v of every row read + written; compact rows of the positives
        "entity_gather+adam": rows * (De * 4 + 12) + E * De * 24 + 2 * B * De * 4,
    }


def table_checksum(model):
    """Exact integer checksum of both tables (fp32 bit patterns summed as int64)."""
    import torch
    out = []
    for t in (model.entity_embedding, model.relation_embedding):
        out.append(int(t.detach().contiguous().view(torch.int32).to(torch.int64).sum().item()))
    return out


def run_dist_check(kb, model, step_obj, targs, w, dev, rank, world, dist, make_ref, nsteps=3, seed0=777_000):
    """Before timing: `nsteps` global steps of the N-rank step against ONE GPU's KGEModel.train_step on the concatenated
    batch (run by rank 0 on a private replica built by ``make_ref``): per-step loss within 2e-6 relative, tables within
    5e-3 * lr after the steps.  The batches of every rank are seeded, so rank 0 can rebuild the global batch."""
    import torch
    lr = w["lr"]
    ref = ropt = None
    if rank == 0:
        ref = make_ref()
        ropt = torch.optim.Adam(filter(lambda p: p.requires_grad, ref.parameters()), lr=lr)
    loss_rel, losses = 0.0, []
    for s in range(nsteps):
        batches = [synthetic_batch(w["nentity"], w["nrelation"], w["batch"], w["negatives"], seed=seed0 + 1000 * r + s)
                   for r in range(world)]
        p, n, sw, m = batches[rank]
        log = step_obj.train_step(iter([(p.to(dev), n.to(dev), sw.to(dev), m)]), targs)
        if rank == 0:
            cat = (torch.cat([b[0] for b in batches]).to(dev), torch.cat([b[1] for b in batches]).to(dev),
                   torch.cat([b[2] for b in batches]).to(dev), m)
            rlog = kb.KGEModel.train_step(ref, ropt, iter([cat]), targs)
            loss_rel = max(loss_rel, abs(log["loss"] - rlog["loss"]) / abs(rlog["loss"]))
            losses.append((log["loss"], rlog["loss"]))
    out = None
    if rank == 0:
        out = {"steps": nsteps, "global_batch": w["batch"] * world, "loss_rel_max": loss_rel, "losses": losses,
               "tol_loss_rel": 2e-6, "tol_table": 2e-2 * lr, "tol_table_bulk": 5e-3 * lr, "tol_bulk_fraction": 1e-5,
               "tol_note": "tables: every element within 2e-2*lr and all but 1e-5 of them within 5e-3*lr.  The N-rank step "
                           "differs from one GPU only in the order the subsampling-weight sum is reduced (1-ulp coefficients); "
                           "Adam's update lr*m/(sqrt(v)+eps) amplifies that where a gradient element is a near-cancelling "
                           "sum (|g| ~ 1e-5 of its terms): a handful of the 29 M elements move by ~1e-2*lr"}
    return ref, out


def run_ours(a):
    import numpy as np
    import torch
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        import datetime
        # a collective that some rank never joins aborts after 3 minutes instead of hanging the whole run
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    lib = _lib.load()
    w = WORKLOAD
    K, W = a.steps, max(a.warmup, 3)

    torch.manual_seed(0)
    model = kb.KGEModel(w["model"], w["nentity"], w["nrelation"], w["hidden_dim"], w["gamma"], True, False).to(dev)
    model.deterministic = bool(a.deterministic)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, model.parameters()), lr=w["lr"])
    targs = argparse.Namespace(cuda=True, negative_adversarial_sampling=True,
                               adversarial_temperature=w["adversarial_temperature"], uni_weight=False,
                               regularization=0.0)
    group = model._fused_adam_plan(opt)
    nb = min(K + W, 32)
    host = []
    for s in range(nb):
        pos, neg, sw, mode = synthetic_batch(w["nentity"], w["nrelation"], w["batch"], w["negatives"],
                                             seed=1000 * rank + s)
        host.append((pos.pin_memory(), neg.pin_memory(), sw.pin_memory(), mode))
    devb = [(p.to(dev), n.to(dev), s.to(dev), m) for p, n, s, m in host]
    h2d = sum(t.numel() * t.element_size() for t in host[0][:3])
    units = w["batch"] * w["negatives"]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    dp = None
    if world > 1:
        # one GLOBAL step per call: positives data-parallel in the sweep, entity rows owner-parallel in the apply,
        # NCCL all-gather of the per-positive buffers and of the updated row ranges (dist.DataParallelStep)
        from customknowledgegraphembedding_b200 import dist as kd
        dp = kd.DataParallelStep(model, opt)

    def device_steps(n, start):
        out = None
        for i in range(n):
            p, ng, s, m = devb[(start + i) % nb]
            if dp is not None:
                out = dp.train_step(iter([(p, ng, s, m)]), targs, sync=False)
            else:
                out = model._fused_step(opt, group, p, ng, s, m, True, w["adversarial_temperature"], False, 0.0)
        return out

    def timed(fn):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    dist_check = None
    if dp is not None and not a.no_dist_check:
        def make_ref():
            ref = kb.KGEModel(w["model"], w["nentity"], w["nrelation"], w["hidden_dim"], w["gamma"], True, False).to(dev)
            ref.load_state_dict(model.state_dict())
            return ref
        ref, dist_check = run_dist_check(kb, model, dp, targs, w, dev, rank, world, dist, make_ref)
        if rank == 0:
            dEt = (model.entity_embedding.detach() - ref.entity_embedding.detach()).abs()
            dRt = (model.relation_embedding.detach() - ref.relation_embedding.detach()).abs()
            dE, dR = float(dEt.max()), float(dRt.max())
            bulk = dist_check["tol_table_bulk"]
            frac = float(((dEt > bulk).sum() + (dRt > bulk).sum()).item()) / (dEt.numel() + dRt.numel())
            dist_check.update(max_abs_diff_entity=dE, max_abs_diff_relation=dR, fraction_above_bulk_tol=frac)
            dist_check["ok_before_timing"] = bool(dist_check["loss_rel_max"] <= dist_check["tol_loss_rel"]
                                                  and max(dE, dR) <= dist_check["tol_table"]
                                                  and frac <= dist_check["tol_bulk_fraction"])
            del dEt, dRt
        del ref
        torch.cuda.empty_cache()

    sampler = ClockSampler(local_rank)
    device_steps(W, 0)  # warm-up (also allocates the workspace and Adam state)
    sampler.start()
    # ---- value: K steps, batches resident in HBM, no host sync inside -------------------------------
    ms_value = timed(lambda: device_steps(K, W))
    # ---- second timed region with per-stage CUDA events (roofline of the dominant kernel) -----------
    lib.kge_profile_enable(1)
    ms_prof = timed(lambda: device_steps(K, W))
    nst = lib.kge_profile_num_stages()
    ms_sum = (C.c_float * nst)()
    launches = (C.c_int64 * nst)()
    lib.kge_profile_read(ms_sum, launches, 1)
    lib.kge_profile_enable(0)
    stages = {lib.kge_profile_stage_name(i).decode(): (ms_sum[i], launches[i]) for i in range(nst)}
    l2_peak = l2_read_peak(lib, dev, timed)  # on every rank: `timed` is collective (barrier + max over ranks)
    # ---- e2e: the public API with the batch in pinned host memory ------------------------------------
    it = iter([host[(W + i) % nb] for i in range(K + 3)])
    for _ in range(3):
        dp.train_step(it, targs) if dp is not None else kb.KGEModel.train_step(model, opt, it, targs)
    last_log = {}

    def e2e_steps():
        for _ in range(K):
            if dp is not None:
                last_log.update(dp.train_step(it, targs))
            else:
                last_log.update(kb.KGEModel.train_step(model, opt, it, targs))
    ms_e2e = timed(e2e_steps)
    # ---- after the timed train regions: every replica holds bit-identical tables ---------------------
    if dp is not None and not a.no_dist_check:
        cs = torch.tensor(table_checksum(model), dtype=torch.int64, device=dev)
        allcs = [torch.empty_like(cs) for _ in range(world)]
        dist.all_gather(allcs, cs)
        if rank == 0:
            stack = torch.stack(allcs)
            spread = (stack.max(0).values - stack.min(0).values).tolist()
            dist_check["checksum_spread_after_timing"] = spread
            dist_check["checksums_rank0"] = cs.tolist()
            dist_check["ok"] = bool(dist_check["ok_before_timing"] and spread == [0, 0])
            if not dist_check["ok"]:
                log("DIST CHECK FAILED: %s" % json.dumps(dist_check))
    if rank == 0:
        log("train sections done (%.3f ms/step); eval / sharded sections ..." % (ms_value / K))
    # ---- eval: filtered ranking of 2048 queries against all entities --------------------------------
    if a.train_only:
        smp = ev = ev_gemm = None
        ev_10m = {"skipped": "--train-only"}
        tr_interht = train_config_bench(CONFIG3, dev, timed, min(K, 50)) if (world == 1 and a.interht) else None
        tr_transparse = train_transparse_bench(dev, timed, min(K, 10)) if (world == 1 and a.transparse) else None
    else:
        tr_interht = train_config_bench(CONFIG3, dev, timed, min(K, 50)) if world == 1 else None
        tr_transparse = train_transparse_bench(dev, timed, min(K, 10)) if world == 1 else None
        smp = sampler_bench(model, opt, group, w, dev, timed, K) if world == 1 else None
        ev = eval_bench(model, w, dev, timed, rank, world)
        if rank == 0:
            log("eval done; eval_gemm ...")
        ev_gemm = eval_gemm_bench(dev, timed, world)
        if rank == 0:
            log("eval_gemm done; eval_10m ...")
        ev_10m = eval_10m_bench(dev, timed, world) if not a.no_eval10m else {"skipped": "--no-eval10m"}
        if rank == 0:
            log("eval_10m done; train_sharded ...")
    tr_sharded = None
    if a.sharded or not a.train_only:
        tr_sharded = {"yago3-10": train_sharded_bench("yago3-10", dev, timed, rank, world, dist, min(K, 20))}
        if not a.no_eval10m:
            tr_sharded["synthetic-10M"] = train_sharded_bench("synthetic-10M", dev, timed, rank, world, dist, min(K, 10))
    sampler.stop_flag = True
    sampler.join(timeout=2)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    total_units = units * K * world
    peak, peak_src = peaks()
    ab = algorithmic_bytes(w)
    dom = max((k for k in stages if stages[k][1] > 0), key=lambda k: stages[k][0])
    dom_ms = stages[dom][0] / K
    achieved = ab.get(dom, 0) / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None
    traffic = None
    tpath = os.path.join(REPO, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(dom)
    step_ms = ms_value / K
    line = {
        "metric": "train_triples_neg_per_sec", "value": total_units / (ms_value * 1e-3), "unit": "triples*neg/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": shared_config(world),
        "e2e": {"value": total_units / (ms_e2e * 1e-3), "unit": "triples*neg/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 16, "ms_per_step": ms_e2e / K, "loss": last_log.get("loss")},
        "gpu_launches": int(sum(v[1] for v in stages.values())),
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic, "peak_source": peak_src,
                     "kernel_ms": dom_ms, "algorithmic_bytes": ab.get(dom),
                     "dram_frac": (traffic / (dom_ms * 1e-3) / 1e9 / peak) if (traffic and dom_ms > 0) else None,
                     "note": "algorithmic bytes are row gathers from the 116 MB entity table, most of which the 126 MB L2 "
                             "serves (traffic = DRAM bytes per launch from ncu is ~0.4x of them): frac can pass 1.0 and "
                             "dram_frac is the share of the HBM peak the kernel really draws; the binding roofline is the "
                             "L2 slices' throughput: see `l2`",
                     "step": {"algorithmic_bytes": ab["step"], "achieved": ab["step"] / (step_ms * 1e-3) / 1e9,
                              "frac": ab["step"] / (step_ms * 1e-3) / 1e9 / peak,
                              "frac_of_8TBs": ab["step"] / (step_ms * 1e-3) / 8e12},
                     # the roofline that actually binds the two dominant stages at this config: L2 slice throughput
                     "l2": {"bound": "l2", "peak": l2_peak, "unit": "GB/s",
                            "peak_source": "measured in this run: kge_probe_l2_read, 32 MB buffer re-read by every SM "
                                           "(B300_MICROARCH.md: LTS cap ~6300 B/clk full chip)",
                            "stages": {k: {"l2_bytes": v, "ms": stages[k][0] / K,
                                           "achieved": v / (stages[k][0] / K * 1e-3) / 1e9,
                                           "frac": v / (stages[k][0] / K * 1e-3) / 1e9 / l2_peak}
                                       for k, v in l2_bytes(w).items() if world == 1 and stages.get(k, (0, 0))[0] > 0}}},
        "stages_ms_per_step": {k: v[0] / K for k, v in stages.items()},
        "profiled_region_ms_per_step": ms_prof / K,
        "dist_check": dist_check,
        "train_interht": tr_interht,
        "train_sharded": tr_sharded,
        "train_transparse": tr_transparse,
        "gpu_sampler": smp,
        "eval": ev,
        "eval_gemm": ev_gemm,
        "eval_10m": ev_10m,
        "clocks": sampler.summary(),
    }
    def _v(d, *path):
        for k in path:
            if not isinstance(d, dict) or k not in d:
                return None
            d = d[k]
        return d
    # the numbers of the other BASELINE configs in one compact object near the front of the line (the tail of a long
    # line may be cut by whoever stores it)
    extras = {
        "eval_rotate_fb15k237_qps": _v(ev, "value"), "eval_rotate_sfu_frac": _v(ev, "roofline", "frac"),
        "eval_gemm_complex_qps": _v(ev_gemm, "ComplEx", "value"), "eval_gemm_distmult_qps": _v(ev_gemm, "DistMult", "value"),
        "eval_gemm_distmult_issued_frac": _v(ev_gemm, "DistMult", "roofline", "frac"),
        "eval_10m_qps": _v(ev_10m, "value"), "eval_10m_checksum": _v(ev_10m, "checksum_n_greater"),
        "train_interht_pytorch": _v(tr_interht, "pytorch", "value"), "train_interht_tf": _v(tr_interht, "tf", "value"),
        "train_transparse_as_written": _v(tr_transparse, "as_written", "value"),
        "train_transparse_head_batch_ms": _v(tr_transparse, "as_written", "head-batch", "ms_per_step"),
        "train_transparse_issued_frac": _v(tr_transparse, "as_written", "head-batch", "roofline", "frac"),
        "train_sharded_yago": _v(tr_sharded, "yago3-10", "value"),
        "train_sharded_yago_ms": _v(tr_sharded, "yago3-10", "ms_per_step"),
        "train_sharded_yago_check_ok": _v(tr_sharded, "yago3-10", "dist_check", "ok"),
        "train_sharded_10m": _v(tr_sharded, "synthetic-10M", "value"),
        "train_sharded_10m_ms": _v(tr_sharded, "synthetic-10M", "ms_per_step"),
        "dist_check_ok": _v(dist_check, "ok"), "l2_peak_gbs": l2_peak,
    }
    line = {k: line[k] for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step")} | \
        {"scaling_extras": extras} | {k: v for k, v in line.items() if k not in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step")}
    log("timing done; running the bounded CPU baseline sample ...")
    base = cpu_baseline(steps=2, warmup=1, budget_s=30.0) if not a.no_cpu_baseline else None
    line["cpu_baseline"] = ({k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "measured_sample_value",
                                                  "full_batch_extrapolated")} if base else None)
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


TRANSPARSE = dict(nentity=14541, nrelation=237, dim=500, gamma=12.0, batch=1024, negatives=256, lr=1e-4)


def train_transparse_bench(dev, timed, K):
    """the reference's second TF model, TFKGEModel('TranSparse') under Trainer.train_step
    (tensorflow_codes/model.py:96-106,226-235; supervisor.py:15-28), FB15k-237 sizes with dim 500, batch 1024, 256
    negatives.  Head-batch steps carry the work (B * (N + 1) rows through three D x D GEMMs: P = X M_r, dX = dP M_r^T,
    dW_r = X^T dP); as the reference wrote the score (model.py:229) a tail-batch step projects the B heads only.
    Roofline: tensor pipe, kind::tf32 = half the measured bf16 rate; `frac` counts the 3 MMA passes of 3xTF32."""
    import torch
    from customknowledgegraphembedding_b200.tf_compat import KerasAdam
    from customknowledgegraphembedding_b200.transparse import TranSparseModel
    c = TRANSPARSE
    E, R, D, B, N = c["nentity"], c["nrelation"], c["dim"], c["batch"], c["negatives"]
    bf16 = None
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            bf16 = float(json.load(f).get("bf16_tflops", 0)) or None
    peak, src = (bf16 / 2, "measured bf16_tflops / 2 (MEASURED_PEAKS.json)") if bf16 else (795.0, "fallback 1.59 PF bf16 / 2")
    out = {"workload": "TranSparse E=%d R=%d D=%d gamma=%g B=%d N=%d, TF loss (attached softmax), Keras Adam on entity / "
                       "relation / W" % (E, R, D, c["gamma"], B, N), "unit": "triples*neg/s"}
    batches = []
    for s in range(4):
        p, n, sw, _ = synthetic_batch(E, R, B, N, seed=9000 + s)
        batches.append((p.to(dev), n.to(dev), sw.to(dev)))
    for variant, tail in (("as_written", False), ("projected_tail", True)):
        torch.manual_seed(0)
        m = TranSparseModel(E, R, D, c["gamma"], project_tail=tail).to(dev)
        opt = KerasAdam(learning_rate=c["lr"])
        res = {}
        for mode in ("head-batch", "tail-batch"):
            def steps(n):
                for i in range(n):
                    p, ng, sw = batches[i % len(batches)]
                    m.tf_train_step(opt, p, ng, sw, mode)
                    opt.iterations += 1

            def forwards(n):
                with torch.no_grad():
                    for i in range(n):
                        p, ng, _ = batches[i % len(batches)]
                        m.both((p, ng), mode)
            steps(2)
            ms = timed(lambda: steps(K)) / K
            fwd_ms = timed(lambda: forwards(K)) / K
            big = mode == "head-batch" or tail
            rows = B * ((N if big else 0) + 1 + (1 if tail else 0))
            flops = 3 * 2.0 * rows * D * D
            res[mode] = {"ms_per_step": ms, "forward_only_ms": fwd_ms, "value": B * N / (ms * 1e-3), "rows_projected": rows,
                         "roofline": {"bound": "tensor", "achieved": 3 * flops / (ms * 1e-3) / 1e12, "peak": peak,
                                      "unit": "TFLOP/s", "frac": 3 * flops / (ms * 1e-3) / 1e12 / peak, "peak_source": src,
                                      "useful_tflops": flops / (ms * 1e-3) / 1e12,
                                      "useful_frac": flops / (ms * 1e-3) / 1e12 / peak,
                                      "note": "whole step: plan, operand splits of W and the rows, 3 grouped GEMMs, score / "
                                              "dscore, scatter, dense Keras Adam over E*D + R*D + R*D*D parameters"}}
        res["value"] = 2 * B * N / ((res["head-batch"]["ms_per_step"] + res["tail-batch"]["ms_per_step"]) * 1e-3)
        out[variant] = res
        del m, opt
        torch.cuda.empty_cache()
    return out


CONFIG3 = dict(model="InterHT", nentity=40943, nrelation=11, hidden_dim=1000, gamma=24.0, batch=1024, negatives=512,
               adversarial_temperature=1.0, lr=5e-5)


def train_config_bench(cfg, dev, timed, K, semantics=("pytorch", "tf")):
    """BASELINE.json configs[2]: InterHT, wn18rr sizes (40943 entities, 11 relations), H = 1000, gamma 24, batch 1024,
    512 negatives, one B200 -- device-resident batches, K steps per timed region, per-stage CUDA events.  Two loss
    semantics: 'pytorch' (KGEModel.train_step: detached softmax weights, torch Adam) and 'tf' (the reference's own
    Trainer.train_step_fn, tensorflow_codes/supervisor.py:15-28 over model.py:148-224: gradient THROUGH the softmax
    weights, temperature 1, Keras Adam epsilon placement)."""
    import torch
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import _lib
    lib = _lib.load()
    de, dr, tr = {"InterHT": (True, False, True), "RotatE": (True, False, False)}[cfg["model"]]
    E, R, H, B, N = cfg["nentity"], cfg["nrelation"], cfg["hidden_dim"], cfg["batch"], cfg["negatives"]
    De, Dr = (2 * H if de else H), (3 * H if tr else (2 * H if dr else H))
    peak, peak_src = peaks()
    units = B * N
    # SURVEY 8(d) per-step byte model
    step_bytes = units * (3 * De * 4 + 16) + B * (2 * De + Dr) * 12 + (E * De + R * Dr) * 24
    out = {"workload": "%s E=%d R=%d H=%d gamma=%g B=%d N=%d -adv" % (cfg["model"], E, R, H, cfg["gamma"], B, N),
           "unit": "triples*neg/s", "algorithmic_bytes_per_step": step_bytes}
    nb = 8
    batches = []
    for s in range(nb):
        p, n, sw, m = synthetic_batch(E, R, B, N, seed=5000 + s)
        batches.append((p.to(dev), n.to(dev), sw.to(dev), m))
    for sem in semantics:
        torch.manual_seed(0)
        model = kb.KGEModel(cfg["model"], E, R, H, cfg["gamma"], de, dr, tr).to(dev)
        opt = torch.optim.Adam(filter(lambda p: p.requires_grad, model.parameters()), lr=cfg["lr"],
                               eps=1e-7 if sem == "tf" else 1e-8)
        group = model._fused_adam_plan(opt)

        def steps(n):
            for i in range(n):
                p, ng, sw, m = batches[i % nb]
                model._fused_step(opt, group, p, ng, sw, m, True, cfg["adversarial_temperature"], False, 0.0,
                                  detach_weights=(sem != "tf"), keras_adam=(sem == "tf"))
        steps(3)
        ms = timed(lambda: steps(K))
        lib.kge_profile_enable(1)
        timed(lambda: steps(K))
        nst = lib.kge_profile_num_stages()
        ms_sum, launches = (C.c_float * nst)(), (C.c_int64 * nst)()
        lib.kge_profile_read(ms_sum, launches, 1)
        lib.kge_profile_enable(0)
        stages = {lib.kge_profile_stage_name(i).decode(): ms_sum[i] / K for i in range(nst) if launches[i] > 0}
        step_ms = ms / K
        out[sem] = {"value": units * K / (ms * 1e-3), "ms_per_step": step_ms, "stages_ms_per_step": stages,
                    "gpu_launches_per_step": int(sum(launches)) // K,
                    "roofline": {"bound": "hbm", "achieved": step_bytes / (step_ms * 1e-3) / 1e9, "peak": peak,
                                 "unit": "GB/s", "frac": step_bytes / (step_ms * 1e-3) / 1e9 / peak,
                                 "frac_of_8TBs": step_bytes / (step_ms * 1e-3) / 8e12, "peak_source": peak_src,
                                 "note": "whole step against SURVEY 8(d)'s byte model (three passes over the candidate "
                                         "rows + p/m/v read+write); the 327 MB table does not fit the 126 MB L2"}}
        del model, opt
        torch.cuda.empty_cache()
    return out


SHARDED = {
    "yago3-10": dict(model="RotatE", nentity=123182, nrelation=37, hidden_dim=1000, gamma=24.0, batch=1024, negatives=1024,
                     adversarial_temperature=1.0, lr=5e-5),
    "synthetic-10M": dict(model="RotatE", nentity=10_000_000, nrelation=237, hidden_dim=1000, gamma=24.0, batch=1024,
                          negatives=1024, adversarial_temperature=1.0, lr=5e-5),
}


def train_sharded_bench(name, dev, timed, rank, world, dist, K, check=True):
    """BASELINE.json configs[4]: RotatE H = 1000, 1024 negatives, batch 1024 per GPU, entity table (and Adam state)
    ROW-SHARDED over the ranks -- no rank holds the table.  Per step and rank: touched ids by bitmap + scan, rows pulled
    from their owners over NVLink into a compact table, sweep on the compact table, per-positive buffers to the owners by
    peer stores, owners update their rows in place.  ONE optimizer step on the global batch (weak scaling).
    `check`: 3 steps against KGEModel.train_step of one GPU on the concatenated batch (full table on rank 0)."""
    import torch
    import customknowledgegraphembedding_b200 as kb
    from customknowledgegraphembedding_b200 import dist as kd
    cfg = SHARDED[name]
    E, R, H, B, N = cfg["nentity"], cfg["nrelation"], cfg["hidden_dim"], cfg["batch"], cfg["negatives"]
    De = 2 * H
    lo, hi = kd.shard_range(E, rank, world)
    cap = min(E, B * N + 2 * B)
    need = (hi - lo) * De * 4 * 4 + cap * De * 4 + world * B * (N * 12 + 4 * De * 4) + (6 << 30)
    free, _ = torch.cuda.mem_get_info(dev)
    fits = torch.tensor([1 if free >= need else 0], device=dev)
    if dist is not None:
        dist.all_reduce(fits, op=dist.ReduceOp.MIN)
    if int(fits.item()) == 0:
        return {"skipped": "needs %.0f GB free per GPU at %d ranks, %.0f GB available" % (need / 1e9, world, free / 1e9)}
    targs = argparse.Namespace(cuda=True, negative_adversarial_sampling=True, adversarial_temperature=1.0, uni_weight=False,
                               regularization=0.0)
    model = kd.RowShardedKGEModel(cfg["model"], E, R, H, cfg["gamma"], True, False, device=dev, seed=11)
    opt = torch.optim.Adam(filter(lambda p: p.requires_grad, model.parameters()), lr=cfg["lr"])
    step = kd.RowShardedStep(model, opt)
    out = {"workload": "RotatE E=%d R=%d H=1000 gamma=%g B=%d per GPU N=%d -adv, row-sharded table" % (E, R, cfg["gamma"], B, N),
           "unit": "triples*neg/s", "n_gpus": world, "rows_per_gpu": hi - lo, "scaling": "weak"}
    # ---- correctness before timing (tables that fit one GPU three times over) ------------------------------------
    if check and E * De * 4 * 4 < 100e9:
        ref = None
        if rank == 0:
            full = kd.RowShardedKGEModel(cfg["model"], E, R, H, cfg["gamma"], True, False, device=dev, seed=11, rank_world=(0, 1))
            ref = kb.KGEModel(cfg["model"], E, R, H, cfg["gamma"], True, False).to(dev)
            with torch.no_grad():
                ref.entity_embedding.copy_(full.entity_embedding)
                ref.relation_embedding.copy_(full.relation_embedding)
            del full

        def make_ref():
            return ref
        wl = dict(cfg)
        ref, dc = run_dist_check(kb, model, step, targs, wl, dev, rank, world, dist, make_ref, seed0=555_000)
        table = ref.entity_embedding.detach().clone() if rank == 0 else torch.empty((E, De), dtype=torch.float32, device=dev)
        rel = ref.relation_embedding.detach().clone() if rank == 0 else torch.empty_like(model.relation_embedding)
        if dist is not None:
            dist.broadcast(table, src=0)
            dist.broadcast(rel, src=0)
        dEt = (model.entity_embedding.detach() - table[lo:hi]).abs() if hi > lo else torch.zeros((1, 1), device=dev)
        dRt = (model.relation_embedding.detach() - rel).abs()
        bulk = 5e-3 * cfg["lr"]
        d = torch.stack([dEt.max(), dRt.max()])
        cnt = torch.stack([(dEt > bulk).sum(), (dRt > bulk).sum() if rank == 0 else torch.zeros((), dtype=torch.int64, device=dev)])
        if dist is not None:
            dist.all_reduce(d, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        if rank == 0:
            frac = float(cnt.sum().item()) / (E * De + dRt.numel())
            dc.update(max_abs_diff_entity=float(d[0]), max_abs_diff_relation=float(d[1]), fraction_above_bulk_tol=frac)
            dc["ok"] = bool(dc["loss_rel_max"] <= dc["tol_loss_rel"] and float(d.max()) <= dc["tol_table"]
                            and frac <= dc["tol_bulk_fraction"])
            out["dist_check"] = dc
        del dEt, dRt
        del table, rel, ref
        torch.cuda.empty_cache()
    nb = 4
    batches = []
    for s in range(nb):
        p, n, sw, m = synthetic_batch(E, R, B, N, seed=9000 + 1000 * rank + s)
        batches.append((p.to(dev), n.to(dev), sw.to(dev), m))

    def steps(k):
        for i in range(k):
            step.train_step(iter([batches[i % nb]]), targs, sync=False)
    steps(3)
    ms = timed(lambda: steps(K))
    step_ms = ms / K
    peak, peak_src = peaks()
    units = B * N
    # per GPU and step: candidate rows once from the compact table in the sweep, once more as the pulled unique rows
    # (read at the owner, written locally), p/m/v read+write of the owned rows, Q rows per use in the apply
    uniq = step.rows_touched()
    bytes_gpu = units * (De * 4 + 12) + uniq * De * 4 * 2 + (hi - lo) * De * 24 + (world * units // world) * (De * 4 + 8)
    out.update({"value": world * units * K / (ms * 1e-3), "ms_per_step": step_ms, "steps": K,
                "unique_rows_per_gpu": uniq, "pulled_bytes_per_gpu": uniq * De * 4,
                "roofline": {"bound": "hbm", "achieved": bytes_gpu / (step_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": bytes_gpu / (step_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_gpu": bytes_gpu,
                             "peak_source": peak_src,
                             "note": "per GPU: B*N candidate rows read in the sweep + unique rows pulled (read + write) + "
                                     "p/m/v of the owned rows (read + write) + one staged Q row per use in the owners' "
                                     "apply; the table does not fit the 126 MB L2"}})
    # the loss every rank reports is that of the global batch: identical on all ranks
    last = step.train_step(iter([batches[0]]), targs)
    t = torch.tensor([last["loss"]], dtype=torch.float64, device=dev)
    if dist is not None:
        allv = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        out["loss_spread_over_ranks"] = float(torch.stack(allv).max() - torch.stack(allv).min())
    out["loss"] = last["loss"]
    del step, model, opt, batches
    torch.cuda.empty_cache()
    return out


def eval_bench(model, w, dev, timed, rank, world=1):
    """Filtered full-entity ranking throughput (queries/s): 1024 head-batch + 1024 tail-batch queries."""
    import numpy as np
    import torch
    import customknowledgegraphembedding_b200 as kb
    rng = np.random.RandomState(7)  # the same queries on every rank: candidates are sharded, not queries
    nq = 1024
    test = np.stack([rng.randint(0, w["nentity"], nq), rng.randint(0, w["nrelation"], nq),
                     rng.randint(0, w["nentity"], nq)], 1).astype(np.int64)
    extra = np.stack([rng.randint(0, w["nentity"], 8 * nq), rng.randint(0, w["nrelation"], 8 * nq),
                      rng.randint(0, w["nentity"], 8 * nq)], 1).astype(np.int64)
    # true triples sharing (h,r) / (r,t) with the test triples so that the filter lists are non-empty
    share = test[rng.randint(0, nq, 4 * nq)].copy()
    share[: 2 * nq, 2] = rng.randint(0, w["nentity"], 2 * nq)
    share[2 * nq:, 0] = rng.randint(0, w["nentity"], 2 * nq)
    all_true = [tuple(int(x) for x in t) for t in np.concatenate([test, extra, share])]
    test_l = [tuple(int(x) for x in t) for t in test]
    maps = model._filter_maps(all_true)
    q = torch.from_numpy(test).to(dev)
    csr = {}
    for mode in ("head-batch", "tail-batch"):
        ip, ix = kb.KGEModel.build_filter_csr(maps, test, mode, w["nentity"])
        csr[mode] = (torch.from_numpy(ip).to(dev), torch.from_numpy(ix).to(dev))

    if world > 1:
        from customknowledgegraphembedding_b200 import dist as kd

    def device_eval():
        if world == 1:  # the two modes side by side on two streams (KGEModel.rank_queries_both_modes)
            model.rank_queries_both_modes(q, csr)
            return
        for mode in ("head-batch", "tail-batch"):
            # each rank scores its entity row range, one all-reduce of the integer counters
            kd.rank_queries_sharded(model, q, mode, *csr[mode])
    device_eval()
    ms = timed(device_eval)
    eargs = argparse.Namespace(nentity=w["nentity"], nrelation=w["nrelation"], countries=False)
    metrics = {}

    def api_eval():
        if world > 1:
            metrics.update(kd.test_step(model, test_l, all_true, eargs))
        else:
            metrics.update(kb.KGEModel.test_step(model, test_l, all_true, eargs))
    api_eval()
    ms_api = timed(api_eval)
    sm_mhz = 1965.0
    pk = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        with open(pk) as f:
            sm_mhz = float(json.load(f).get("sm_max_mhz", sm_mhz))
    coords = 2.0 * nq * w["nentity"] * w["hidden_dim"] / (ms * 1e-3) / max(world, 1)
    sfu_peak = 148 * 16 * sm_mhz * 1e6
    return {"metric": "eval_queries_per_sec", "value": 2 * nq / (ms * 1e-3), "unit": "queries/s",
            "queries": 2 * nq, "candidates_per_query": w["nentity"], "ms": ms,
            "roofline": {"bound": "sfu", "achieved": coords, "peak": sfu_peak, "unit": "pair-coordinates/s per GPU",
                         "frac": coords / sfu_peak,
                         "note": "RotatE needs one MUFU.SQRT per (query, candidate, complex coordinate); the SFU pipe "
                                 "issues 16 per clock per SM: 148 SMs x 16 x %.0f MHz (max SM clock).  The whole call is "
                                 "counted (staging, true scores, tiles, filter pass); the tile grid is 912 CTAs per mode = "
                                 "6.16 waves on 148 SMs" % sm_mhz},
            "sharding": "none" if world == 1 else "entity rows over %d ranks, all-reduce of int64 counters" % world,
            "e2e": {"value": 2 * nq / (ms_api * 1e-3), "unit": "queries/s", "ms": ms_api,
                    "note": "KGEModel.test_step from Python triple lists: CSR filter build on host, H2D, D2H of ranks"},
            "mrr_random_init": metrics.get("MRR")}


def sampler_bench(model, opt, group, w, dev, timed, K):
    """the batch produced on the GPU (kge_sample_negatives + gathers) instead of copied from the host:
    sampler alone, and the train step fed by it.  272,115 synthetic training triples (FB15k-237's train size)."""
    import numpy as np
    import torch
    from customknowledgegraphembedding_b200.sampler import TrainTripleIndex, BidirectionalGpuIterator
    rng = np.random.RandomState(21)
    T = 272115
    triples = np.stack([rng.randint(0, w["nentity"], T), rng.randint(0, w["nrelation"], T), rng.randint(0, w["nentity"], T)], 1)
    idx = TrainTripleIndex(triples, w["nentity"], w["nrelation"], device=dev)
    it = BidirectionalGpuIterator(idx, w["negatives"], w["batch"], seed=3)
    for _ in range(3):
        next(it)

    def only_sampler():
        for _ in range(K):
            next(it)
    ms_s = timed(only_sampler)

    def fed_steps():
        for _ in range(K):
            p, ng, sw, m = next(it)
            model._fused_step(opt, group, p, ng, sw, m, True, w["adversarial_temperature"], False, 0.0)
    fed_steps()
    ms_t = timed(fed_steps)
    units = w["batch"] * w["negatives"]
    return {"sampler_only": {"value": units * K / (ms_s * 1e-3), "unit": "negatives/s", "ms_per_batch": ms_s / K},
            "train_fed_by_gpu_sampler": {"value": units * K / (ms_t * 1e-3), "unit": "triples*neg/s", "ms_per_step": ms_t / K,
                                         "h2d_bytes_per_step": 0},
            "train_triples": T}


def eval_10m_bench(dev, timed, world=1, nq=256):
    """BASELINE.json configs[4] / north_star: filtered full-entity eval on a synthetic 10M-entity graph, RotatE
    H = 1000 (table 80 GB fp32), entity rows SHARDED over the ranks (strong scaling: the same queries, E / N rows per
    GPU, query rows exchanged, one all-reduce of the int64 counters)."""
    import numpy as np
    import torch
    from customknowledgegraphembedding_b200 import dist as kd
    E, R, H = 10_000_000, 237, 1000
    lo, hi = kd.shard_range(E, int(os.environ.get("RANK", "0")), world)
    need = (hi - lo) * 2 * H * 4 + (4 << 30)
    free, _ = torch.cuda.mem_get_info(dev)
    if free < need:
        return {"skipped": "needs %.1f GB free on the device, %.1f GB available" % (need / 1e9, free / 1e9)}
    m = kd.RowShardedKGEModel("RotatE", E, R, H, 9.0, True, False, device=dev, seed=5)
    rng = np.random.RandomState(3)
    q = torch.from_numpy(np.stack([rng.randint(0, E, nq), rng.randint(0, R, nq), rng.randint(0, E, nq)], 1)).to(dev)
    lens = rng.randint(0, 6, nq)
    indptr = np.zeros(nq + 1, dtype=np.int64)
    np.cumsum(lens, out=indptr[1:])
    idx = torch.from_numpy(rng.randint(0, E, int(indptr[-1])).astype(np.int64)).to(dev)
    indptr = torch.from_numpy(indptr).to(dev)
    res = {}

    def run():
        for mode in ("head-batch", "tail-batch"):
            res[mode] = m.rank_queries(q, mode, indptr, idx)
    run()
    ms = timed(run)
    ng = res["tail-batch"][1]
    out = {"metric": "eval_queries_per_sec", "value": 2 * nq / (ms * 1e-3), "unit": "queries/s", "queries": 2 * nq,
           "candidates_per_query": E, "rows_per_gpu": hi - lo, "ms": ms, "scaling": "strong",
           "pair_coordinates_per_s": 2.0 * nq * E * H / (ms * 1e-3),
           "roofline": {"bound": "sfu", "achieved": 2.0 * nq * E * H / (ms * 1e-3) / world, "peak": 148 * 16 * 1965e6,
                        "unit": "pair-coordinates/s per GPU", "frac": 2.0 * nq * E * H / (ms * 1e-3) / world / (148 * 16 * 1965e6)},
           "checksum_n_greater": int(ng.sum().item()),  # identical for every N: the ranks do not depend on the sharding
           "sharding": "entity rows over %d ranks; query rows exchanged (all-reduce of [Q,2,De]); all-reduce of int64 "
                       "counters" % world}
    del m
    torch.cuda.empty_cache()
    return out


def eval_gemm_bench(dev, timed, world=1):
    """BASELINE.json configs[3]: ComplEx / DistMult on FB15k sizes (14951 entities, 1345 relations), H = 1000,
    filtered full-entity ranking as tcgen05 GEMM tiles (3xTF32).  Roofline: tensor pipe; peak = half the measured
    dense bf16 rate (kind::tf32 runs at half the bf16 rate); achieved counts the 3 MMA passes actually issued."""
    import numpy as np
    import torch
    import customknowledgegraphembedding_b200 as kb
    E, R, H, nq = 14951, 1345, 1000, 4096
    bf16 = None
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            bf16 = float(json.load(f).get("bf16_tflops", 0)) or None
    peak, src = (bf16 / 2, "measured bf16_tflops / 2 (MEASURED_PEAKS.json)") if bf16 else (795.0, "fallback 1.59 PF bf16 / 2")
    out = {}
    if world > 1:
        from customknowledgegraphembedding_b200 import dist as kd
    for name, de, dr in (("ComplEx", True, True), ("DistMult", False, False)):
        torch.manual_seed(1)
        m = kb.KGEModel(name, E, R, H, 12.0, de, dr).to(dev)
        rng = np.random.RandomState(11)
        q = torch.from_numpy(np.stack([rng.randint(0, E, nq), rng.randint(0, R, nq), rng.randint(0, E, nq)], 1)).to(dev)
        lens = rng.randint(0, 6, nq)
        indptr = np.zeros(nq + 1, dtype=np.int64)
        np.cumsum(lens, out=indptr[1:])
        idx = torch.from_numpy(rng.randint(0, E, int(indptr[-1])).astype(np.int64)).to(dev)
        indptr = torch.from_numpy(indptr).to(dev)

        def run():
            for mode in ("head-batch", "tail-batch"):
                if world > 1:
                    kd.rank_queries_sharded(m, q, mode, indptr, idx)
                else:
                    m.rank_queries(q, mode, indptr, idx)
        run()
        ms = timed(run)
        flops = 2.0 * (2 * nq) * E * m.entity_dim
        out[name] = {"value": 2 * nq / (ms * 1e-3), "unit": "queries/s", "queries": 2 * nq, "candidates_per_query": E,
                     "ms": ms, "roofline": {"bound": "tensor", "achieved": 3 * flops / (ms * 1e-3) / 1e12 / world, "peak": peak,
                                            "unit": "TFLOP/s per GPU", "frac": 3 * flops / (ms * 1e-3) / 1e12 / peak / world,
                                            "peak_source": src, "useful_tflops": flops / (ms * 1e-3) / 1e12 / world,
                                            "useful_frac": flops / (ms * 1e-3) / 1e12 / peak / world,
                                            "note": "whole call: operand split, true-score tile, sort of the filter "
                                                    "lists and the counting tiles; 3 tf32 MMA passes per product"}}
        del m
    return out


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else a library prints (NCCL's version
    banner, torchrun notices) was redirected to stderr in main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)  # fd 1 -> stderr for the rest of the run
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--deterministic", action="store_true", help="use the fixed-order backward variant")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eval10m", action="store_true", help="skip the 10M-entity sharded eval section (80 GB table)")
    ap.add_argument("--train-only", action="store_true", help="tuning loops: only the headline train step sections")
    ap.add_argument("--sharded", action="store_true", help="with --train-only: also the row-sharded train sections (configs[4])")
    ap.add_argument("--transparse", action="store_true", help="with --train-only: also time the TranSparse train step (8(f) N4)")
    ap.add_argument("--interht", action="store_true", help="with --train-only: also time BASELINE configs[2] (InterHT)")
    ap.add_argument("--no-dist-check", action="store_true", help="N > 1: skip the correctness check around the timed region")
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)  # K steps as asked; the positives per step shrink until the run fits ~3 minutes (cpu_baseline)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
