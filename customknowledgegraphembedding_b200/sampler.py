"""On-GPU producer of the train batch: the ``TrainDataset`` + ``BidirectionalOneShotIterator`` of the
PyTorch codes (TF twin: /root/reference/compress_data/main.py:58-94) without the CPU workers and
without the per-step host->device copy of ``(positive, negative, weight)``.

* subsampling weights ``sqrt(1 / (count(h, r) + count(t, -r-1)))`` with counts starting at 4 -- computed once per
  training triple (integer counts, then the same two fp32 operations upstream applies);
* negatives: ``kge_sample_negatives`` (csrc/kge_sampler.cu) -- uniform over the entities that do not form a known
  training triple, Philox4x32-10 keyed by (seed, step, global positive index, slot);
* the iterator alternates head-batch / tail-batch every step and walks an independent random permutation of the
  training triples per mode, like the two shuffling ``DataLoader``s upstream.

Yields ``(positive_sample [B,3] i64, negative_sample [B,N] i64, subsampling_weight [B] f32, mode)`` on the device:
exactly what ``KGEModel.train_step`` / ``DataParallelStep.train_step`` consume.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _lib, ops


class TrainTripleIndex:
    """Device-resident index of the training triples: sorted true-set composites per mode + per-triple weights."""

    def __init__(self, train_triples, nentity: int, nrelation: int, device="cuda", count_start: int = 4):
        T = torch.as_tensor(np.asarray(train_triples, dtype=np.int64).reshape(-1, 3))
        self.nentity, self.nrelation = int(nentity), int(nrelation)
        if T.numel() and (int(T[:, [0, 2]].max()) >= nentity or int(T[:, 1].max()) >= nrelation or int(T.min()) < 0):
            raise ValueError("triple id out of range")
        if nrelation * nentity * nentity >= 2 ** 63:
            raise ValueError("nrelation * nentity^2 must fit int64")
        dev = torch.device(device)
        self.triples = T.to(dev)
        h, r, t = self.triples[:, 0], self.triples[:, 1], self.triples[:, 2]
        E = self.nentity
        # true_head[(r, t)] / true_tail[(h, r)] as one sorted composite array each (duplicates removed, like set())
        self.true_sorted = {
            "head-batch": torch.unique((r * E + t) * E + h),
            "tail-batch": torch.unique((r * E + h) * E + t),
        }
        # count[(h, r)] and count[(t, -r-1)]: `start` + number of occurrences - 1 + 1 (upstream starts at `start`)
        def counts(key):
            _, inv, cnt = torch.unique(key, return_inverse=True, return_counts=True)
            return cnt[inv]
        c = (counts(h * self.nrelation + r) + (count_start - 1)) + (counts(t * self.nrelation + r) + (count_start - 1))
        # torch.sqrt(1 / torch.Tensor([c])): fp32 division, fp32 sqrt (both correctly rounded on the GPU as on the CPU)
        self.weights = torch.sqrt(1.0 / c.to(torch.float32))

    def __len__(self):
        return int(self.triples.shape[0])


def sample_negatives(index: TrainTripleIndex, positive: torch.Tensor, negative_sample_size: int, mode: str, seed: int,
                     step: int, first_index: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[B,N] int64 negatives for ``positive`` [B,3] (device): thin call of ``kge_sample_negatives``."""
    if mode not in ("head-batch", "tail-batch"):
        raise ValueError("Training batch mode %s not supported" % mode)
    lib = _lib.load()
    positive = ops._need_cuda("positive_sample", positive, torch.int64)
    B = positive.shape[0]
    if out is None:
        out = torch.empty((B, negative_sample_size), dtype=torch.int64, device=positive.device)
    ts = index.true_sorted[mode]
    with torch.cuda.device(positive.device):
        _lib.check(lib.kge_sample_negatives(ops._ptr(positive), B, int(negative_sample_size), ops.mode_id(mode),
                                            index.nentity, ops._ptr(ts), ts.numel(), C.c_uint64(seed & (2 ** 64 - 1)),
                                            C.c_uint64(first_index), C.c_uint32(step & 0xFFFFFFFF), ops._ptr(out),
                                            ops._stream()), "kge_sample_negatives")
    return out


class BidirectionalGpuIterator:
    """``BidirectionalOneShotIterator(train_dataloader_head, train_dataloader_tail)`` on the device.

    ``rank`` / ``world``: every rank walks the same permutations and takes its contiguous share of each global batch,
    so the union over ranks is the batch a single process would have produced (the sampler is keyed by the global
    positive index).  With ``world > 1`` the short last batch of an epoch is dropped (``drop_last`` defaults to True
    there): the multi-GPU steps exchange fixed-shape per-positive buffers and need the same B on every rank."""

    def __init__(self, index: TrainTripleIndex, negative_sample_size: int, batch_size: int, seed: int = 0,
                 rank: int = 0, world: int = 1, drop_last: Optional[bool] = None):
        self.index, self.N, self.B = index, int(negative_sample_size), int(batch_size)
        if drop_last is None:
            drop_last = int(world) > 1
        if int(world) > 1 and not drop_last:
            raise ValueError("BidirectionalGpuIterator: world > 1 needs drop_last=True (a short last batch would give the "
                             "ranks unequal, possibly empty, shares)")
        if int(world) > 1 and len(index) < self.B * int(world):
            raise ValueError("fewer training triples (%d) than one global batch (%d x %d)" % (len(index), self.B, world))
        self.seed, self.rank, self.world, self.drop_last = int(seed), int(rank), int(world), bool(drop_last)
        self.step = 0
        self._state = {m: {"epoch": -1, "perm": None, "cursor": 0, "served": 0} for m in ("head-batch", "tail-batch")}

    def _next_positions(self, mode):
        st = self._state[mode]
        n = len(self.index)
        if st["perm"] is None or st["cursor"] >= n or (self.drop_last and st["cursor"] + self.B * self.world > n):
            st["epoch"] += 1
            g = torch.Generator(device=self.index.triples.device)
            g.manual_seed(self.seed * 1000003 + 2 * st["epoch"] + (0 if mode == "head-batch" else 1))
            st["perm"] = torch.randperm(n, device=self.index.triples.device, generator=g)
            st["cursor"] = 0
        lo = st["cursor"]
        hi = min(lo + self.B * self.world, n)
        st["cursor"] = hi
        first = st["served"]
        st["served"] += hi - lo
        share = (hi - lo + self.world - 1) // self.world
        a, b = min(lo + self.rank * share, hi), min(lo + (self.rank + 1) * share, hi)
        return st["perm"][a:b], first + (a - lo)

    def __iter__(self):
        return self

    def __next__(self):
        mode = "head-batch" if self.step % 2 == 0 else "tail-batch"
        pos_idx, first = self._next_positions(mode)
        positive = self.index.triples[pos_idx].contiguous()
        weight = self.index.weights[pos_idx].contiguous()
        negative = sample_negatives(self.index, positive, self.N, mode, self.seed, self.step, first_index=first)
        self.step += 1
        return positive, negative, weight, mode
