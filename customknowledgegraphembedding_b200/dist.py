"""Multi-GPU plumbing: one process per GPU, ``torch.distributed`` (NCCL) between the kernel calls.

The path shards twice:

* **eval** -- candidates are row-range sharded: each rank counts ``score > true`` over its entity range for ALL
  queries (``kge_eval_rank(entity_lo, entity_hi)``) and the integer counters are summed with one all-reduce.
* **train** -- positives are data-parallel in the sweep (``kge_train_sweep``), entity rows are owner-parallel in
  the apply (``kge_train_apply``): the per-positive buffers are all-gathered, each rank gathers gradients and
  applies Adam for the rows it owns, the updated row ranges are broadcast back.  The loss and every update are
  those of ONE step on the global batch (sum of subsampling weights all-reduced first), bit-identical on every
  rank and independent of which rank swept which positive.

The reference's own mechanism is replicas with an implicit gradient all-reduce under ``tf.distribute``
(/root/reference/tensorflow_codes/run.py:8-17, supervisor.py:26-30); this keeps its semantics (replicated
tables, one optimizer step per global batch) and replaces the dense gradient all-reduce.

The two kernel calls are the methods ``_sweep`` / ``_apply`` (and ``local_fn`` for eval) so that the host logic
is testable on CPU over ``gloo`` with stand-ins (tests/test_dist_cpu.py).
"""
from __future__ import annotations

import contextlib
import os
import ctypes as C
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import _lib, ops
from ._lib import AdamOpts, LossOpts, StepBuffers


def shard_range(n: int, rank: int, world: int):
    """Contiguous row range of ``rank``: ceil(n / world) rows each, the tail ranks may be short or empty."""
    rows = (n + world - 1) // world
    lo = min(rank * rows, n)
    return lo, min(lo + rows, n)


def _world(group=None):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _coalesced(group, device, async_ops=False):
    """Batch several NCCL calls into one group launch when torch exposes the coalescing manager.  With
    ``async_ops`` the context yields a handle whose ``wait()`` makes the current stream wait for the batch."""
    cm = getattr(dist, "_coalescing_manager", None)
    if cm is None or device.type != "cuda":
        return contextlib.nullcontext()
    try:
        return cm(group=group, device=device, async_ops=async_ops)
    except TypeError:  # signature differs in this torch: fall back to one launch per call
        return contextlib.nullcontext()


# --------------------------------------------------------------------------------------------------
# eval
# --------------------------------------------------------------------------------------------------
def rank_queries_sharded(model, queries, mode, filter_indptr=None, filter_idx=None, group=None, local_fn=None):
    """(true_score, n_greater, n_equal) over ALL entities, each rank scoring only its row range."""
    rank, world = _world(group)
    lo, hi = shard_range(model.nentity, rank, world)
    fn = local_fn if local_fn is not None else model.rank_queries
    true_score, n_greater, n_equal = fn(queries, mode, filter_indptr, filter_idx, entity_lo=lo, entity_hi=hi)
    if world > 1:
        counts = torch.stack([n_greater, n_equal])
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
        n_greater, n_equal = counts[0], counts[1]
    return true_score, n_greater, n_equal


def test_step(model, test_triples, all_true_triples, args, group=None, local_fn=None):
    """``KGEModel.test_step`` with the candidates sharded over the ranks of ``group``; same metrics on every rank."""
    def rank_fn(queries, mode, indptr, idx):
        return rank_queries_sharded(model, queries, mode, indptr, idx, group=group, local_fn=local_fn)
    return type(model).test_step(model, test_triples, all_true_triples, args, rank_fn=rank_fn)


# --------------------------------------------------------------------------------------------------
# train
# --------------------------------------------------------------------------------------------------
_BUFFERS = ("S", "Q", "GH", "GT", "GR", "GM", "stats", "terms")


def _on(dev):
    """Device guard only when the tensor's device is not the current one (the guard costs ~10 us of host time)."""
    if dev.type != "cuda" or torch.cuda.current_device() == dev.index:
        return contextlib.nullcontext()
    return torch.cuda.device(dev)


class DataParallelStep:
    """``step.train_step(train_iterator, args)`` == ``KGEModel.train_step`` on the concatenation of every rank's batch."""

    def __init__(self, model, optimizer, group=None):
        self.model, self.optimizer, self.group = model, optimizer, group
        self.rank, self.world = _world(group)
        self.lo, self.hi = shard_range(model.nentity, self.rank, self.world)
        self._local: Dict[str, torch.Tensor] = {}
        self._gathered: Dict[str, torch.Tensor] = {}
        self._ws: Optional[torch.Tensor] = None
        self._shape = None
        # per-positive buffers exchanged by peer stores (symmetric memory) instead of an NCCL all-gather
        self.peer_exchange = os.environ.get("KGE_DP_PEER_EXCHANGE", "1") != "0"
        self._peer = None
        self._rows_peer = None  # entity table in symmetric memory: updated rows are stored into the other replicas
        self._wsum_ahead = {}   # id(staged batch) -> all-reduced weight sum queued one step early

    # ---- buffers -----------------------------------------------------------------------------------
    def _row_sizes(self, N):
        m = self.model
        planes = {"ComplEx": 2, "RotatE": 2, "InterHT": 3}.get(m.model_name, 1)
        return {"S": N, "Q": planes * m.hidden_dim, "GH": m.entity_dim, "GT": m.entity_dim, "GR": m.relation_dim,
                "GM": 1, "stats": 2, "terms": 2}

    def _ensure(self, B, N, dev):
        if self._shape == (B, N, dev):
            return
        if B <= 0 or N <= 0:
            raise ValueError("train_step: empty batch on rank %d (B=%d, N=%d)" % (self.rank, B, N))
        if self.world > 1 and dist.is_initialized():
            # gathered buffers, symmetric-memory offsets and B_total are all world * B: every rank must bring the same
            # shape.  One small all-reduce per shape CHANGE (not per step); a mismatch raises on every rank.
            shp = torch.tensor([B, -B, N, -N], dtype=torch.int64, device=dev)
            dist.all_reduce(shp, op=dist.ReduceOp.MAX, group=self.group)
            bmax, bmin, nmax, nmin = int(shp[0]), -int(shp[1]), int(shp[2]), -int(shp[3])
            if bmax != bmin or nmax != nmin:
                raise ValueError("train_step: ranks bring different batch shapes (B %d..%d, N %d..%d); use "
                                 "drop_last=True / equal shares per rank" % (bmin, bmax, nmin, nmax))
        sizes = self._row_sizes(N)
        self._local = {k: torch.empty(B * s, dtype=torch.float32, device=dev) for k, s in sizes.items()}
        self._peer = None
        if self.world > 1 and dev.type == "cuda" and self.peer_exchange:
            self._peer = self._peer_buffers(B, N, dev, sizes)
            if self._peer is not None and self._rows_peer is None and self.world <= 16:
                self._rows_peer = self._peer_table()
        if self._peer is None:
            self._gathered = {k: torch.empty(self.world * B * s, dtype=torch.float32, device=dev) for k, s in sizes.items()}
            self._gathered["pos"] = torch.empty((self.world * B, 3), dtype=torch.int64, device=dev)
            self._gathered["neg"] = torch.empty((self.world * B, N), dtype=torch.int64, device=dev)
        self._shape = (B, N, dev)
        self._structs = {}  # pointer structs of the buffers above, built on first use

    def _bufs(self, which):
        """StepBuffers of the local / gathered buffers (cached: building one costs ~10 us of host time per call)."""
        src = self._local if which == "local" else self._gathered
        key = (which,) + tuple(src[k].data_ptr() for k in ("S", "Q"))
        hit = self._structs.get(which)
        if hit is None or hit[0] != key:
            hit = self._structs[which] = (key, StepBuffers(*[C.c_void_p(src[k].data_ptr()) for k in _BUFFERS]))
        return hit[1]

    def _peer_buffers(self, B, N, dev, sizes):
        """The gathered buffers carved from ONE symmetric-memory block, so that every rank can store into every other
        rank's copy (kge_train_exchange).  Returns the peer pointer tables, or None when this torch has no symmetric
        memory (the NCCL all-gather is used then).  Collective: every rank takes the same branch."""
        try:
            import torch.distributed._symmetric_memory as symm
        except ImportError:
            return None
        G = self.world
        items = [(k, G * B * sizes[k] * 4, torch.float32, None) for k in _BUFFERS]
        items += [("pos", G * B * 3 * 8, torch.int64, (G * B, 3)), ("neg", G * B * N * 8, torch.int64, (G * B, N))]
        offsets, total = {}, 0
        for k, nbytes, _, _ in items:
            offsets[k] = total
            total += (nbytes + 255) & ~255
        block = symm.empty(total, dtype=torch.uint8, device=dev)
        hdl = symm.rendezvous(block, self.group if self.group is not None else dist.group.WORLD)
        self._gathered = {}
        for k, nbytes, dt, shape in items:
            t = block[offsets[k]:offsets[k] + nbytes].view(dt)
            self._gathered[k] = t.view(shape) if shape is not None else t
        bases = [int(p) for p in hdl.buffer_ptrs]
        tables = (StepBuffers * G)(*[StepBuffers(*[C.c_void_p(b + offsets[k]) for k in _BUFFERS]) for b in bases])
        pos_ptrs = (C.c_void_p * G)(*[C.c_void_p(b + offsets["pos"]) for b in bases])
        neg_ptrs = (C.c_void_p * G)(*[C.c_void_p(b + offsets["neg"]) for b in bases])
        return {"hdl": hdl, "block": block, "tables": tables, "pos": pos_ptrs, "neg": neg_ptrs}

    def _peer_table(self):
        """Move the (replicated) entity table into symmetric memory, once: every rank can then store the rows it
        owns straight into the other replicas from the Adam epilogue (kge_train_apply_peers)."""
        import torch.distributed._symmetric_memory as symm
        ent = self.model.entity_embedding
        block = symm.empty(ent.numel(), dtype=torch.float32, device=ent.device)
        block.copy_(ent.data.reshape(-1))
        hdl = symm.rendezvous(block, self.group if self.group is not None else dist.group.WORLD)
        ent.data = block.view(ent.shape)
        bases = [int(p) for p in hdl.buffer_ptrs]
        others = [bases[r] for r in range(self.world) if r != self.rank]
        # NVSwitch multicast mapping of the block (0 where the fabric has none): one multimem.st reaches every replica
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0) if os.environ.get("KGE_DP_MULTICAST", "1") != "0" else 0
        return {"hdl": hdl, "block": block, "ptrs": (C.c_void_p * len(others))(*[C.c_void_p(b) for b in others]),
                "mc": C.c_void_p(mc) if mc else None}

    def _exchange(self, pos, neg):
        """kge_train_exchange on the current stream; the caller follows it with the barrier that orders every rank's
        apply after every rank's stores."""
        B, N = neg.shape
        local = self._bufs("local")
        with _on(pos.device):
            _lib.check(_lib.load().kge_train_exchange(
                C.byref(self.model._desc()), ops._ptr(pos), ops._ptr(neg), B, N, C.byref(local), self.rank, self.world,
                self._peer["tables"], self._peer["pos"], self._peer["neg"], ops._stream()), "kge_train_exchange")

    # ---- the two kernel calls (overridable: tests/test_dist_cpu.py) ------------------------------------
    def _sweep(self, mode, pos, neg, w, wsum, B_total, loss_opts):
        m = self.model
        B, N = neg.shape
        bufs = self._bufs("local")
        sws = self._sweep_ws(m._desc(), B, N, pos.device)
        with _on(pos.device):
            _lib.check(_lib.load().kge_train_sweep(
                C.byref(m._desc()), ops.mode_id(mode), ops._ptr(m.entity_embedding), ops._ptr(m.relation_embedding),
                ops._ptr(m._modulus()), ops._ptr(pos), ops._ptr(neg), ops._ptr(w), B, N, C.byref(loss_opts),
                ops._ptr(wsum), B_total, C.byref(bufs), ops._ptr(sws), sws.numel(), ops._stream()), "kge_train_sweep")

    def _sweep_ws(self, desc, B, N, dev):
        """Workspace of the split (warp-per-row) sweep: per-item partial dL/dq + item counter."""
        need = ops.workspace_bytes(_lib.WS_TRAIN_SWEEP, desc, B, N)
        ws = getattr(self, "_sws", None)
        if ws is None or ws.numel() < need or ws.device != dev:
            ws = torch.empty(need, dtype=torch.uint8, device=dev)
            self._sws = ws
        return ws

    def _apply(self, mode, pos_all, neg_all, loss_opts, adam, out, row_peers=None, multicast=None):
        """kge_train_apply_peers on the rows this rank owns; ``row_peers``: ctypes array of the other ranks' table
        pointers / ``multicast``: the NVSwitch multicast mapping of the table -- updated rows are stored there too."""
        m, opt = self.model, self.optimizer
        B_all, N = neg_all.shape
        ent, rel, mod = m.entity_embedding, m.relation_embedding, m._modulus()
        st_e, st_r = opt.state[ent], opt.state[rel]
        st_m = opt.state[mod] if mod is not None else None
        desc = m._desc()
        need = ops.workspace_bytes(_lib.WS_TRAIN_APPLY, desc, B_all, N)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=ent.device)
        bufs = self._bufs("gathered")
        P = ops._ptr
        with _on(ent.device):
            _lib.check(_lib.load().kge_train_apply_peers(
                C.byref(desc), ops.mode_id(mode), P(ent), P(rel), P(mod), P(st_e["exp_avg"]), P(st_e["exp_avg_sq"]),
                P(st_r["exp_avg"]), P(st_r["exp_avg_sq"]), P(st_m["exp_avg"]) if st_m else None,
                P(st_m["exp_avg_sq"]) if st_m else None, P(pos_all), P(neg_all), B_all, N, C.byref(loss_opts),
                C.byref(adam), C.byref(bufs), self.lo, self.hi, 0 if row_peers is None else len(row_peers), row_peers,
                multicast, P(out), C.c_void_p(out.data_ptr() + 12), P(self._ws), self._ws.numel(), ops._stream()),
                "kge_train_apply_peers")

    # ---- one global step -----------------------------------------------------------------------------
    def train_step(self, train_iterator, args, sync: bool = True):
        m, opt = self.model, self.optimizer
        m.train()
        # input lookahead, as in KGEModel.train_step: the batch of this call was taken from the iterator and copied H2D
        # on the copy stream while the previous step ran
        ahead, staged = m._ahead.pop(id(train_iterator), None) if m._ahead else None, None
        if ahead is not None:
            (positive_sample, negative_sample, subsampling_weight, mode), staged = ahead[1], ahead[2]
        else:
            positive_sample, negative_sample, subsampling_weight, mode = next(train_iterator)
        if mode not in ("head-batch", "tail-batch"):
            raise ValueError("mode %s not supported" % mode)
        group_opts = m._fused_adam_plan(opt)
        if group_opts is None:
            raise RuntimeError("DataParallelStep needs torch.optim.Adam / FusedAdam over exactly the model's parameters")
        dev = m.entity_embedding.device
        adv = bool(getattr(args, "negative_adversarial_sampling", False))
        uni = bool(getattr(args, "uni_weight", False))
        reg = float(getattr(args, "regularization", 0.0))
        loss_opts = LossOpts(int(adv), int(uni), 1, 0, float(getattr(args, "adversarial_temperature", 1.0)), reg)
        lookahead = (sync and m.input_lookahead and dev.type == "cuda" and not positive_sample.is_cuda
                     and torch.cuda.current_device() == dev.index)
        if lookahead:
            if staged is None or staged[2] is None or staged[0].device != dev:
                staged = m._stage_host_batch(_lib.load(), positive_sample, negative_sample, subsampling_weight, False, dev)
            _lib.check(_lib.load().kge_stream_wait_event(ops._stream(), staged[6]), "kge_stream_wait_event")
            pos, neg, w = staged[0], staged[1], staged[2]
        else:
            pos = positive_sample.to(dev, non_blocking=True).contiguous()
            neg = negative_sample.to(dev, non_blocking=True).contiguous()
            w = subsampling_weight.to(dev, dtype=torch.float32, non_blocking=True).contiguous()
        B, N = neg.shape
        self._ensure(B, N, dev)
        # global sum of the subsampling weights: the loss is that of ONE step on the global batch.  With the lookahead
        # it was queued behind the previous step (see below), so this step starts with the sweep.
        wsum = self._wsum_ahead.pop(id(staged), None) if lookahead else None
        if wsum is None:
            wsum = w.sum().reshape(1)
            if self.world > 1:
                dist.all_reduce(wsum, op=dist.ReduceOp.SUM, group=self.group)
        if self.world > 1 and self._peer is not None:
            self._sweep(mode, pos, neg, w, wsum, self.world * B, loss_opts)
            self._exchange(pos, neg)  # peer stores over NVLink: GH / GT rows go to their owners only
            self._peer["hdl"].barrier()
            pos_all, neg_all = self._gathered["pos"], self._gathered["neg"]
        elif self.world > 1:
            self._sweep(mode, pos, neg, w, wsum, self.world * B, loss_opts)
            with _coalesced(self.group, dev):
                dist.all_gather_into_tensor(self._gathered["pos"], pos, group=self.group)
                dist.all_gather_into_tensor(self._gathered["neg"], neg, group=self.group)
                for k in _BUFFERS:
                    dist.all_gather_into_tensor(self._gathered[k], self._local[k], group=self.group)
            pos_all, neg_all = self._gathered["pos"], self._gathered["neg"]
        else:
            self._sweep(mode, pos, neg, w, wsum, self.world * B, loss_opts)
            pos_all, neg_all = pos, neg
            for k in _BUFFERS:
                self._gathered[k] = self._local[k]
        states = [m._adam_state(opt, p) for p in (m.entity_embedding, m.relation_embedding)]
        if m._modulus() is not None:
            states.append(m._adam_state(opt, m._modulus()))
        for s in states:
            s["step"] += 1
        b1, b2 = group_opts["betas"]
        adam = AdamOpts(float(group_opts["lr"]), float(b1), float(b2), float(group_opts["eps"]), int(states[0]["step"]), 0)
        out = torch.zeros(4, dtype=torch.float32, device=dev)
        plain = type(self)._apply is DataParallelStep._apply and type(self)._share_rows is DataParallelStep._share_rows
        if self.world > 1 and plain and self._rows_peer is not None and reg == 0.0:
            # the broadcast of the owned rows rides on the Adam epilogue as NVLink stores; one barrier closes the step
            self._apply(mode, pos_all, neg_all, loss_opts, adam, out, row_peers=self._rows_peer["ptrs"],
                        multicast=self._rows_peer["mc"])
            self._rows_peer["hdl"].barrier()
        else:
            self._apply(mode, pos_all, neg_all, loss_opts, adam, out)
            self._share_rows()
        if not sync:
            return out
        if lookahead:
            m._take_ahead(_lib.load(), train_iterator, False, dev)  # next(iterator) + its H2D copies under this step
            nxt = m._ahead.get(id(train_iterator))
            self._wsum_ahead.clear()
            if nxt is not None and nxt[2] is not None:
                # the weight sum of batch k+1 and its all-reduce go behind step k on the main stream: every rank queues
                # the same collectives in the same order (all ranks run the same number of steps)
                _lib.check(_lib.load().kge_stream_wait_event(ops._stream(), nxt[2][6]), "kge_stream_wait_event")
                nsum = nxt[2][2].sum().reshape(1)
                if self.world > 1:
                    dist.all_reduce(nsum, op=dist.ReduceOp.SUM, group=self.group)
                self._wsum_ahead[id(nxt[2])] = nsum
        pos_l, neg_l, loss, regv = out.tolist()
        if lookahead and m._stage is not None and m._stage["key"] == staged[5]:
            m._stage["free"].append(staged)  # the step is over: its staging set can take a new batch
        log = {"regularization": regv} if reg != 0.0 else {}
        log.update({"positive_sample_loss": pos_l, "negative_sample_loss": neg_l, "loss": loss})
        return log

    def _share_rows(self):
        """Every rank broadcasts the entity rows it owns (its Adam update) to the others."""
        if self.world <= 1:
            return
        ent = self.model.entity_embedding.data
        with _coalesced(self.group, ent.device):
            for r in range(self.world):
                lo, hi = shard_range(self.model.nentity, r, self.world)
                if hi > lo:
                    src = dist.get_global_rank(self.group, r) if self.group is not None else r
                    dist.broadcast(ent[lo:hi], src=src, group=self.group)


# --------------------------------------------------------------------------------------------------
# row-sharded entity table (BASELINE configs[4]: 10M entities x 2000 floats = 80 GB)
# --------------------------------------------------------------------------------------------------
from .model import KGEModel  # noqa: E402  (after the replicated-table helpers on purpose)


class RowShardedKGEModel(KGEModel):
    """``KGEModel`` whose ``entity_embedding`` holds only the rows ``[lo, hi)`` this rank owns.

    ``nentity`` stays the GLOBAL entity count (ids in triples are global); the relation table is replicated.
    ``rank_queries`` / ``test_step`` give the same integers as the unsharded model: the rows the queries name
    are exchanged (owners fill ``[Q,2,De]``, one all-reduce over zeros elsewhere -- exact), every rank counts
    ``score > true`` over its own rows with ``kge_eval_rank_shard`` and the counters are summed.
    """

    INIT_BLOCK = 250_000  # rows per initialisation stream

    def __init__(self, model_name, nentity, nrelation, hidden_dim, gamma, double_entity_embedding=False,
                 double_relation_embedding=False, triple_relation_embedding=False, group=None, device=None, seed=0,
                 rank_world=None):
        # rank_world=(r, w): shard geometry without a process group (one-GPU emulation in the tests)
        rank, world = rank_world if rank_world is not None else _world(group)
        lo, hi = shard_range(nentity, rank, world)
        # build the small tensors through the parent with a 1-row placeholder table, then install the shard
        super().__init__(model_name, 1, nrelation, hidden_dim, gamma, double_entity_embedding,
                         double_relation_embedding, triple_relation_embedding)
        self.nentity = nentity
        self.group, self.rank, self.world, self.lo, self.hi = group, rank, world, lo, hi
        dev = torch.device(device) if device is not None else torch.device("cpu")
        # shard-local generation that does not depend on the sharding: the table is defined in blocks
        # of INIT_BLOCK rows, block i drawn from its own stream; a rank draws the blocks that overlap its rows.
        r = self.embedding_range.item()
        shard = torch.empty((hi - lo, self.entity_dim), dtype=torch.float32, device=dev)
        blk = self.INIT_BLOCK
        for b0 in range((lo // blk) * blk, hi, blk):
            b1 = min(b0 + blk, nentity)
            g = torch.Generator(device=dev).manual_seed(int(seed) * 1000003 + b0 // blk)
            rows = torch.empty((b1 - b0, self.entity_dim), dtype=torch.float32, device=dev).uniform_(-r, r, generator=g)
            s0, s1 = max(b0, lo), min(b1, hi)
            shard[s0 - lo:s1 - lo] = rows[s0 - b0:s1 - b0]
            del rows
        self.entity_embedding = torch.nn.Parameter(shard)
        g2 = torch.Generator().manual_seed(int(seed))  # the replicated relation table: same on every rank
        rel = torch.empty((nrelation, self.relation_dim), dtype=torch.float32).uniform_(-r, r, generator=g2)
        self.relation_embedding = torch.nn.Parameter(rel)
        if device is not None:
            self.to(dev)

    @classmethod
    def from_full(cls, model: KGEModel, group=None, rank_world=None):
        """Shard an existing (replicated) model: this rank keeps rows [lo, hi).  For tests and checkpoints."""
        de = model.entity_dim == 2 * model.hidden_dim
        tr = model.relation_dim == 3 * model.hidden_dim
        dr = model.relation_dim == 2 * model.hidden_dim
        dev = model.entity_embedding.device
        m = cls(model.model_name, model.nentity, model.nrelation, model.hidden_dim, model.gamma.item(), de, dr, tr,
                group=group, device=dev, rank_world=rank_world)
        with torch.no_grad():
            m.entity_embedding.copy_(model.entity_embedding[m.lo:m.hi])
            m.relation_embedding.copy_(model.relation_embedding)
            if model._modulus() is not None:
                m.modulus.copy_(model.modulus)
        m.interht_eps = model.interht_eps
        return m

    def forward(self, sample, mode="single"):
        raise NotImplementedError("RowShardedKGEModel: forward on a row-sharded table goes through "
                                  "RowShardedStep / rank_queries (the rows have to be exchanged first)")

    # ---- row exchange --------------------------------------------------------------------------------
    def gather_rows(self, ids: torch.Tensor) -> torch.Tensor:
        """Rows of the GLOBAL table for ``ids`` (same tensor on every rank) -> [len(ids), De] on every rank."""
        ent = self.entity_embedding.detach()
        ids = ids.to(ent.device)
        mine = (ids >= self.lo) & (ids < self.hi)
        out = torch.zeros((ids.numel(), self.entity_dim), dtype=torch.float32, device=ent.device)
        sel = torch.nonzero(mine).view(-1)
        if sel.numel():
            out[sel] = ent[ids[sel] - self.lo]
        if self.world > 1 and dist.is_initialized():
            dist.all_reduce(out, op=dist.ReduceOp.SUM, group=self.group)  # exactly one rank holds each row: x + 0
        return out

    # ---- eval ------------------------------------------------------------------------------------------
    @torch.no_grad()
    def rank_queries(self, queries, mode, filter_indptr=None, filter_idx=None, entity_lo=None, entity_hi=None,
                     local_fn=None):
        dev = self.entity_embedding.device
        queries = queries.to(dev).contiguous()
        Q = queries.shape[0]
        qrows = self.gather_rows(queries[:, [0, 2]].reshape(-1)).view(Q, 2, self.entity_dim)
        to = lambda t: None if t is None else t.to(dev, non_blocking=True)
        fn = local_fn if local_fn is not None else self._rank_local
        true_score, n_greater, n_equal = fn(queries, qrows, mode, to(filter_indptr), to(filter_idx))
        if self.world > 1 and dist.is_initialized():
            counts = torch.stack([n_greater, n_equal])
            dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=self.group)
            n_greater, n_equal = counts[0], counts[1]
        return true_score, n_greater, n_equal

    def _rank_local(self, queries, qrows, mode, indptr, idx):
        mod = self._modulus()
        return ops.eval_rank_shard(self._desc(), mode, self.entity_embedding.detach(), self.lo,
                                   self.relation_embedding.detach(), None if mod is None else mod.detach(), queries,
                                   qrows, self.lo, self.hi, indptr, idx)


class RowShardedStep(DataParallelStep):
    """Train step over a ``RowShardedKGEModel``: no rank holds the full entity table or its Adam state.

    Per step (north_star "all-gather of the touched rows ... sparse gradients to the owners"):

    1. the unique entity ids of the LOCAL batch are requested from their owners (all-to-all of id lists, all-to-all of
       rows) into a compact table ``[U, De]``; the batch is remapped to compact ids;
    2. ``kge_train_sweep`` on the compact table: scores, loss terms, dL/dq, compact per-positive gradient rows;
    3. all-gather of the per-positive buffers and of the batches (global ids), exactly as ``DataParallelStep``;
    4. ``kge_train_apply_shard`` on the rows this rank owns: gradient gather in fixed order + dense Adam in place.

    The update is that of ONE step on the global batch, bit-identical to ``DataParallelStep`` on replicated tables
    (the sweep reads the same row values; the apply is the same kernel on a rebased pointer).
    """

    def __init__(self, model, optimizer, group=None):
        if not isinstance(model, RowShardedKGEModel):
            raise TypeError("RowShardedStep needs a RowShardedKGEModel")
        super().__init__(model, optimizer, group)
        self.rank, self.world = model.rank, model.world
        self.lo, self.hi = model.lo, model.hi
        self.last_exchange = {}
        # kernel path (kge_shard_compact + kge_shard_fetch_rows): touched ids by bitmap + scan, rows pulled from the
        # owners' shards through peer mappings -- no torch.unique, no all-to-all, no host sync.  Needs CUDA and, with
        # world > 1, the shards in symmetric memory; otherwise (gloo tests, KGE_SHARD_KERNELS=0) the torch / NCCL
        # all-to-all plumbing below is used.
        self.use_kernels = os.environ.get("KGE_SHARD_KERNELS", "1") != "0"
        self.shard_ptrs = None      # ctypes array of `world` device pointers (tests inject emulated shards here)
        self._shard_hdl = None
        self._cstate = None         # compact-table buffers of the kernel path

    # ---- kernel path -----------------------------------------------------------------------------------------
    def _peer_table(self):
        """(called by _ensure once the peer buffers exist) Move this rank's SHARD into symmetric memory so that every
        rank can read the rows it needs straight from their owner.  No replica of the table exists anywhere."""
        import torch.distributed._symmetric_memory as symm
        ent = self.model.entity_embedding
        rows_per = (self.model.nentity + self.world - 1) // self.world
        block = symm.empty(max(rows_per, 1) * self.model.entity_dim, dtype=torch.float32, device=ent.device)
        block[:ent.numel()].copy_(ent.data.reshape(-1))
        hdl = symm.rendezvous(block, self.group if self.group is not None else dist.group.WORLD)
        ent.data = block[:ent.numel()].view(ent.shape)
        self.shard_ptrs = (C.c_void_p * self.world)(*[C.c_void_p(int(p)) for p in hdl.buffer_ptrs])
        self._shard_hdl = hdl
        hdl.barrier()  # every shard is in place before anybody's first pull
        return {"hdl": hdl, "block": block, "ptrs": None, "mc": None}

    def _kernel_path(self, dev) -> bool:
        if not self.use_kernels or dev.type != "cuda" or self.model.entity_dim % 4:
            return False
        if self.shard_ptrs is None and self.world == 1:
            self.shard_ptrs = (C.c_void_p * 1)(C.c_void_p(self.model.entity_embedding.data_ptr()))
        return self.shard_ptrs is not None

    def _compact_state(self, B, N, dev):
        m = self.model
        cap = min(m.nentity, B * N + 2 * B)
        key = (B, N, dev, cap)
        if self._cstate is None or self._cstate["key"] != key:
            need = _lib.load().kge_shard_compact_workspace_bytes(m.nentity)
            self._cstate = {
                "key": key, "cap": cap,
                "cpos": torch.empty((B, 3), dtype=torch.int64, device=dev),
                "cneg": torch.empty((B, N), dtype=torch.int64, device=dev),
                "uniq": torch.empty(cap, dtype=torch.int64, device=dev),
                "n": torch.zeros(1, dtype=torch.int32, device=dev),
                "table": torch.empty((cap, m.entity_dim), dtype=torch.float32, device=dev),
                "ws": torch.empty(need, dtype=torch.uint8, device=dev)}
        return self._cstate

    def _sweep_kernels(self, mode, pos, neg, w, wsum, B_total, loss_opts):
        m, lib = self.model, _lib.load()
        B, N = neg.shape
        cs = self._compact_state(B, N, pos.device)
        P = ops._ptr
        nshards = len(self.shard_ptrs)
        rows_per = (m.nentity + nshards - 1) // nshards  # shard_range: contiguous blocks of ceil(E / world) rows
        if nshards == 1 and self._shard_hdl is None:  # (the parameter may have been re-allocated: .to(), load_state_dict)
            self.shard_ptrs[0] = C.c_void_p(m.entity_embedding.data_ptr())
        with _on(pos.device):
            stream = ops._stream()
            _lib.check(lib.kge_shard_compact(P(pos), P(neg), B, N, m.nentity, P(cs["cpos"]), P(cs["cneg"]), P(cs["uniq"]),
                                             cs["cap"], P(cs["n"]), P(cs["ws"]), cs["ws"].numel(), stream), "kge_shard_compact")
            _lib.check(lib.kge_shard_fetch_rows(P(cs["uniq"]), P(cs["n"]), cs["cap"], m.entity_dim, self.shard_ptrs,
                                                nshards, rows_per, P(cs["table"]), stream), "kge_shard_fetch_rows")
        self._run_sweep(mode, cs["table"], cs["cpos"], cs["cneg"], w, wsum, B_total, loss_opts)

    def rows_touched(self) -> int:
        """Unique rows of the last local batch (reads the device counter: a sync -- for reports, not for the step)."""
        return int(self._cstate["n"].item()) if self._cstate is not None else self.last_exchange.get("rows_requested", 0)

    # ---- 1: rows of the local batch from their owners ------------------------------------------------------
    def fetch_rows(self, ids_sorted: torch.Tensor) -> torch.Tensor:
        """``ids_sorted``: sorted unique global ids this rank needs -> their rows ``[U, De]`` (same order)."""
        m = self.model
        ent = m.entity_embedding.detach()
        dev = ids_sorted.device
        if self.world == 1 or not dist.is_initialized():
            return ent[ids_sorted - self.lo]
        rows_per = (m.nentity + self.world - 1) // self.world
        bounds = torch.arange(1, self.world, device=dev, dtype=torch.int64) * rows_per
        cut = torch.searchsorted(ids_sorted, bounds)  # owners are contiguous id ranges and the ids are sorted
        req_counts = torch.diff(torch.cat([cut.new_zeros(1), cut, cut.new_full((1,), ids_sorted.numel())]))
        got_counts = torch.empty_like(req_counts)
        dist.all_to_all_single(got_counts, req_counts, group=self.group)
        req_l, got_l = req_counts.tolist(), got_counts.tolist()  # host sync: the split sizes of the next two calls
        asked = torch.empty(sum(got_l), dtype=torch.int64, device=dev)
        dist.all_to_all_single(asked, ids_sorted, got_l, req_l, group=self.group)
        served = ent[asked - self.lo]  # every asked id is ours
        rows = torch.empty((ids_sorted.numel(), m.entity_dim), dtype=torch.float32, device=dev)
        dist.all_to_all_single(rows, served, req_l, got_l, group=self.group)
        self.last_exchange = {"rows_requested": int(ids_sorted.numel()), "rows_served": int(asked.numel())}
        return rows

    def _sweep(self, mode, pos, neg, w, wsum, B_total, loss_opts):
        m = self.model
        B, N = neg.shape
        if self._kernel_path(pos.device):
            return self._sweep_kernels(mode, pos, neg, w, wsum, B_total, loss_opts)
        ids = torch.cat([pos[:, 0], pos[:, 2], neg.reshape(-1)])
        uniq, inv = torch.unique(ids, sorted=True, return_inverse=True)
        compact = self.fetch_rows(uniq).contiguous()
        cpos = torch.stack([inv[:B], pos[:, 1], inv[B:2 * B]], dim=1).contiguous()
        cneg = inv[2 * B:].view(B, N).contiguous()
        self._run_sweep(mode, compact, cpos, cneg, w, wsum, B_total, loss_opts)

    def _run_sweep(self, mode, compact, cpos, cneg, w, wsum, B_total, loss_opts):
        m = self.model
        B, N = cneg.shape
        desc = m._desc()
        desc.nentity = compact.shape[0]
        bufs = StepBuffers(*[C.c_void_p(self._local[k].data_ptr()) for k in _BUFFERS])
        sws = self._sweep_ws(desc, B, N, compact.device)
        with _on(compact.device):
            _lib.check(_lib.load().kge_train_sweep(
                C.byref(desc), ops.mode_id(mode), ops._ptr(compact), ops._ptr(m.relation_embedding),
                ops._ptr(m._modulus()), ops._ptr(cpos), ops._ptr(cneg), ops._ptr(w), B, N, C.byref(loss_opts),
                ops._ptr(wsum), B_total, C.byref(bufs), ops._ptr(sws), sws.numel(), ops._stream()), "kge_train_sweep")

    # ---- 4: owners gather gradients and apply Adam on their shard ---------------------------------------------
    def _apply(self, mode, pos_all, neg_all, loss_opts, adam, out):
        m, opt = self.model, self.optimizer
        B_all, N = neg_all.shape
        ent, rel, mod = m.entity_embedding, m.relation_embedding, m._modulus()
        st_e, st_r = opt.state[ent], opt.state[rel]
        st_m = opt.state[mod] if mod is not None else None
        desc = m._desc()
        need = ops.workspace_bytes(_lib.WS_TRAIN_APPLY, desc, B_all, N)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=ent.device)
        bufs = StepBuffers(*[C.c_void_p(self._gathered[k].data_ptr()) for k in _BUFFERS])
        P = ops._ptr
        with _on(ent.device):
            _lib.check(_lib.load().kge_train_apply_shard(
                C.byref(desc), ops.mode_id(mode), P(ent), self.lo, ent.shape[0], P(rel), P(mod), P(st_e["exp_avg"]),
                P(st_e["exp_avg_sq"]), P(st_r["exp_avg"]), P(st_r["exp_avg_sq"]), P(st_m["exp_avg"]) if st_m else None,
                P(st_m["exp_avg_sq"]) if st_m else None, P(pos_all), P(neg_all), B_all, N, C.byref(loss_opts),
                C.byref(adam), C.byref(bufs), self.lo, self.hi, P(out), C.c_void_p(out.data_ptr() + 12), P(self._ws),
                self._ws.numel(), ops._stream()), "kge_train_apply_shard")

    def _share_rows(self):
        """Rows stay with their owners: nothing to broadcast.  With the shards in symmetric memory one barrier closes the
        step: every owner's update is complete before anybody pulls rows for the next batch."""
        if self._shard_hdl is not None:
            self._shard_hdl.barrier()
