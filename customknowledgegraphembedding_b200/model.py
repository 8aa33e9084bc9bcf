"""``KGEModel`` -- drop-in for the PyTorch ``codes/model.py::KGEModel`` of the KnowledgeGraphEmbedding
codes the reference wraps (absent submodule @2e440e0f), extended with the
reference's own InterHT score function and ``-tr`` flag (/root/reference/tensorflow_codes/model.py:49-112,
207-224).  Same constructor, ``forward(sample, mode)``, static ``train_step`` / ``test_step`` and
parameter names (``entity_embedding``, ``relation_embedding``, ``gamma``, ``embedding_range``,
``modulus`` -- the naming contract of /root/reference/Checker/copy_weights.py:9-10).

All arithmetic runs in the hand-written sm_100a kernels of ``libkge_b200.so``; torch provides device
memory, streams and the autograd graph edge.  There is no CPU path: a model left on the CPU raises.
"""
from __future__ import annotations

import contextlib

import ctypes as C
import logging
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn as nn

from . import _lib, ops
from ._lib import AdamOpts, LossOpts

MODEL_NAMES = ("TransE", "DistMult", "ComplEx", "RotatE", "pRotatE", "InterHT")


class _ScoreFunction(torch.autograd.Function):
    """scores = K1(entity, relation, modulus; sample, mode);  backward = K3 into dense gradients."""

    @staticmethod
    def forward(ctx, entity, relation, modulus, positive, negative, desc, mode, deterministic):
        mod = modulus if modulus is not None and modulus.numel() else None
        scores = ops.score_forward(desc, mode, entity.detach(), relation.detach(),
                                   None if mod is None else mod.detach(), positive, negative)
        ctx.save_for_backward(entity, relation, modulus, positive, negative)
        ctx.desc, ctx.mode, ctx.deterministic, ctx.has_mod = desc, mode, deterministic, mod is not None
        return scores

    @staticmethod
    def backward(ctx, dscore):
        entity, relation, modulus, positive, negative = ctx.saved_tensors
        g_ent = torch.zeros_like(entity)
        g_rel = torch.zeros_like(relation)
        g_mod = torch.zeros_like(modulus) if ctx.has_mod else None
        ops.score_backward(ctx.desc, ctx.mode, entity.detach(), relation.detach(),
                           modulus.detach() if ctx.has_mod else None, positive, negative,
                           dscore.contiguous(), g_ent, g_rel, g_mod, deterministic=ctx.deterministic)
        return g_ent, g_rel, g_mod, None, None, None, None, None


class KGEModel(nn.Module):
    def __init__(self, model_name, nentity, nrelation, hidden_dim, gamma,
                 double_entity_embedding=False, double_relation_embedding=False,
                 triple_relation_embedding=False):
        super().__init__()
        self.model_name = model_name
        self.nentity = nentity
        self.nrelation = nrelation
        self.hidden_dim = hidden_dim
        self.epsilon = 2.0

        self.gamma = nn.Parameter(torch.Tensor([gamma]), requires_grad=False)
        self.embedding_range = nn.Parameter(
            torch.Tensor([(self.gamma.item() + self.epsilon) / hidden_dim]), requires_grad=False)

        self.entity_dim = hidden_dim * 2 if double_entity_embedding else hidden_dim
        self.relation_dim = hidden_dim * 2 if double_relation_embedding else hidden_dim
        if triple_relation_embedding:  # tensorflow_codes/model.py:75-78
            self.relation_dim = hidden_dim * 3

        self.entity_embedding = nn.Parameter(torch.zeros(nentity, self.entity_dim))
        nn.init.uniform_(tensor=self.entity_embedding, a=-self.embedding_range.item(), b=self.embedding_range.item())
        self.relation_embedding = nn.Parameter(torch.zeros(nrelation, self.relation_dim))
        nn.init.uniform_(tensor=self.relation_embedding, a=-self.embedding_range.item(), b=self.embedding_range.item())

        if model_name == "pRotatE":
            self.modulus = nn.Parameter(torch.Tensor([[0.5 * self.embedding_range.item()]]))

        # Do not forget to modify this line when you add a new model in the "forward" function
        if model_name not in MODEL_NAMES:
            raise ValueError("model %s not supported" % model_name)
        if model_name == "RotatE" and (not double_entity_embedding or double_relation_embedding):
            raise ValueError("RotatE should use --double_entity_embedding")
        if model_name == "ComplEx" and (not double_entity_embedding or not double_relation_embedding):
            raise ValueError("ComplEx should use --double_entity_embedding and --double_relation_embedding")
        if model_name == "InterHT" and (not double_entity_embedding or not triple_relation_embedding):
            raise ValueError("InterHT should use --double_entity_embedding and --triple_relation_embedding")

        # B200 knobs (not in the reference API)
        self.deterministic = False  # fixed-order backward (exact-check runs)
        self.interht_eps = 0.0      # 0: tf.linalg.normalize; 1e-12: F.normalize
        self.check_ids = False      # range-check every batch's ids (IndexError like index_select); synchronises
        self._ws: Dict[tuple, torch.Tensor] = {}
        self._scalars: Dict[str, tuple] = {}
        self._plan_key, self._plan_ok = None, False
        self._filter_cache = None
        self.last_eval = {}
        # train_step input lookahead: while the GPU runs step k, batch k+1 is taken from the iterator and copied to the
        # device on a side stream (see _fused_step).  Set to False for the strictly sequential upstream behaviour.
        self.input_lookahead = True
        self._ahead = {}     # id(iterator) -> (iterator, batch, staged device copy or None); holding the iterator pins its id
        self._stage = None   # copy stream and the pool of staging buffer sets
        self._hot = None     # argument cache of the staged train step (pointers, structs, step counters)
        self._eval_streams = None  # two side streams: the head-batch and tail-batch rankings of a chunk run concurrently

    def __getstate__(self):
        # streams, events, staged batches and workspaces are per-process scratch: never pickled / deep-copied
        d = self.__dict__.copy()
        d.update(_ahead={}, _stage=None, _ws={}, _hot=None, _eval_streams=None)
        return d

    # ------------------------------------------------------------------------------------
    def _scalar(self, name: str) -> float:
        """gamma / embedding_range as Python floats without a device sync per call: `.item()` on a CUDA parameter
        waits for everything queued on the stream, which would serialise the host and the GPU every step.  The cached
        value is refreshed when the parameter is replaced (`.to(device)`) or written in place (`load_state_dict`)."""
        p = getattr(self, name)
        key = (id(p), p.data_ptr(), p._version)
        hit = self._scalars.get(name)
        if hit is None or hit[0] != key:
            hit = (key, float(p.item()))
            self._scalars[name] = hit
        return hit[1]

    def _desc(self):
        return ops.make_desc(self.model_name, self.nentity, self.nrelation, self.hidden_dim, self.entity_dim,
                             self.relation_dim, self._scalar("gamma"), self._scalar("embedding_range"), self.interht_eps)

    def _modulus(self) -> Optional[torch.Tensor]:
        return self.modulus if self.model_name == "pRotatE" else None

    def _workspace(self, kind: int, B: int, N: int) -> torch.Tensor:
        dev = self.entity_embedding.device
        need = ops.workspace_bytes(kind, self._desc(), B, N)
        key = (kind, dev)
        ws = self._ws.get(key)
        if ws is None or ws.numel() < need:
            ws = torch.empty(need, dtype=torch.uint8, device=dev)
            self._ws[key] = ws
        return ws

    def _check_batch_ids(self, positive, negative):
        """``check_ids``: ids against the table sizes -- on the host for host tensors (no copy, no sync), by
        ``kge_check_ids`` for device tensors (synchronises the stream)."""
        if positive.is_cuda:
            ops.check_ids(positive, negative, self.nentity, self.nrelation)
        else:
            from .wire import check_id_ranges
            check_id_ranges(positive, negative if negative is not None else positive[:, :1], self.nentity, self.nrelation)

    def forward(self, sample, mode="single"):
        """Raw scores ``[B,1]`` ('single') or ``[B,N]`` ('head-batch' / 'tail-batch') -- U2."""
        if mode == "single":
            positive, negative = sample, None
        elif mode in ("head-batch", "tail-batch"):
            positive, negative = sample
        else:
            raise ValueError("mode %s not supported" % mode)
        dev = self.entity_embedding.device
        if dev.type != "cuda":
            raise RuntimeError("KGEModel is on %s: the B200 kernels need the model on a CUDA device "
                               "(no CPU fallback)" % dev)
        if self.check_ids:
            self._check_batch_ids(positive, negative)
        positive = positive.to(dev, non_blocking=True)
        if negative is not None:
            negative = negative.to(dev, non_blocking=True)
        mod = self._modulus()
        return _ScoreFunction.apply(self.entity_embedding, self.relation_embedding,
                                    mod if mod is not None else torch.empty(0, device=dev),
                                    positive.contiguous(), None if negative is None else negative.contiguous(),
                                    self._desc(), mode, self.deterministic)

    # ------------------------------------------------------------------------------------
    # train_step (U9; loss body restated by the reference at tensorflow_codes/model.py:4-44)
    # ------------------------------------------------------------------------------------
    def _fused_adam_plan(self, optimizer):
        """The fused K1-K4 step applies when `optimizer` is plain Adam over exactly this model's trainable
        parameters (what the upstream run.py builds).  Returns the param group or None."""
        from .optim import FusedAdam
        if not isinstance(optimizer, (torch.optim.Adam, FusedAdam)) or len(optimizer.param_groups) != 1:
            return None
        g = optimizer.param_groups[0]
        if g.get("amsgrad") or g.get("weight_decay", 0) != 0 or g.get("maximize") or g.get("differentiable"):
            return None
        key = (id(optimizer), id(g), len(g["params"]))  # the parameter-identity walk is done once per optimizer
        if self._plan_key != key:
            mine = [p for p in self.parameters() if p.requires_grad]
            theirs = [p for p in g["params"] if p.requires_grad]  # Adam skips parameters that never get a gradient
            same = len(theirs) == len(mine) and all(a is b for a, b in zip(sorted(theirs, key=id), sorted(mine, key=id)))
            self._plan_key, self._plan_ok = key, same
        return g if self._plan_ok else None

    @staticmethod
    def _adam_state(optimizer, p):
        st = optimizer.state[p]
        if len(st) == 0:  # same layout torch.optim.Adam creates lazily
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @staticmethod
    def train_step(model, optimizer, train_iterator, args):
        """A single train step: apply back-propagation and return the loss log (same keys as upstream)."""
        model.train()
        ahead, staged = model._ahead.pop(id(train_iterator), None) if model._ahead else None, None
        if ahead is not None:  # taken from this iterator during its previous step
            (positive_sample, negative_sample, subsampling_weight, mode), staged = ahead[1], ahead[2]
        else:
            positive_sample, negative_sample, subsampling_weight, mode = next(train_iterator)
        if mode not in ("head-batch", "tail-batch"):
            raise ValueError("mode %s not supported" % mode)
        if model.check_ids:
            model._check_batch_ids(positive_sample, negative_sample)
        adv = bool(getattr(args, "negative_adversarial_sampling", False))
        alpha = float(getattr(args, "adversarial_temperature", 1.0))
        uni = bool(getattr(args, "uni_weight", False))
        reg = float(getattr(args, "regularization", 0.0))
        dev = model.entity_embedding.device
        if dev.type != "cuda":
            raise RuntimeError("KGEModel.train_step: model is on %s; move it to a CUDA device (no CPU fallback)" % dev)

        group = model._fused_adam_plan(optimizer)
        if group is None:
            return model._train_step_generic(optimizer, positive_sample, negative_sample, subsampling_weight, mode,
                                             adv, alpha, uni, reg)

        if (model.input_lookahead and not positive_sample.is_cuda and torch.cuda.current_device() == dev.index):
            out = model._fused_step_staged(optimizer, group, positive_sample, negative_sample, subsampling_weight, mode,
                                           adv, alpha, uni, reg, staged, train_iterator)
        else:
            out = model._fused_step(optimizer, group, positive_sample, negative_sample, subsampling_weight, mode,
                                    adv, alpha, uni, reg)
        pos_l, neg_l, loss, regv = out.tolist() if isinstance(out, torch.Tensor) else (float(x) for x in out)
        log = {"regularization": regv} if reg != 0.0 else {}
        log.update({"positive_sample_loss": pos_l, "negative_sample_loss": neg_l, "loss": loss})
        return log

    # ---- input lookahead -------------------------------------------------------------------------
    def _stage_host_batch(self, lib, positive_sample, negative_sample, subsampling_weight, uni, dev):
        """H2D copies of one host batch on the copy stream into a free set of staging buffers (kge_stage_batch);
        returns (pos, neg, w or None, event, host sources).  A set goes back to the pool when the step that read it
        has had its loss read back (a stream synchronisation): one set per batch in flight plus the one in use."""
        pos = positive_sample.contiguous()
        neg = negative_sample.contiguous()
        if pos.dtype != torch.int64 or neg.dtype != torch.int64:
            raise TypeError("positive_sample / negative_sample must be int64")
        w = None if uni else subsampling_weight.to(torch.float32).contiguous()
        B, N = neg.shape
        st = self._stage
        key = (B, N, dev, bool(uni))
        if st is None or st["dev"] != dev:
            stream = torch.cuda.Stream(dev)
            st = self._stage = {"dev": dev, "stream": stream, "stream_ptr": C.c_void_p(stream.cuda_stream), "key": key, "free": []}
        if st["key"] != key:
            st["key"], st["free"] = key, []
        if st["free"]:
            dpos, dneg, dw, ev = st["free"].pop()[:4]
        else:
            dpos = torch.empty((B, 3), dtype=torch.int64, device=dev)
            dneg = torch.empty((B, N), dtype=torch.int64, device=dev)
            dw = None if uni else torch.empty(B, dtype=torch.float32, device=dev)
            ev = torch.cuda.Event()
        _lib.check(lib.kge_stage_batch(C.c_void_p(pos.data_ptr()), C.c_void_p(neg.data_ptr()),
                                       None if w is None else C.c_void_p(w.data_ptr()), ops._ptr(dpos), ops._ptr(dneg),
                                       ops._ptr(dw), B, N, st["stream_ptr"]), "kge_stage_batch")
        ev.record(st["stream"])
        # (the host sources stay alive with the tuple; the raw event handle exists once the event has been recorded)
        return dpos, dneg, dw, ev, (pos, neg, w), key, C.c_void_p(ev.cuda_event)

    def _take_ahead(self, lib, iterator, uni, dev):
        """``next(iterator)`` one step early, its H2D copies queued behind nothing but themselves."""
        try:
            batch = next(iterator)
        except StopIteration:
            return  # an exhausted iterator raises again on the next call's own next(): nothing to remember
        staged = None
        if (isinstance(batch[0], torch.Tensor) and not batch[0].is_cuda and batch[3] in ("head-batch", "tail-batch")):
            staged = self._stage_host_batch(lib, batch[0], batch[1], batch[2], uni, dev)
        self._ahead[id(iterator)] = (iterator, batch, staged)

    def _fused_step_staged(self, optimizer, group, positive_sample, negative_sample, subsampling_weight, mode,
                           adv, alpha, uni, reg, staged, iterator):
        """The host-batch step of ``train_step`` with the input lookahead: wait for the staged copy of THIS batch, queue
        the step (kge_train_step), take the NEXT batch from the iterator and start its H2D copies on the copy stream
        while the GPU works, then read the loss back (kge_fetch_floats).  Between the end of one step and the first
        kernel of the next the GPU idles for exactly the Python that runs here before the C call, so everything that does
        not change from step to step (pointers, structs, workspace) is cached in ``_hot`` and revalidated by address."""
        lib = _lib.load()
        ent, rel, mod = self.entity_embedding, self.relation_embedding, self._modulus()
        dev = ent.device
        st_e, st_r = self._adam_state(optimizer, ent), self._adam_state(optimizer, rel)
        st_m = self._adam_state(optimizer, mod) if mod is not None else None
        key = (id(optimizer), ent.data_ptr(), rel.data_ptr(), st_e["exp_avg"].data_ptr(), st_e["exp_avg_sq"].data_ptr(),
               st_r["exp_avg"].data_ptr(), st_r["exp_avg_sq"].data_ptr(), id(st_e["step"]), id(st_r["step"]),
               0 if mod is None else mod.data_ptr(), 0 if st_m is None else st_m["exp_avg"].data_ptr())
        hot = self._hot
        if hot is None or hot["key"] != key:
            P = ops._ptr
            steps = [st["step"] for st in (st_e, st_r, st_m) if st is not None]
            if any(t.is_cuda for t in steps):  # capturable optimizers keep `step` on the device: general path
                return self._fused_step(optimizer, group, positive_sample, negative_sample, subsampling_weight, mode,
                                        adv, alpha, uni, reg)
            adam = AdamOpts(0.0, 0.9, 0.999, 1e-8, 1, 0)
            host = torch.empty(4, dtype=torch.float32).pin_memory()
            out_dev = torch.empty(4, dtype=torch.float32, device=dev)
            hot = self._hot = {
                "key": key, "steps": [t.numpy() for t in steps], "keep": (steps, host, out_dev),
                "tables": (P(ent), P(rel), P(mod), P(st_e["exp_avg"]), P(st_e["exp_avg_sq"]), P(st_r["exp_avg"]),
                           P(st_r["exp_avg_sq"]), P(st_m["exp_avg"]) if st_m else None, P(st_m["exp_avg_sq"]) if st_m else None),
                "adam": adam, "adam_ref": C.byref(adam), "loss": {}, "ws": {},
                "out": host.numpy(), "out_host": C.c_void_p(host.data_ptr()), "out_dev": P(out_dev),
                "out_dev3": C.c_void_p(out_dev.data_ptr() + 12)}
        for v in hot["steps"]:
            v += 1  # the optimizer's own `step` tensors (0-d views), as torch.optim.Adam.step() would
        adam = hot["adam"]
        adam.lr, (adam.beta1, adam.beta2), adam.eps, adam.step = group["lr"], group["betas"], group["eps"], int(hot["steps"][0])
        lkey = (adv, uni, alpha, reg)
        lo = hot["loss"].get(lkey)
        if lo is None:
            lo = LossOpts(int(adv), int(uni), 1, 0, alpha, reg)
            lo = hot["loss"][lkey] = (lo, C.byref(lo))
        desc = self._desc()
        B, N = negative_sample.shape
        wkey = (B, N, desc.gamma, desc.embedding_range, desc.interht_eps)
        ws = hot["ws"].get(wkey)
        if ws is None:
            t = self._workspace(_lib.WS_TRAIN_STEP, B, N)
            ws = hot["ws"][wkey] = (t, ops._ptr(t), t.numel())
        if staged is None or (staged[2] is None) != bool(uni) or staged[0].device != dev:
            staged = self._stage_host_batch(lib, positive_sample, negative_sample, subsampling_weight, uni, dev)
        stream = C.c_void_p(torch._C._cuda_getCurrentRawStream(dev.index))
        _lib.check(lib.kge_stream_wait_event(stream, staged[6]), "kge_stream_wait_event")
        _lib.check(lib.kge_train_step(
            C.byref(desc), ops.mode_id(mode), *hot["tables"], ops._ptr(staged[0]), ops._ptr(staged[1]), ops._ptr(staged[2]),
            B, N, lo[1], hot["adam_ref"], int(self.deterministic), hot["out_dev"], hot["out_dev3"], ws[1], ws[2], stream),
            "kge_train_step")
        if iterator is not None:
            self._take_ahead(lib, iterator, uni, dev)  # host work + H2D of batch k+1 under step k
        _lib.check(lib.kge_fetch_floats(hot["out_host"], hot["out_dev"], 4, stream), "kge_fetch_floats")
        if self._stage is not None and self._stage["key"] == staged[5]:
            self._stage["free"].append(staged)  # the step is over: its staging set can take a new batch
        return hot["out"]

    def _fused_step(self, optimizer, group, positive_sample, negative_sample, subsampling_weight, mode,
                    adv, alpha, uni, reg, detach_weights=True, keras_adam=False):
        """One call into kge_train_step(_host).  Batch on the host -> copies inside the C call, returns a numpy
        [pos_loss, neg_loss, loss, reg]; batch on the device -> fully asynchronous, returns a device tensor."""
        model = self
        lib = _lib.load()
        dev = model.entity_embedding.device
        B, N = negative_sample.shape
        ent, rel, mod = model.entity_embedding, model.relation_embedding, model._modulus()
        st_e, st_r = model._adam_state(optimizer, ent), model._adam_state(optimizer, rel)
        st_m = model._adam_state(optimizer, mod) if mod is not None else None
        states = [s for s in (st_e, st_r, st_m) if s is not None]
        for s in states:
            s["step"] += 1
        step = int(st_e["step"])
        beta1, beta2 = group["betas"]
        adam = AdamOpts(float(group["lr"]), float(beta1), float(beta2), float(group["eps"]), step, int(keras_adam))
        loss_opts = LossOpts(int(adv), int(uni), int(detach_weights), 0, alpha, reg)
        desc = model._desc()
        host_batch = not positive_sample.is_cuda
        P = ops._ptr
        common_state = (P(st_e["exp_avg"]), P(st_e["exp_avg_sq"]), P(st_r["exp_avg"]), P(st_r["exp_avg_sq"]),
                        P(st_m["exp_avg"]) if st_m else None, P(st_m["exp_avg_sq"]) if st_m else None)
        # (the device guard costs ~10 us of host time: only taken when the model is not on the current device)
        with (torch.cuda.device(dev) if torch.cuda.current_device() != dev.index else contextlib.nullcontext()):
            stream = ops._stream()
            if host_batch:
                pos = positive_sample.contiguous()
                neg = negative_sample.contiguous()
                w = None if uni else subsampling_weight.to(torch.float32).contiguous()
                if pos.dtype != torch.int64 or neg.dtype != torch.int64:
                    raise TypeError("positive_sample / negative_sample must be int64")
                ws = model._workspace(_lib.WS_TRAIN_STEP_HOST, B, N)
                out = np.empty(4, dtype=np.float32)
                _lib.check(lib.kge_train_step_host(
                    C.byref(desc), ops.mode_id(mode), P(ent), P(rel), P(mod), *common_state,
                    C.c_void_p(pos.data_ptr()), C.c_void_p(neg.data_ptr()),
                    None if w is None else C.c_void_p(w.data_ptr()), B, N, C.byref(loss_opts), C.byref(adam),
                    int(model.deterministic), out.ctypes.data_as(C.c_void_p), P(ws), ws.numel(), stream),
                    "kge_train_step_host")
                return out
            pos = ops._need_cuda("positive_sample", positive_sample, torch.int64)
            neg = ops._need_cuda("negative_sample", negative_sample, torch.int64)
            w = None if uni else ops._need_cuda("subsampling_weight", subsampling_weight.to(dev), torch.float32)
            ws = model._workspace(_lib.WS_TRAIN_STEP, B, N)
            out = torch.empty(4, dtype=torch.float32, device=dev)  # [pos, neg, loss, reg]: all four are written
            _lib.check(lib.kge_train_step(
                C.byref(desc), ops.mode_id(mode), P(ent), P(rel), P(mod), *common_state, P(pos), P(neg), P(w),
                B, N, C.byref(loss_opts), C.byref(adam), int(model.deterministic), P(out),
                C.c_void_p(out.data_ptr() + 12), P(ws), ws.numel(), stream), "kge_train_step")
            return out

    def _train_step_generic(self, optimizer, positive_sample, negative_sample, subsampling_weight, mode,
                            adv, alpha, uni, reg):
        """Any other optimizer: K1 forward, K2 loss, K3 backward into .grad, then optimizer.step()."""
        dev = self.entity_embedding.device
        optimizer.zero_grad()
        positive_sample = positive_sample.to(dev)
        negative_sample = negative_sample.to(dev)
        subsampling_weight = subsampling_weight.to(dev)
        negative_score = self((positive_sample, negative_sample), mode=mode)
        positive_score = self(positive_sample)
        loss3, dneg, dpos = ops.loss_forward_backward(
            negative_score.detach(), positive_score.detach(), subsampling_weight,
            negative_adversarial_sampling=adv, adversarial_temperature=alpha, uni_weight=uni)
        torch.autograd.backward([negative_score, positive_score], [dneg, dpos.view(-1, 1)])
        pos_l, neg_l, loss = loss3.tolist()
        log = {}
        if reg != 0.0:
            regv = reg * (ops.l3_sum_cubes(self.entity_embedding.detach())
                          + ops.l3_sum_cubes(self.relation_embedding.detach())).item()
            with torch.no_grad():
                for p in (self.entity_embedding, self.relation_embedding):
                    p.grad.add_(3.0 * reg * p * p.abs())
            loss += regv
            log["regularization"] = regv
        optimizer.step()
        log.update({"positive_sample_loss": pos_l, "negative_sample_loss": neg_l, "loss": loss})
        return log

    # ------------------------------------------------------------------------------------
    # test_step (U10): filtered full-entity ranking, rank by counting instead of argsort
    # ------------------------------------------------------------------------------------
    def _filter_maps(self, all_true_triples):
        """Sorted (r,t)->heads and (h,r)->tails tables over the unique true triples, built once per list."""
        key = (id(all_true_triples), len(all_true_triples))
        if self._filter_cache is not None and self._filter_cache[0] == key:
            return self._filter_cache[1]
        A = np.unique(np.asarray(all_true_triples, dtype=np.int64).reshape(-1, 3), axis=0)
        E = np.int64(self.nentity)
        maps = {}
        for mode, kcols, vcol in (("head-batch", (1, 2), 0), ("tail-batch", (1, 0), 2)):
            k = A[:, kcols[0]] * E + A[:, kcols[1]]
            order = np.argsort(k, kind="stable")
            maps[mode] = (k[order], A[order, vcol])
        self._filter_cache = (key, maps)
        return maps

    @staticmethod
    def build_filter_csr(maps, queries: np.ndarray, mode: str, nentity: int):
        """CSR (indptr [Q+1], idx) of the candidate ids upstream's TestDataset gives bias -1 for each query."""
        keys, vals = maps[mode]
        E = np.int64(nentity)
        other = queries[:, 2] if mode == "head-batch" else queries[:, 0]
        qk = queries[:, 1] * E + other
        lo = np.searchsorted(keys, qk, side="left")
        hi = np.searchsorted(keys, qk, side="right")
        lens = hi - lo
        indptr = np.zeros(len(queries) + 1, dtype=np.int64)
        np.cumsum(lens, out=indptr[1:])
        total = int(indptr[-1])
        if total == 0:
            return indptr, np.zeros(0, dtype=np.int64)
        pos_in = np.arange(total, dtype=np.int64) - np.repeat(indptr[:-1], lens)
        idx = vals[np.repeat(lo, lens) + pos_in]
        return indptr, idx.astype(np.int64, copy=False)

    @torch.no_grad()
    def rank_queries(self, queries: torch.Tensor, mode: str, filter_indptr=None, filter_idx=None,
                     entity_lo: int = 0, entity_hi: Optional[int] = None):
        """(true_score, n_greater, n_equal) for a [Q,3] int64 query tensor (any device) -- thin K5 call."""
        dev = self.entity_embedding.device
        entity_hi = self.nentity if entity_hi is None else entity_hi
        to = lambda t: None if t is None else t.to(dev, non_blocking=True)
        return ops.eval_rank(self._desc(), mode, self.entity_embedding.detach(), self.relation_embedding.detach(),
                             None if self._modulus() is None else self._modulus().detach(), to(queries),
                             entity_lo, entity_hi, to(filter_indptr), to(filter_idx))

    @torch.no_grad()
    def rank_queries_both_modes(self, queries: torch.Tensor, filters: Dict[str, tuple]):
        """``{mode: rank_queries(queries, mode, *filters[mode])}`` for head-batch and tail-batch with the two rankings on
        two side streams: the tile kernel holds one CTA per SM and a ranking of Q x E is a whole number of 148-CTA waves
        plus a partial one (FB15k-237, 1024 queries: 6.16) -- run together, the partial wave of one mode fills with tiles
        of the other, and the small staging / true-score / filter kernels of one hide under the tiles of the other."""
        dev = self.entity_embedding.device
        cur = torch.cuda.current_stream(dev)
        if self._eval_streams is None or self._eval_streams[0].device != dev:
            self._eval_streams = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        queries = queries.to(dev, non_blocking=True)
        filters = {m: tuple(None if t is None else t.to(dev, non_blocking=True) for t in f) for m, f in filters.items()}
        start = torch.cuda.Event()
        start.record(cur)
        out = {}
        for side, mode in zip(self._eval_streams, ("head-batch", "tail-batch")):
            side.wait_event(start)
            with torch.cuda.stream(side):
                out[mode] = self.rank_queries(queries, mode, *filters[mode])
        for side, mode in zip(self._eval_streams, ("head-batch", "tail-batch")):
            cur.wait_stream(side)
            for t in out[mode]:
                t.record_stream(cur)  # produced on the side stream, consumed on the caller's
        return out

    @staticmethod
    def test_step(model, test_triples, all_true_triples, args, rank_fn=None):
        """Evaluate the model on test or valid datasets: filtered MRR / MR / HITS@1,3,10 over
        head-batch then tail-batch queries (same query order and metric arithmetic as upstream).
        ``rank_fn(queries, mode, filter_indptr, filter_idx)`` replaces the local K5 call (dist.test_step)."""
        model.eval()
        dev = model.entity_embedding.device
        if getattr(args, "countries", False):
            # Countries S* datasets are evaluated with AUC-PR, which is calculated using scikit-learn (upstream branch):
            # every test triple is scored against the regions as candidate tails -- K1 in 'single' mode.
            from sklearn.metrics import average_precision_score
            sample, y_true = [], []
            for head, relation, tail in test_triples:
                for candidate_region in args.regions:
                    y_true.append(1 if candidate_region == tail else 0)
                    sample.append((head, relation, candidate_region))
            sample = torch.LongTensor(sample).to(dev)
            with torch.no_grad():
                y_score = model(sample).squeeze(1).cpu().numpy()
            return {"auc_pr": average_precision_score(np.array(y_true), y_score)}
        if dev.type != "cuda" and rank_fn is None:
            raise RuntimeError("KGEModel.test_step: model is on %s; move it to a CUDA device (no CPU fallback)" % dev)
        nentity = getattr(args, "nentity", None) or model.nentity
        T = np.asarray(test_triples, dtype=np.int64).reshape(-1, 3)
        maps = model._filter_maps(all_true_triples)
        chunk = int(getattr(model, "eval_chunk", 4096))
        ranks, ties = {"head-batch": [], "tail-batch": []}, 0
        both = rank_fn is None and dev.type == "cuda" and getattr(model, "eval_two_streams", True)
        if rank_fn is None:
            rank_fn = model.rank_queries
        for lo in range(0, len(T), chunk):
            q = T[lo:lo + chunk]
            csr = {mode: tuple(torch.from_numpy(a) for a in KGEModel.build_filter_csr(maps, q, mode, nentity))
                   for mode in ("head-batch", "tail-batch")}
            if both:  # the two modes of a chunk side by side on two streams (same integers either way)
                res = model.rank_queries_both_modes(torch.from_numpy(q), csr)
            else:
                res = {mode: rank_fn(torch.from_numpy(q), mode, *csr[mode]) for mode in ("head-batch", "tail-batch")}
            for mode in ("head-batch", "tail-batch"):
                _, n_greater, n_equal = res[mode]
                ranks[mode].append((1 + n_greater).cpu().numpy())
                ties += int(n_equal.sum().item())
        # upstream's query order: every head-batch query, then every tail-batch query
        ranks = ranks["head-batch"] + ranks["tail-batch"]
        ranks = np.concatenate(ranks) if ranks else np.zeros(0, dtype=np.int64)
        model.last_eval = {"ranks": ranks, "n_tied_candidates": ties}
        if ties:
            logging.info("test_step: %d candidates tie exactly with a true score; ranks are 1 + #greater", ties)
        # upstream: per-query logs, then `sum([log[metric] for log in logs]) / len(logs)` -- the same builtin
        # sum over Python floats in query order, so equal ranks give bit-equal metrics
        logs = []
        for ranking in ranks.tolist():
            logs.append({
                "MRR": 1.0 / ranking,
                "MR": float(ranking),
                "HITS@1": 1.0 if ranking <= 1 else 0.0,
                "HITS@3": 1.0 if ranking <= 3 else 0.0,
                "HITS@10": 1.0 if ranking <= 10 else 0.0,
            })
        metrics = {}
        if logs:
            for metric in logs[0].keys():
                metrics[metric] = sum([log[metric] for log in logs]) / len(logs)
        return metrics
