"""TranSparse -- the second score function of the reference's ``TFKGEModel``
(/root/reference/tensorflow_codes/model.py:96-106 variables, :139-142,161-164,187-190 gathers, :226-235 score),
SURVEY.md 8(f) N4.

Every relation owns a projection ``W[r] [D, D]`` under a fixed 0/1 ``mask[r]``; with ``M = mask[r] * W[r]``

    score = gamma - || normalize(head @ M) * normalize(relation) - normalize(X @ M) ||_1

``project_tail=False`` is the reference AS WRITTEN: ``model.py:229`` projects ``head`` a second time (``X = head``), the
tail entity is never read and tail-batch scores are ``[B, 1]``.  ``project_tail=True`` uses ``X = tail``.

All arithmetic runs in ``libkge_b200.so`` (``csrc/kge_transparse.cu``: the projections as 3xTF32 tiles on tcgen05, one
grouped GEMM per pass over the relation-sorted rows); torch provides memory, streams and the autograd edge.  No CPU
path: a model left on the CPU raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.nn as nn

from . import _lib, ops
from ._lib import TranSparseDesc


class _Projection(torch.autograd.Function):
    """(positive_score [B], negative_score [B, nout]) = kge_transparse_forward; backward = kge_transparse_backward."""

    @staticmethod
    def forward(ctx, entity, relation, W, model, positive, negative, mode):
        need_grad = any(ctx.needs_input_grad[:3])  # (grad mode is off inside forward: ask the node)
        pos_s, neg_s, ws = model._forward_raw(positive, negative, mode, with_backward=need_grad)
        ctx.model, ctx.mode, ctx.ws, ctx.shape = model, mode, ws, (positive.shape[0], 0 if negative is None else negative.shape[1])
        ctx.save_for_backward(positive)
        ctx.mark_non_differentiable(*(() if need_grad else (pos_s, neg_s)))
        return pos_s, neg_s

    @staticmethod
    def backward(ctx, dpos, dneg):
        (positive,) = ctx.saved_tensors
        m = ctx.model
        B, N = ctx.shape
        nout = m.nout(ctx.mode, N)
        dev = positive.device
        dpos = torch.zeros(B, device=dev) if dpos is None else dpos.contiguous().float()
        dneg = torch.zeros(B, max(nout, 1), device=dev) if dneg is None else dneg.contiguous().float()
        g_ent = torch.empty_like(m.entity_embedding)
        g_rel = torch.empty_like(m.relation_embedding)
        g_W = torch.empty_like(m.W)
        m._backward_raw(positive, ctx.mode, B, N, dpos, dneg, g_ent, g_rel, g_W, ctx.ws)
        m._release(ctx.ws)
        ctx.ws = None
        return g_ent, g_rel, g_W, None, None, None, None


class TranSparseModel(nn.Module):
    def __init__(self, nentity, nrelation, hidden_dim, gamma, project_tail=False):
        super().__init__()
        self.model_name = "TranSparse"
        self.nentity, self.nrelation, self.hidden_dim = nentity, nrelation, hidden_dim
        self.entity_dim = self.relation_dim = hidden_dim  # model.py:227 multiplies [.., entity_dim] by [relation_dim, relation_dim]
        self.epsilon = 2.0
        self.project_tail = bool(project_tail)
        self.gamma = nn.Parameter(torch.Tensor([gamma]), requires_grad=False)
        rng = (float(gamma) + self.epsilon) / hidden_dim
        self.embedding_range = nn.Parameter(torch.Tensor([rng]), requires_grad=False)
        self.entity_embedding = nn.Parameter(torch.empty(nentity, hidden_dim).uniform_(-rng, rng))
        self.relation_embedding = nn.Parameter(torch.empty(nrelation, hidden_dim).uniform_(-rng, rng))
        self.W = nn.Parameter(torch.empty(nrelation, hidden_dim, hidden_dim).uniform_(-rng, rng))     # model.py:105-106
        prob = torch.empty(nrelation, hidden_dim, hidden_dim).uniform_(1, 100)                           # model.py:101
        self.register_buffer("mask", (prob >= 50).float())                                              # model.py:99-103
        self._gamma_py = float(gamma)
        self.check_ids = False  # range-check every batch's ids on the device (kge_check_ids; synchronises)
        self._free = {}      # (mode, B, N, with_backward) -> workspaces not in use
        self._grads = None   # persistent dense gradient buffers of the fused train steps

    # ---- C-ABI plumbing --------------------------------------------------------------------------
    def _desc(self) -> TranSparseDesc:
        return TranSparseDesc(self.nentity, self.nrelation, self.hidden_dim, self._gamma_py, int(self.project_tail))

    def nout(self, mode: str, N: int) -> int:
        return int(_lib.load().kge_transparse_nout(C.byref(self._desc()), ops.mode_id(mode), N))

    def _take(self, mode, B, N, with_backward):
        key = (mode, B, N, bool(with_backward))
        free = self._free.setdefault(key, [])
        if free:
            return free.pop()
        need = _lib.load().kge_transparse_workspace_bytes(C.byref(self._desc()), ops.mode_id(mode), B, max(N, 1), int(with_backward))
        if need == 0:
            raise ValueError("TranSparse: sizes not supported (B=%d, N=%d)" % (B, N))
        ws = torch.empty(need, dtype=torch.uint8, device=self.entity_embedding.device)
        ws._kge_key = key
        return ws

    def _release(self, ws):
        self._free.setdefault(ws._kge_key, []).append(ws)

    def _check(self, positive, negative, mode):
        dev = self.entity_embedding.device
        if dev.type != "cuda":
            raise RuntimeError("TranSparseModel: model is on %s; move it to a CUDA device (no CPU fallback)" % dev)
        positive = positive.to(dev, dtype=torch.int64).contiguous()
        if positive.dim() != 2 or positive.shape[1] != 3:
            raise ValueError("positive_sample must be [B, 3]")
        if mode != "single":
            if negative is None:
                raise ValueError("mode %s needs negative samples" % mode)
            negative = negative.to(dev, dtype=torch.int64).contiguous()
            if negative.dim() != 2 or negative.shape[0] != positive.shape[0]:
                raise ValueError("negative_sample must be [B, N]")
        else:
            negative = None
        if self.check_ids:
            ops.check_ids(positive, negative, self.nentity, self.nrelation)
        return positive, negative

    def _forward_raw(self, positive, negative, mode, with_backward):
        B = positive.shape[0]
        N = 0 if negative is None else negative.shape[1]
        dev = positive.device
        ws = self._take(mode, B, N, with_backward)
        pos_s = torch.empty(B, dtype=torch.float32, device=dev)
        neg_s = torch.empty(B, max(self.nout(mode, N), 0), dtype=torch.float32, device=dev)
        P = ops._ptr
        with torch.cuda.device(dev):
            _lib.check(_lib.load().kge_transparse_forward(
                C.byref(self._desc()), ops.mode_id(mode), P(self.entity_embedding), P(self.relation_embedding), P(self.W),
                P(self.mask), P(positive), P(negative), B, max(N, 1), P(pos_s), P(neg_s) if neg_s.numel() else None,
                int(with_backward), P(ws), ws.numel(), ops._stream()), "kge_transparse_forward")
        if not with_backward:
            self._release(ws)  # stream-ordered reuse: the next call on this stream runs after this one
            ws = None
        return pos_s, neg_s, ws

    def _backward_raw(self, positive, mode, B, N, dpos, dneg, g_ent, g_rel, g_W, ws):
        P = ops._ptr
        with torch.cuda.device(positive.device):
            _lib.check(_lib.load().kge_transparse_backward(
                C.byref(self._desc()), ops.mode_id(mode), P(self.relation_embedding), P(self.mask), P(positive), B, max(N, 1),
                P(dpos), P(dneg), P(g_ent), P(g_rel), P(g_W), P(ws), ws.numel(), ops._stream()), "kge_transparse_backward")

    # ---- the model API ---------------------------------------------------------------------------
    def both(self, sample, mode):
        """(positive_score [B, 1], negative_score [B, nout]) of one batch in ONE pass (one autograd node)."""
        positive, negative = self._check(sample[0], sample[1], mode)
        pos_s, neg_s = _Projection.apply(self.entity_embedding, self.relation_embedding, self.W, self, positive, negative, mode)
        return pos_s.view(-1, 1), neg_s

    def forward(self, sample, mode="single"):
        """``KGEModel.forward`` contract: ``sample [B,3]`` with mode 'single' -> ``[B,1]``; ``(positive [B,3],
        negative [B,N])`` with 'head-batch' / 'tail-batch' -> ``[B,N]`` (``[B,1]`` for tail-batch as the reference
        wrote it, see the module docstring)."""
        if mode == "single":
            positive, _ = self._check(sample, None, mode)
            pos_s, _ = _Projection.apply(self.entity_embedding, self.relation_embedding, self.W, self, positive, None, mode)
            return pos_s.view(-1, 1)
        if mode not in ("head-batch", "tail-batch"):
            raise ValueError("mode %s not supported" % mode)
        return self.both(sample, mode)[1]

    # ---- train steps -----------------------------------------------------------------------------
    def _grad_buffers(self):
        if self._grads is None or self._grads[0].device != self.entity_embedding.device:
            self._grads = tuple(torch.empty_like(p) for p in (self.entity_embedding, self.relation_embedding, self.W))
        return self._grads

    def _step_grads(self, positive, negative, weight, mode, adv, alpha, uni, detach_weights):
        """forward -> K2 loss (kge_loss_forward_backward) -> backward; returns (loss3, (g_ent, g_rel, g_W))."""
        positive, negative = self._check(positive, negative, mode)
        B, N = negative.shape
        pos_s, neg_s, ws = self._forward_raw(positive, negative, mode, with_backward=True)
        loss3, dneg, dpos = ops.loss_forward_backward(
            neg_s, pos_s, None if uni else weight.to(positive.device), negative_adversarial_sampling=adv,
            adversarial_temperature=alpha, uni_weight=uni, detach_weights=detach_weights)
        grads = self._grad_buffers()
        self._backward_raw(positive, mode, B, N, dpos, dneg, *grads, ws)
        self._release(ws)
        return loss3, grads

    @staticmethod
    def train_step(model, optimizer, train_iterator, args):
        """``KGEModel.train_step`` contract (log dict with the upstream keys) for any torch optimizer over the model's
        parameters: the gradients land in ``.grad`` and ``optimizer.step()`` applies them."""
        model.train()
        positive_sample, negative_sample, subsampling_weight, mode = next(train_iterator)
        if mode not in ("head-batch", "tail-batch"):
            raise ValueError("mode %s not supported" % mode)
        adv = bool(getattr(args, "negative_adversarial_sampling", False))
        alpha = float(getattr(args, "adversarial_temperature", 1.0))
        uni = bool(getattr(args, "uni_weight", False))
        reg = float(getattr(args, "regularization", 0.0))
        optimizer.zero_grad()
        loss3, grads = model._step_grads(positive_sample, negative_sample, subsampling_weight.reshape(-1), mode, adv, alpha,
                                         uni, True)
        pos_l, neg_l, loss = loss3.tolist()
        log = {}
        for p, g in zip((model.entity_embedding, model.relation_embedding, model.W), grads):
            p.grad = g.clone()
        if reg != 0.0:  # model.py:33-39: L3 over the two embedding tables
            regv = reg * (ops.l3_sum_cubes(model.entity_embedding.detach()) + ops.l3_sum_cubes(model.relation_embedding.detach())).item()
            with torch.no_grad():
                for p in (model.entity_embedding, model.relation_embedding):
                    p.grad.add_(3.0 * reg * p * p.abs())
            loss += regv
            log["regularization"] = regv
        optimizer.step()
        log.update({"positive_sample_loss": pos_l, "negative_sample_loss": neg_l, "loss": loss})
        return log

    def tf_train_step(self, optimizer, positive_sample, negative_sample, subsampling_weight, mode: str) -> torch.Tensor:
        """supervisor.py:13-30 for this model: both ``call``s (attached softmax weights, temperature 1), the
        subsampling-weighted means, /2, gradients of the three variables, Keras Adam (dense, as TF applies the
        gathered-variable gradients).  ``optimizer``: ``tf_compat.KerasAdam``.  Returns the loss (device scalar)."""
        loss3, grads = self._step_grads(positive_sample, negative_sample, subsampling_weight.reshape(-1), mode, True, 1.0,
                                        False, False)
        group = optimizer.param_groups[0]
        step = optimizer.iterations + 1
        for p, g in zip((self.entity_embedding, self.relation_embedding, self.W), grads):
            st = optimizer.state[p]
            if not st:
                st["exp_avg"], st["exp_avg_sq"] = torch.zeros_like(p), torch.zeros_like(p)
            ops.adam_dense(p.data, g, st["exp_avg"], st["exp_avg_sq"], lr=float(group["lr"]), beta1=group["betas"][0],
                           beta2=group["betas"][1], eps=group["eps"], step=step, keras_epsilon_mode=True)
        return loss3[2]
