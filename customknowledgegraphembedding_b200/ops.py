"""Torch-facing wrappers of the C-ABI: torch is used for device memory and streams only.

Every function takes CUDA tensors, checks dtype / contiguity / device, and enqueues the kernels of
``libkge_b200.so`` on ``torch.cuda.current_stream()``.  CPU tensors are refused -- there is no
fallback path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import AdamOpts, LossOpts, ModelDesc


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(name: str, t: torch.Tensor, dtype) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise RuntimeError("%s is on %s: the B200 KGE kernels need CUDA tensors (no CPU fallback); "
                           "move the model and batch to the GPU" % (name, t.device))
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    return t if t.is_contiguous() else t.contiguous()


def make_desc(model_name: str, nentity: int, nrelation: int, hidden_dim: int, entity_dim: int, relation_dim: int,
              gamma: float, embedding_range: float, interht_eps: float = 0.0) -> ModelDesc:
    if model_name not in _lib.MODEL_IDS:
        raise ValueError("model %s not supported" % model_name)
    return ModelDesc(_lib.MODEL_IDS[model_name], hidden_dim, entity_dim, relation_dim, nentity, nrelation,
                     gamma, embedding_range, interht_eps, 0)


def mode_id(mode: str) -> int:
    if mode not in _lib.MODE_IDS:
        raise ValueError("mode %s not supported" % mode)
    return _lib.MODE_IDS[mode]


def workspace_bytes(kind: int, desc: ModelDesc, B: int, N: int) -> int:
    return int(_lib.load().kge_workspace_bytes(kind, C.byref(desc), B, N))


def score_forward(desc: ModelDesc, mode: str, entity, relation, modulus, positive, negative) -> torch.Tensor:
    lib = _lib.load()
    m = mode_id(mode)
    entity = _need_cuda("entity_embedding", entity, torch.float32)
    relation = _need_cuda("relation_embedding", relation, torch.float32)
    positive = _need_cuda("positive_sample", positive, torch.int64)
    B = positive.shape[0]
    if m == 0:
        N, negative = 1, None
    else:
        negative = _need_cuda("negative_sample", negative, torch.int64)
        if negative.dim() != 2 or negative.shape[0] != B:
            raise ValueError("negative_sample must be [B,N] with B=%d, got %s" % (B, tuple(negative.shape)))
        N = negative.shape[1]
    if modulus is not None:
        modulus = _need_cuda("modulus", modulus, torch.float32)
    scores = torch.empty((B, N), dtype=torch.float32, device=entity.device)
    with torch.cuda.device(entity.device):
        _lib.check(lib.kge_score_forward(C.byref(desc), m, _ptr(entity), _ptr(relation), _ptr(modulus),
                                         _ptr(positive), _ptr(negative), B, N, _ptr(scores), _stream()),
                   "kge_score_forward")
    return scores


def score_backward(desc: ModelDesc, mode: str, entity, relation, modulus, positive, negative, dscore,
                   grad_entity, grad_relation, grad_modulus, deterministic: bool = False) -> None:
    """Accumulates into grad_entity / grad_relation / grad_modulus (dense, same shapes as the tables)."""
    lib = _lib.load()
    m = mode_id(mode)
    B = positive.shape[0]
    N = 1 if m == 0 else negative.shape[1]
    dscore = _need_cuda("dscore", dscore, torch.float32)
    ws, ws_bytes = None, 0
    if deterministic:
        ws_bytes = workspace_bytes(_lib.WS_BACKWARD, desc, B, N)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=entity.device)
    with torch.cuda.device(entity.device):
        _lib.check(lib.kge_score_backward(C.byref(desc), m, _ptr(entity), _ptr(relation), _ptr(modulus),
                                          _ptr(positive), _ptr(negative if m != 0 else None), B, N, _ptr(dscore),
                                          _ptr(grad_entity), _ptr(grad_relation), _ptr(grad_modulus),
                                          1 if deterministic else 0, _ptr(ws), ws_bytes, _stream()),
                   "kge_score_backward")


def loss_forward_backward(negative_score, positive_score, subsampling_weight, *, negative_adversarial_sampling: bool,
                          adversarial_temperature: float = 1.0, uni_weight: bool = False,
                          detach_weights: bool = True, want_grad: bool = True
                          ) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor]]:
    """-> (loss3 = [positive_sample_loss, negative_sample_loss, loss], dL/dneg [B,N], dL/dpos [B])."""
    lib = _lib.load()
    negative_score = _need_cuda("negative_score", negative_score, torch.float32)
    positive_score = _need_cuda("positive_score", positive_score.reshape(-1), torch.float32)
    B, N = negative_score.shape
    w = None
    if not uni_weight:
        w = _need_cuda("subsampling_weight", subsampling_weight.reshape(-1), torch.float32)
    dev = negative_score.device
    loss3 = torch.empty(3, dtype=torch.float32, device=dev)
    dneg = torch.empty_like(negative_score) if want_grad else None
    dpos = torch.empty(B, dtype=torch.float32, device=dev) if want_grad else None
    ws = torch.empty(2 * B + 8, dtype=torch.float32, device=dev)
    opts = LossOpts(int(negative_adversarial_sampling), int(uni_weight), int(detach_weights), 0,
                    float(adversarial_temperature), 0.0)
    with torch.cuda.device(dev):
        _lib.check(lib.kge_loss_forward_backward(_ptr(negative_score), _ptr(positive_score), _ptr(w), B, N,
                                                 C.byref(opts), _ptr(loss3), _ptr(dneg), _ptr(dpos), _ptr(ws),
                                                 ws.numel() * 4, _stream()), "kge_loss_forward_backward")
    return loss3, dneg, dpos


def check_ids(positive, negative, nentity: int, nrelation: int) -> None:
    """Device-side range check of a batch (kge_check_ids): IndexError as ``index_select`` / ``tf.gather`` would raise.
    Synchronises the stream -- a debugging switch (``model.check_ids = True``), not part of the timed path."""
    lib = _lib.load()
    positive = _need_cuda("positive_sample", positive, torch.int64)
    B = positive.shape[0]
    N = 0
    if negative is not None:
        negative = _need_cuda("negative_sample", negative, torch.int64)
        N = negative.shape[1]
    scratch = torch.empty(4, dtype=torch.int64, device=positive.device)
    with torch.cuda.device(positive.device):
        rc = lib.kge_check_ids(_ptr(positive), _ptr(negative), B, N, int(nentity), int(nrelation), _ptr(scratch), _stream())
    if rc == _lib.KGE_ERR_INVALID_ARGUMENT:
        raise IndexError(lib.kge_last_error().decode())
    _lib.check(rc, "kge_check_ids")


def l3_sum_cubes(x: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    x = _need_cuda("x", x, torch.float32)
    out = torch.empty(1, dtype=torch.float32, device=x.device)
    ws = torch.empty(1024, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.kge_l3_sum_cubes(_ptr(x), x.numel(), _ptr(out), _ptr(ws), 4096, _stream()), "kge_l3_sum_cubes")
    return out


def adam_dense(p, g, m, v, *, lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, step: int,
               keras_epsilon_mode: bool = False, l3_lambda: float = 0.0, zero_grad: bool = False) -> None:
    lib = _lib.load()
    for name, t in (("p", p), ("m", m), ("v", v)):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError("adam_dense: %s must be a contiguous CUDA float32 tensor" % name)
    if g is not None and not (g.is_cuda and g.dtype == torch.float32 and g.is_contiguous()):
        raise RuntimeError("adam_dense: g must be a contiguous CUDA float32 tensor")
    opts = AdamOpts(lr, beta1, beta2, eps, int(step), int(keras_epsilon_mode))
    with torch.cuda.device(p.device):
        _lib.check(lib.kge_adam_dense(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), C.byref(opts),
                                      float(l3_lambda), int(zero_grad), _stream()), "kge_adam_dense")


def eval_rank(desc: ModelDesc, mode: str, entity, relation, modulus, queries, entity_lo: int, entity_hi: int,
              filter_indptr=None, filter_idx=None, out=None, accumulate: bool = False):
    """-> (true_score [Q] f32, n_greater [Q] i64, n_equal [Q] i64) over candidates [entity_lo, entity_hi)."""
    lib = _lib.load()
    m = mode_id(mode)
    entity = _need_cuda("entity_embedding", entity, torch.float32)
    relation = _need_cuda("relation_embedding", relation, torch.float32)
    queries = _need_cuda("queries", queries, torch.int64)
    Q = queries.shape[0]
    dev = entity.device
    if filter_indptr is not None:
        filter_indptr = _need_cuda("filter_indptr", filter_indptr, torch.int64)
        filter_idx = _need_cuda("filter_idx", filter_idx, torch.int64)
        nflt = filter_idx.numel()
    else:
        nflt = 0
    if out is None:
        true_score = torch.empty(Q, dtype=torch.float32, device=dev)
        n_greater = torch.zeros(Q, dtype=torch.int64, device=dev)
        n_equal = torch.zeros(Q, dtype=torch.int64, device=dev)
    else:
        true_score, n_greater, n_equal = out
    ws_bytes = workspace_bytes(_lib.WS_EVAL, desc, Q, nflt)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.kge_eval_rank(C.byref(desc), m, _ptr(entity), _ptr(relation), _ptr(modulus), _ptr(queries), Q,
                                     entity_lo, entity_hi, _ptr(filter_indptr), _ptr(filter_idx), _ptr(true_score),
                                     _ptr(n_greater), _ptr(n_equal), int(accumulate), _ptr(ws), ws_bytes, _stream()),
                   "kge_eval_rank")
    return true_score, n_greater, n_equal


def eval_rank_shard(desc: ModelDesc, mode: str, entity_shard, shard_row0: int, relation, modulus, queries, query_rows,
                    entity_lo: int, entity_hi: int, filter_indptr=None, filter_idx=None, out=None,
                    accumulate: bool = False):
    """kge_eval_rank_shard: `entity_shard` holds rows [shard_row0, shard_row0 + len) of the global table,
    `query_rows` [Q,2,De] the (head, tail) rows of the queries.  -> (true_score, n_greater, n_equal) of the shard."""
    lib = _lib.load()
    m = mode_id(mode)
    entity_shard = _need_cuda("entity_shard", entity_shard, torch.float32)
    relation = _need_cuda("relation_embedding", relation, torch.float32)
    queries = _need_cuda("queries", queries, torch.int64)
    query_rows = _need_cuda("query_rows", query_rows, torch.float32)
    Q = queries.shape[0]
    if query_rows.numel() != Q * 2 * desc.entity_dim:
        raise ValueError("query_rows must be [Q,2,entity_dim]")
    dev = entity_shard.device
    if filter_indptr is not None:
        filter_indptr = _need_cuda("filter_indptr", filter_indptr, torch.int64)
        filter_idx = _need_cuda("filter_idx", filter_idx, torch.int64)
        nflt = filter_idx.numel()
    else:
        nflt = 0
    if out is None:
        true_score = torch.empty(Q, dtype=torch.float32, device=dev)
        n_greater = torch.zeros(Q, dtype=torch.int64, device=dev)
        n_equal = torch.zeros(Q, dtype=torch.int64, device=dev)
    else:
        true_score, n_greater, n_equal = out
    if Q == 0:
        return true_score, n_greater, n_equal
    ws_bytes = workspace_bytes(_lib.WS_EVAL, desc, Q, nflt)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    # an empty shard still passes a valid (never dereferenced) pointer
    ent_ptr = _ptr(entity_shard) if entity_shard.numel() else _ptr(relation)
    with torch.cuda.device(dev):
        _lib.check(lib.kge_eval_rank_shard(C.byref(desc), m, ent_ptr, int(shard_row0), int(entity_shard.shape[0]),
                                           _ptr(relation), _ptr(modulus), _ptr(queries), _ptr(query_rows), Q,
                                           int(entity_lo), int(entity_hi), _ptr(filter_indptr), _ptr(filter_idx),
                                           _ptr(true_score), _ptr(n_greater), _ptr(n_equal), int(accumulate), _ptr(ws),
                                           ws_bytes, _stream()), "kge_eval_rank_shard")
    return true_score, n_greater, n_equal
