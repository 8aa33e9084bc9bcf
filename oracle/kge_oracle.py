"""CPU oracle (PyTorch, fp32 with an fp64 twin) for the KGE hot path.

TEST INFRASTRUCTURE -- see ``oracle/__init__.py`` for who may import this.

Every function cites the reference lines it restates.  ``/root/reference`` paths:

* ``tensorflow_codes/model.py:4-44``     eager ``loss()``  (PyTorch loss semantics)
* ``tensorflow_codes/model.py:49-112``   ``TFKGEModel.__init__``  (dims, init range)
* ``tensorflow_codes/model.py:114-205``  ``TFKGEModel.call``  (modes 0/1/3, fused terms)
* ``tensorflow_codes/model.py:207-224``  ``InterHT``
* ``tensorflow_codes/model.py:96-106,226-235`` ``TranSparse`` (``OracleTranSparse``)
* ``tensorflow_codes/supervisor.py:13-30`` ``Trainer.train_step``
* SURVEY.md section 8(a) rows U1-U11 -- the absent PyTorch submodule
  (``KnowledgeGraphEmbedding`` @ 2e440e0f, "parity unpinned").

The module is deliberately plain: materialise the gathers, run the elementwise
chain, let autograd and ``torch.optim.Adam`` define gradients and updates, and
let ``torch.argsort(descending=True)`` define ranks.
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

MODEL_NAMES = ("TransE", "DistMult", "ComplEx", "RotatE", "pRotatE", "InterHT")
MODES = ("single", "head-batch", "tail-batch")

# TF wire encoding of the mode (compress_data/main.py:65,69; supervisor.py:18; model.py:124,203)
TF_MODE_HEAD_BATCH = 0
TF_MODE_TAIL_BATCH = 1
TF_MODE_SINGLE = 3


class OracleKGEModel(nn.Module):
    """Restatement of the PyTorch ``KGEModel`` plus InterHT
    (``tensorflow_codes/model.py:207-224``) and the ``-tr`` flag (``model.py:75-78``)."""

    def __init__(self, model_name, nentity, nrelation, hidden_dim, gamma,
                 double_entity_embedding=False, double_relation_embedding=False,
                 triple_relation_embedding=False, dtype=torch.float32,
                 interht_eps: float = 0.0):
        super().__init__()
        self.model_name = model_name
        self.nentity = nentity
        self.nrelation = nrelation
        self.hidden_dim = hidden_dim
        self.epsilon = 2.0
        self.u = 1.0  # model.py:94-95
        # eps of the L2 normalisation in InterHT: tf.linalg.normalize has none
        # (model.py:215-218); F.normalize clamps at 1e-12 (U8).  Identical away from 0 rows.
        self.interht_eps = interht_eps

        # U1: gamma / embedding_range are fp32 [1] non-trainable parameters whatever the table dtype
        self.gamma = nn.Parameter(torch.tensor([gamma], dtype=torch.float32), requires_grad=False)
        self.embedding_range = nn.Parameter(
            torch.tensor([(self.gamma.item() + self.epsilon) / hidden_dim], dtype=torch.float32),
            requires_grad=False)

        self.entity_dim = hidden_dim * 2 if double_entity_embedding else hidden_dim
        self.relation_dim = hidden_dim * 2 if double_relation_embedding else hidden_dim
        if triple_relation_embedding:
            self.relation_dim = hidden_dim * 3

        self.entity_embedding = nn.Parameter(torch.zeros(nentity, self.entity_dim, dtype=dtype))
        nn.init.uniform_(self.entity_embedding, -self.embedding_range.item(), self.embedding_range.item())
        self.relation_embedding = nn.Parameter(torch.zeros(nrelation, self.relation_dim, dtype=dtype))
        nn.init.uniform_(self.relation_embedding, -self.embedding_range.item(), self.embedding_range.item())

        if model_name == "pRotatE":
            self.modulus = nn.Parameter(torch.tensor([[0.5 * self.embedding_range.item()]], dtype=dtype))

        if model_name not in MODEL_NAMES:
            raise ValueError("model %s not supported" % model_name)
        if model_name == "RotatE" and (not double_entity_embedding or double_relation_embedding):
            raise ValueError("RotatE should use --double_entity_embedding")
        if model_name == "ComplEx" and (not double_entity_embedding or not double_relation_embedding):
            raise ValueError("ComplEx should use --double_entity_embedding and --double_relation_embedding")
        if model_name == "InterHT" and (not double_entity_embedding or not triple_relation_embedding):
            raise ValueError("InterHT should use --double_entity_embedding and --triple_relation_embedding")

    # ---- U2 ------------------------------------------------------------------------------
    def gather(self, sample, mode="single"):
        if mode == "single":
            head = torch.index_select(self.entity_embedding, 0, sample[:, 0]).unsqueeze(1)
            relation = torch.index_select(self.relation_embedding, 0, sample[:, 1]).unsqueeze(1)
            tail = torch.index_select(self.entity_embedding, 0, sample[:, 2]).unsqueeze(1)
        elif mode == "head-batch":
            tail_part, head_part = sample
            b, n = head_part.size(0), head_part.size(1)
            head = torch.index_select(self.entity_embedding, 0, head_part.reshape(-1)).view(b, n, -1)
            relation = torch.index_select(self.relation_embedding, 0, tail_part[:, 1]).unsqueeze(1)
            tail = torch.index_select(self.entity_embedding, 0, tail_part[:, 2]).unsqueeze(1)
        elif mode == "tail-batch":
            head_part, tail_part = sample
            b, n = tail_part.size(0), tail_part.size(1)
            head = torch.index_select(self.entity_embedding, 0, head_part[:, 0]).unsqueeze(1)
            relation = torch.index_select(self.relation_embedding, 0, head_part[:, 1]).unsqueeze(1)
            tail = torch.index_select(self.entity_embedding, 0, tail_part.reshape(-1)).view(b, n, -1)
        else:
            raise ValueError("mode %s not supported" % mode)
        return head, relation, tail

    def forward(self, sample, mode="single"):
        head, relation, tail = self.gather(sample, mode)
        fn = getattr(self, self.model_name)
        return fn(head, relation, tail, mode)

    # ---- U3 ------------------------------------------------------------------------------
    def TransE(self, head, relation, tail, mode):
        if mode == "head-batch":
            score = head + (relation - tail)
        else:
            score = (head + relation) - tail
        return self.gamma.item() - torch.norm(score, p=1, dim=2)

    # ---- U4 ------------------------------------------------------------------------------
    def DistMult(self, head, relation, tail, mode):
        if mode == "head-batch":
            score = head * (relation * tail)
        else:
            score = (head * relation) * tail
        return score.sum(dim=2)

    # ---- U5 ------------------------------------------------------------------------------
    def ComplEx(self, head, relation, tail, mode):
        re_head, im_head = torch.chunk(head, 2, dim=2)
        re_relation, im_relation = torch.chunk(relation, 2, dim=2)
        re_tail, im_tail = torch.chunk(tail, 2, dim=2)
        if mode == "head-batch":
            re_score = re_relation * re_tail + im_relation * im_tail
            im_score = re_relation * im_tail - im_relation * re_tail
            score = re_head * re_score + im_head * im_score
        else:
            re_score = re_head * re_relation - im_head * im_relation
            im_score = re_head * im_relation + im_head * re_relation
            score = re_score * re_tail + im_score * im_tail
        return score.sum(dim=2)

    # ---- U6 ------------------------------------------------------------------------------
    def RotatE(self, head, relation, tail, mode):
        pi = 3.14159265358979323846
        re_head, im_head = torch.chunk(head, 2, dim=2)
        re_tail, im_tail = torch.chunk(tail, 2, dim=2)
        phase_relation = relation / (self.embedding_range.item() / pi)
        re_relation = torch.cos(phase_relation)
        im_relation = torch.sin(phase_relation)
        if mode == "head-batch":
            re_score = re_relation * re_tail + im_relation * im_tail
            im_score = re_relation * im_tail - im_relation * re_tail
            re_score = re_score - re_head
            im_score = im_score - im_head
        else:
            re_score = re_head * re_relation - im_head * im_relation
            im_score = re_head * im_relation + im_head * re_relation
            re_score = re_score - re_tail
            im_score = im_score - im_tail
        score = torch.stack([re_score, im_score], dim=0)
        score = score.norm(dim=0)
        return self.gamma.item() - score.sum(dim=2)

    # ---- U7 ------------------------------------------------------------------------------
    def pRotatE(self, head, relation, tail, mode):
        pi = 3.14159262358979323846  # sic (upstream constant)
        phase_head = head / (self.embedding_range.item() / pi)
        phase_relation = relation / (self.embedding_range.item() / pi)
        phase_tail = tail / (self.embedding_range.item() / pi)
        if mode == "head-batch":
            score = phase_head + (phase_relation - phase_tail)
        else:
            score = (phase_head + phase_relation) - phase_tail
        score = torch.sin(score)
        score = torch.abs(score)
        return self.gamma.item() - score.sum(dim=2) * self.modulus

    # ---- P5 / U8 -------------------------------------------------------------------------
    def _l2_normalize(self, x):
        norm = torch.sqrt((x * x).sum(dim=-1, keepdim=True))  # tf.linalg.normalize(ord=2), model.py:215-218
        if self.interht_eps > 0.0:
            norm = norm.clamp_min(self.interht_eps)            # F.normalize(x, 2, -1)
        return x / norm

    def InterHT(self, head, relation, tail, mode):
        a_head, b_head = torch.chunk(head, 2, dim=2)            # model.py:208
        _, re_mid, _ = torch.chunk(relation, 3, dim=2)          # model.py:209
        a_tail, b_tail = torch.chunk(tail, 2, dim=2)            # model.py:210
        a_head = self._l2_normalize(a_head)
        a_tail = self._l2_normalize(a_tail)
        b_head = self._l2_normalize(b_head)
        b_tail = self._l2_normalize(b_tail)
        b_head = b_head + self.u                                 # model.py:219
        b_tail = b_tail + self.u                                 # model.py:220
        score = a_head * b_tail - a_tail * b_head + re_mid      # model.py:222
        return self.gamma.item() - score.abs().sum(dim=2)       # model.py:223 (tf.norm ord=1)


class OracleTranSparse(nn.Module):
    """``TFKGEModel('TranSparse', ...)`` (``tensorflow_codes/model.py:96-106`` variables, ``:139-142,161-164,187-190``
    gathers, ``:226-235`` score): per-relation projection ``W [R, D, D]`` under a fixed 0/1 ``mask``.

    ``project_tail=False`` is the reference AS WRITTEN: ``:229`` projects ``head`` a second time, so ``p_tail`` is
    ``p_head`` and the tail entity is never read -- the score is ``gamma - sum_j |p_j| * |r_j - 1|`` with ``p`` the
    normalised projected head and ``r`` the normalised relation row; in tail-batch mode the result is ``[B, 1]``
    (the head side has one row per positive).  ``project_tail=True`` projects the tail (what ``:229`` evidently meant).
    ``mask`` follows ``:97-103`` in distribution only (0/1 with rate 0.5, drawn here with torch's generator): tests
    copy the tensor they compare on, as the Checker does for every variable."""

    def __init__(self, nentity, nrelation, hidden_dim, gamma, project_tail=False, dtype=torch.float32):
        super().__init__()
        self.model_name = "TranSparse"
        self.nentity, self.nrelation, self.hidden_dim = nentity, nrelation, hidden_dim
        self.entity_dim = self.relation_dim = hidden_dim
        self.epsilon = 2.0
        self.project_tail = bool(project_tail)
        self.gamma = nn.Parameter(torch.tensor([gamma], dtype=torch.float32), requires_grad=False)
        rng = (float(gamma) + self.epsilon) / hidden_dim
        self.embedding_range = nn.Parameter(torch.tensor([rng], dtype=torch.float32), requires_grad=False)
        self.entity_embedding = nn.Parameter(torch.empty(nentity, hidden_dim, dtype=dtype).uniform_(-rng, rng))
        self.relation_embedding = nn.Parameter(torch.empty(nrelation, hidden_dim, dtype=dtype).uniform_(-rng, rng))
        self.W = nn.Parameter(torch.empty(nrelation, hidden_dim, hidden_dim, dtype=dtype).uniform_(-rng, rng))
        prob = torch.empty(nrelation, hidden_dim, hidden_dim).uniform_(1, 100)           # model.py:101
        self.register_buffer("mask", (prob >= 50).to(dtype))                            # model.py:99-103

    gather = OracleKGEModel.gather

    @staticmethod
    def _l2_normalize(x):
        return x / torch.sqrt((x * x).sum(dim=-1, keepdim=True))                         # tf.linalg.normalize(ord=2)

    def forward(self, sample, mode="single"):
        head, relation, tail = self.gather(sample, mode)
        rel_ids = (sample if mode == "single" else sample[0])[:, 1]
        m = self.mask[rel_ids] * self.W[rel_ids]                                          # model.py:139-141, 227
        p_head = self._l2_normalize(torch.matmul(head, m))                                # model.py:227-228
        p_tail = self._l2_normalize(torch.matmul(tail if self.project_tail else head, m))  # model.py:229-230
        relation = self._l2_normalize(relation)                                           # model.py:231
        score = p_head * relation - p_tail                                                # model.py:233
        return self.gamma.item() - score.abs().sum(dim=2)                                 # model.py:234


def transparse_to_double(model: "OracleTranSparse") -> "OracleTranSparse":
    twin = OracleTranSparse(model.nentity, model.nrelation, model.hidden_dim, model.gamma.item(),
                            model.project_tail, dtype=torch.float64)
    with torch.no_grad():
        for k in ("entity_embedding", "relation_embedding", "W"):
            getattr(twin, k).copy_(getattr(model, k).double())
        twin.mask.copy_(model.mask.double())
    return twin


# ------------------------------------------------------------------------------------------
# Loss (tensorflow_codes/model.py:4-44  ==  upstream train_step body, U9)
# ------------------------------------------------------------------------------------------
def kge_loss(negative_score, positive_score, subsampling_weight, *,
             negative_adversarial_sampling: bool, adversarial_temperature: float = 1.0,
             uni_weight: bool = False, detach_weights: bool = True):
    """``negative_score [B,N]``, ``positive_score [B,1]`` raw scores -> (pos_loss, neg_loss, loss).

    model.py:12-18 (adv / mean), :20-22 (positive), :24-29 (uni / subsampling), :31 (/2).
    ``detach_weights=False`` is the TF graph model's behaviour (model.py:168-171: the
    softmax is not stop-gradient'ed)."""
    if negative_adversarial_sampling:
        p = F.softmax(negative_score * adversarial_temperature, dim=1)
        if detach_weights:
            p = p.detach()
        neg = (p * F.logsigmoid(-negative_score)).sum(dim=1)
    else:
        neg = F.logsigmoid(-negative_score).mean(dim=1)
    pos = F.logsigmoid(positive_score).squeeze(dim=1)
    if uni_weight:
        positive_sample_loss = -pos.mean()
        negative_sample_loss = -neg.mean()
    else:
        positive_sample_loss = -(subsampling_weight * pos).sum() / subsampling_weight.sum()
        negative_sample_loss = -(subsampling_weight * neg).sum() / subsampling_weight.sum()
    loss = (positive_sample_loss + negative_sample_loss) / 2
    return positive_sample_loss, negative_sample_loss, loss


def l3_regularization(model, regularization: float):
    """model.py:33-39: ``lambda * (||E||_3^3 + || ||R||_3 ||_3^3)``."""
    return regularization * (model.entity_embedding.norm(p=3) ** 3
                             + model.relation_embedding.norm(p=3).norm(p=3) ** 3)


class Args:
    """Minimal stand-in for the upstream argparse namespace (U9/U10 flags)."""

    def __init__(self, **kw):
        self.cuda = False
        self.negative_adversarial_sampling = False
        self.adversarial_temperature = 1.0
        self.uni_weight = False
        self.regularization = 0.0
        self.test_batch_size = 4
        self.cpu_num = 0
        self.test_log_steps = 10 ** 9
        self.countries = False
        self.nentity = None
        self.nrelation = None
        self.__dict__.update(kw)


def train_step(model: OracleKGEModel, optimizer, train_iterator, args) -> Dict[str, float]:
    """Upstream ``KGEModel.train_step`` (U9; loss body == tensorflow_codes/model.py:4-44)."""
    model.train()
    optimizer.zero_grad()
    positive_sample, negative_sample, subsampling_weight, mode = next(train_iterator)
    negative_score = model((positive_sample, negative_sample), mode=mode)
    positive_score = model(positive_sample)
    pos_l, neg_l, loss = kge_loss(
        negative_score, positive_score, subsampling_weight,
        negative_adversarial_sampling=args.negative_adversarial_sampling,
        adversarial_temperature=args.adversarial_temperature,
        uni_weight=args.uni_weight)
    if args.regularization != 0.0:
        reg = l3_regularization(model, args.regularization)
        loss = loss + reg
        reg_log = {"regularization": reg.item()}
    else:
        reg_log = {}
    loss.backward()
    optimizer.step()
    return {**reg_log, "positive_sample_loss": pos_l.item(),
            "negative_sample_loss": neg_l.item(), "loss": loss.item()}


# ------------------------------------------------------------------------------------------
# Filtered full-entity ranking (upstream test_step + TestDataset, U10)
# ------------------------------------------------------------------------------------------
def test_dataset_row(triple, all_true: set, nentity: int, mode: str):
    """Upstream ``TestDataset.__getitem__`` for one triple: candidates that form a known
    true triple are replaced by the true entity id with bias -1 (U10)."""
    head, relation, tail = triple
    ids = np.arange(nentity, dtype=np.int64)
    bias = np.zeros(nentity, dtype=np.float32)
    if mode == "head-batch":
        for e in range(nentity):
            if (e, relation, tail) in all_true:
                ids[e] = head
                bias[e] = -1.0
        ids[head] = head
        bias[head] = 0.0
    elif mode == "tail-batch":
        for e in range(nentity):
            if (head, relation, e) in all_true:
                ids[e] = tail
                bias[e] = -1.0
        ids[tail] = tail
        bias[tail] = 0.0
    else:
        raise ValueError("negative batch mode %s not supported" % mode)
    return (torch.tensor([head, relation, tail], dtype=torch.int64),
            torch.from_numpy(ids), torch.from_numpy(bias), mode)


def build_true_maps(all_true_triples: Iterable[Tuple[int, int, int]]):
    """(h,r)->sorted tails and (r,t)->sorted heads over all true triples."""
    hr2t: Dict[Tuple[int, int], List[int]] = {}
    rt2h: Dict[Tuple[int, int], List[int]] = {}
    for h, r, t in all_true_triples:
        hr2t.setdefault((h, r), []).append(t)
        rt2h.setdefault((r, t), []).append(h)
    return hr2t, rt2h


def test_dataset_row_fast(triple, hr2t, rt2h, nentity: int, mode: str):
    """Same output as :func:`test_dataset_row` using the true-maps instead of an E-long scan."""
    head, relation, tail = triple
    ids = np.arange(nentity, dtype=np.int64)
    bias = np.zeros(nentity, dtype=np.float32)
    if mode == "head-batch":
        flt = np.asarray(rt2h.get((relation, tail), []), dtype=np.int64)
        ids[flt] = head
        bias[flt] = -1.0
        ids[head] = head
        bias[head] = 0.0
    else:
        flt = np.asarray(hr2t.get((head, relation), []), dtype=np.int64)
        ids[flt] = tail
        bias[flt] = -1.0
        ids[tail] = tail
        bias[tail] = 0.0
    return (torch.tensor([head, relation, tail], dtype=torch.int64),
            torch.from_numpy(ids), torch.from_numpy(bias), mode)


def metrics_from_ranks(ranks: Sequence[int]) -> Dict[str, float]:
    """Upstream per-query log -> averaged metrics (U10)."""
    logs = []
    for ranking in ranks:
        logs.append({
            "MRR": 1.0 / ranking,
            "MR": float(ranking),
            "HITS@1": 1.0 if ranking <= 1 else 0.0,
            "HITS@3": 1.0 if ranking <= 3 else 0.0,
            "HITS@10": 1.0 if ranking <= 10 else 0.0,
        })
    metrics = {}
    for metric in logs[0].keys():
        metrics[metric] = sum(log[metric] for log in logs) / len(logs)
    return metrics


@torch.no_grad()
def test_step(model: OracleKGEModel, test_triples, all_true_triples, args,
              return_ranks: bool = False, return_scores: bool = False):
    """Upstream ``KGEModel.test_step`` (U10): for head-batch then tail-batch, score every test
    triple against all entities, add the filter bias, ``argsort(descending=True)``, rank = 1 +
    position of the true entity.  Query order: all head-batch queries, then all tail-batch."""
    model.eval()
    nentity = args.nentity if args.nentity is not None else model.nentity
    hr2t, rt2h = build_true_maps(all_true_triples)
    ranks: List[int] = []
    all_scores = []
    bs = args.test_batch_size
    for mode in ("head-batch", "tail-batch"):
        for lo in range(0, len(test_triples), bs):
            rows = [test_dataset_row_fast(t, hr2t, rt2h, nentity, mode) for t in test_triples[lo:lo + bs]]
            positive_sample = torch.stack([r[0] for r in rows], 0)
            negative_sample = torch.stack([r[1] for r in rows], 0)
            filter_bias = torch.stack([r[2] for r in rows], 0)
            score = model((positive_sample, negative_sample), mode)
            score = score + filter_bias.to(score.dtype)
            argsort = torch.argsort(score, dim=1, descending=True)
            positive_arg = positive_sample[:, 0] if mode == "head-batch" else positive_sample[:, 2]
            for i in range(positive_sample.size(0)):
                ranking = (argsort[i, :] == positive_arg[i]).nonzero()
                assert ranking.size(0) == 1
                ranks.append(1 + ranking.item())
            if return_scores:
                all_scores.append(score)
    metrics = metrics_from_ranks(ranks)
    out = (metrics,)
    if return_ranks:
        out = out + (np.asarray(ranks, dtype=np.int64),)
    if return_scores:
        out = out + (torch.cat(all_scores, 0),)
    return out[0] if len(out) == 1 else out


# ------------------------------------------------------------------------------------------
# TF-shaped model / trainer semantics (tensorflow_codes/model.py:114-205, supervisor.py:13-30)
# ------------------------------------------------------------------------------------------
def tf_call(model: OracleKGEModel, sample, mode_int: int):
    """``TFKGEModel.call((sample, mode))`` -> fused per-positive term ``[B,1]``.

    mode 3 -> ``log_sigmoid(single score)`` (model.py:127-146, 203-205);
    mode 0 / other -> ``sum_n softmax(s)_n * log_sigmoid(-s_n)`` over head-batch / tail-batch
    scores, temperature 1, gradient flowing through the softmax (model.py:148-199, 121-125)."""
    positive_sample, negative_sample = sample
    if int(mode_int) == TF_MODE_SINGLE:
        s = model(positive_sample, "single")
        return F.logsigmoid(s)
    mode = "head-batch" if int(mode_int) == TF_MODE_HEAD_BATCH else "tail-batch"
    s = model((positive_sample, negative_sample), mode)
    return (F.softmax(s * 1, dim=1) * F.logsigmoid(-s)).sum(dim=1, keepdim=True)


def tf_train_loss(model: OracleKGEModel, positive_sample, negative_sample, subsampling_weight, mode):
    """supervisor.py:17-23: two model calls, subsampling-weighted means, /2.
    ``subsampling_weight`` is ``[B,1]`` (run.py:63), ``mode`` is ``[B]`` and only mode[0] is read."""
    negative_score = tf_call(model, (positive_sample, negative_sample), int(mode[0]))
    positive_score = tf_call(model, (positive_sample, negative_sample), TF_MODE_SINGLE)
    positive_sample_loss = -(subsampling_weight * positive_score).sum() / subsampling_weight.sum()
    negative_sample_loss = -(subsampling_weight * negative_score).sum() / subsampling_weight.sum()
    return (positive_sample_loss + negative_sample_loss) / 2


class KerasAdam:
    """``tf.keras.optimizers.Adam`` defaults (run.py:111): beta1 .9, beta2 .999, **epsilon 1e-7**,
    update ``p -= lr * sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps)`` (eps outside the bias
    correction -- not the same as ``torch.optim.Adam``).  Third-party algorithm restated from
    the published Keras formula; TensorFlow is absent here."""

    def __init__(self, params, lr=1e-3, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        self.params = list(params)
        self.lr, self.b1, self.b2, self.eps = lr, beta_1, beta_2, epsilon
        self.t = 0
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]

    @torch.no_grad()
    def apply_gradients(self, grads_and_vars):
        self.t += 1
        lr = self.lr(self.t - 1) if callable(self.lr) else self.lr
        alpha = lr * math.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
        for (g, p), m, v in zip(grads_and_vars, self.m, self.v):
            m.mul_(self.b1).add_(g, alpha=1.0 - self.b1)
            v.mul_(self.b2).addcmul_(g, g, value=1.0 - self.b2)
            p.sub_(alpha * m / (v.sqrt() + self.eps))


def tf_train_step(model: OracleKGEModel, optimizer: KerasAdam, batch):
    """supervisor.py:15-28 for one replica."""
    positive_sample, negative_sample, subsampling_weight, mode = batch
    params = [model.entity_embedding, model.relation_embedding]
    if hasattr(model, "W"):
        params.append(model.W)  # TranSparse: model.py:105 (trainable); the mask is not (model.py:103)
    loss = tf_train_loss(model, positive_sample, negative_sample, subsampling_weight, mode)
    grads = torch.autograd.grad(loss, params)
    optimizer.apply_gradients(zip(grads, params))
    return loss.detach()


# ------------------------------------------------------------------------------------------
# Seeded synthetic inputs (SURVEY 8(d) "Synthetic inputs")
# ------------------------------------------------------------------------------------------
def synthetic_batch(nentity, nrelation, batch, negatives, seed, mode=None):
    g = torch.Generator().manual_seed(int(seed))
    pos = torch.stack([
        torch.randint(0, nentity, (batch,), generator=g),
        torch.randint(0, nrelation, (batch,), generator=g),
        torch.randint(0, nentity, (batch,), generator=g)], dim=1)
    neg = torch.randint(0, nentity, (batch, negatives), generator=g)
    c = torch.randint(8, 201, (batch,), generator=g).to(torch.float32)
    w = torch.sqrt(1.0 / c)
    if mode is None:
        mode = "head-batch" if seed % 2 == 0 else "tail-batch"
    return pos, neg, w, mode


def model_flags(model_name):
    """(-de, -dr, -tr) each model needs (U1 validation; InterHT: README.md:10-20)."""
    return {
        "TransE": (False, False, False),
        "DistMult": (False, False, False),
        "ComplEx": (True, True, False),
        "RotatE": (True, False, False),
        "pRotatE": (False, False, False),
        "InterHT": (True, False, True),
    }[model_name]


def make_model(model_name, nentity, nrelation, hidden_dim, gamma, seed=0, dtype=torch.float32):
    de, dr, tr = model_flags(model_name)
    torch.manual_seed(seed)
    return OracleKGEModel(model_name, nentity, nrelation, hidden_dim, gamma, de, dr, tr, dtype=dtype)


def to_double(model: OracleKGEModel) -> OracleKGEModel:
    """fp64 twin with identical weights (gamma / embedding_range stay fp32 values)."""
    de = model.entity_dim == 2 * model.hidden_dim
    dr = model.relation_dim == 2 * model.hidden_dim
    tr = model.relation_dim == 3 * model.hidden_dim
    twin = OracleKGEModel(model.model_name, model.nentity, model.nrelation, model.hidden_dim,
                          model.gamma.item(), de, dr, tr, dtype=torch.float64,
                          interht_eps=model.interht_eps)
    with torch.no_grad():
        twin.entity_embedding.copy_(model.entity_embedding.double())
        twin.relation_embedding.copy_(model.relation_embedding.double())
        if hasattr(model, "modulus"):
            twin.modulus.copy_(model.modulus.double())
    return twin


# --------------------------------------------------------------------------------------------------
# U11: the producer of the train batch -- TrainDataset / BidirectionalOneShotIterator of the PyTorch codes
# (absent submodule; SURVEY 8(a) U11 -- restated from the published algorithm; the TF twin builds the same
# generators at /root/reference/compress_data/main.py:64-90 with mode 0 = head-batch, 1 = tail-batch).
# --------------------------------------------------------------------------------------------------
def count_frequency(triples, start: int = 4):
    """Frequency of the partial triples (head, relation) and (tail, -relation-1), counts start at `start`."""
    count = {}
    for head, relation, tail in triples:
        count[(head, relation)] = count.get((head, relation), start - 1) + 1
        count[(tail, -relation - 1)] = count.get((tail, -relation - 1), start - 1) + 1
    return count


def get_true_head_and_tail(triples):
    """true_head[(relation, tail)] / true_tail[(head, relation)] as unique numpy arrays (order of np.array(set))."""
    true_head, true_tail = {}, {}
    for head, relation, tail in triples:
        true_tail.setdefault((head, relation), []).append(tail)
        true_head.setdefault((relation, tail), []).append(head)
    true_head = {k: np.array(list(set(v))) for k, v in true_head.items()}
    true_tail = {k: np.array(list(set(v))) for k, v in true_tail.items()}
    return true_head, true_tail


def subsampling_weight_of(triple, count):
    head, relation, tail = triple
    w = count[(head, relation)] + count[(tail, -relation - 1)]
    return torch.sqrt(1 / torch.Tensor([w]))  # fp32: 1/w rounded, then sqrt


def train_dataset_item(triple, mode: str, nentity: int, negative_sample_size: int, count, true_head, true_tail,
                       rng: "np.random.RandomState"):
    """One `TrainDataset.__getitem__`: draw 2N uniform ids at a time, drop the true heads / tails (np.in1d),
    until N negatives are collected.  -> (positive [3] i64, negative [N] i64, weight [1] f32, mode)."""
    head, relation, tail = triple
    weight = subsampling_weight_of(triple, count)
    negative_sample_list, size = [], 0
    while size < negative_sample_size:
        negative_sample = rng.randint(nentity, size=negative_sample_size * 2)
        if mode == "head-batch":
            mask = np.in1d(negative_sample, true_head[(relation, tail)], assume_unique=True, invert=True)
        elif mode == "tail-batch":
            mask = np.in1d(negative_sample, true_tail[(head, relation)], assume_unique=True, invert=True)
        else:
            raise ValueError("Training batch mode %s not supported" % mode)
        negative_sample = negative_sample[mask]
        negative_sample_list.append(negative_sample)
        size += negative_sample.size
    negative_sample = np.concatenate(negative_sample_list)[:negative_sample_size]
    return torch.LongTensor(triple), torch.LongTensor(negative_sample), weight, mode


# --------------------------------------------------------------------------------------------------
# Checker for the on-GPU sampler (csrc/kge_sampler.cu): Philox4x32-10 (Salmon et al., SC'11; the known-answer
# vectors of Random123's kat_vectors are checked in tests/test_oracle_golden.py) and the kernel's draw scheme.
# --------------------------------------------------------------------------------------------------
def philox4x32_10(counter, key):
    """counter: 4 uint32, key: 2 uint32 -> 4 uint32 (python ints)."""
    M0, M1, W0, W1, MASK = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85, 0xFFFFFFFF
    c0, c1, c2, c3 = [int(x) & MASK for x in counter]
    k0, k1 = [int(x) & MASK for x in key]
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        c0, c1, c2, c3 = (p1 >> 32) ^ c1 ^ k0, p1 & MASK, (p0 >> 32) ^ c3 ^ k1, p0 & MASK
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c0, c1, c2, c3


def sample_negatives_reference(positive, negative_sample_size, mode, nentity, train_triples, seed, step, first_index=0):
    """Pure-Python restatement of kge_sample_negatives: same counters, same acceptance rule -> identical ids."""
    true = set((int(h), int(r), int(t)) for h, r, t in train_triples)
    E = int(nentity)
    out = np.zeros((len(positive), negative_sample_size), dtype=np.int64)
    for b, (h, r, t) in enumerate(np.asarray(positive).tolist()):
        gidx = first_index + b
        for slot in range(negative_sample_size):
            e, done = 0, False
            for attempt in range(256):
                p = philox4x32_10((gidx & 0xFFFFFFFF, slot, attempt, step & 0xFFFFFFFF),
                                  (seed & 0xFFFFFFFF, ((seed >> 32) ^ (gidx >> 32)) & 0xFFFFFFFF))
                for half in range(2):
                    x = (p[2 * half + 1] << 32) | p[2 * half]
                    e = (x * E) >> 64
                    known = ((e, r, t) in true) if mode == "head-batch" else ((h, r, e) in true)
                    if not known:
                        done = True
                        break
                if done:
                    break
            out[b, slot] = e
    return out
