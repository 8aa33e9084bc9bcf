"""Execute the reference's TensorFlow hot-path files under ``oracle/tf_shim`` and write golden vectors.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Run HERE (the build container), where
``/root/reference`` is mounted::

    python oracle/run_reference_tf.py            # rewrites tests/golden/ref_tf_*.npz

It imports, unmodified and from where they lie,

* ``/root/reference/tensorflow_codes/model.py``      (``TFKGEModel``, ``InterHT``, ``loss``)
* ``/root/reference/tensorflow_codes/supervisor.py`` (``Trainer.train_step``)

with ``import tensorflow`` resolved to the torch-backed shim.  What is recorded:

``ref_tf_interht.npz``   weights + batch -> raw ``InterHT`` scores for the 3 gathers, and the
                         fused ``call`` outputs for modes 0 / 1 / 3 (model.py:114-224).
``ref_tf_trainstep.npz`` 3 consecutive ``Trainer.train_step`` calls (supervisor.py:13-30) with the
                         Keras-Adam shim: per-step loss and the tables after each step.
``ref_tf_transparse.npz`` the same three things for ``TFKGEModel('TranSparse', ...)`` (model.py:96-106,226-235):
                         variables incl. ``W`` and the 0/1 ``mask``, raw scores of the three gathers, ``call`` outputs,
                         3 ``Trainer.train_step`` calls (entity / relation / W after each).
``ref_tf_loss.npz``      the eager ``loss()`` (model.py:4-44) evaluated on recorded score matrices for
                         adv / non-adv x uni / subsampling x reg / no-reg.

The GPU box has no ``/root/reference``; tests there read only the committed ``.npz`` files.
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = os.environ.get("KGE_REFERENCE_ROOT", "/root/reference")
GOLDEN = os.path.join(REPO, "tests", "golden")


def _import_reference():
    shim = os.path.join(HERE, "tf_shim")
    if shim not in sys.path:
        sys.path.insert(0, shim)
    mods = {}
    for name in ("model", "supervisor"):
        path = os.path.join(REF, "tensorflow_codes", name + ".py")
        spec = importlib.util.spec_from_file_location("ref_tf_" + name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mods[name] = mod
    import tensorflow as tf  # the shim
    return mods["model"], mods["supervisor"], tf


def _batch(E, R, B, N, seed):
    g = torch.Generator().manual_seed(seed)
    pos = torch.stack([torch.randint(0, E, (B,), generator=g),
                       torch.randint(0, R, (B,), generator=g),
                       torch.randint(0, E, (B,), generator=g)], 1)
    neg = torch.randint(0, E, (B, N), generator=g)
    w = torch.sqrt(1.0 / torch.randint(8, 201, (B, 1), generator=g).float())
    return pos, neg, w


def golden_interht(ref_model, tf):
    E, R, H, B, N, gamma = 97, 7, 24, 6, 11, 24.0
    torch.manual_seed(1234)
    m = ref_model.TFKGEModel("InterHT", E, R, H, gamma, double_entity_embedding=True,
                             triple_relation_embedding=True)
    pos, neg, w = _batch(E, R, B, N, 7)
    out = {
        "E": E, "R": R, "H": H, "gamma": gamma,
        "entity_embedding": m.entity_embedding.numpy().copy(),
        "relation_embedding": m.relation_embedding.numpy().copy(),
        "positive_sample": pos.numpy(), "negative_sample": neg.numpy(),
    }
    # raw InterHT scores on the three gathers the reference builds (model.py:127-199)
    ent, rel = m.entity_embedding, m.relation_embedding
    h1 = tf.expand_dims(tf.gather(ent, pos[:, 0]), 1)
    r1 = tf.expand_dims(tf.gather(rel, pos[:, 1]), 1)
    t1 = tf.expand_dims(tf.gather(ent, pos[:, 2]), 1)
    hN = tf.reshape(tf.gather(ent, tf.reshape(neg, [-1])), [B, N, -1])
    tN = tf.reshape(tf.gather(ent, tf.reshape(neg, [-1])), [B, N, -1])
    out["score_single"] = m.InterHT(h1, r1, t1, 3).numpy()
    out["score_head_batch"] = m.InterHT(hN, r1, t1, 0).numpy()
    out["score_tail_batch"] = m.InterHT(h1, r1, tN, 1).numpy()
    for mode in (0, 1, 3):
        out["call_mode%d" % mode] = m(((pos, neg), torch.tensor(mode))).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "ref_tf_interht.npz"), **out)
    return out


def golden_trainstep(ref_model, ref_sup, tf):
    E, R, H, B, N, gamma, lr = 53, 4, 12, 5, 9, 24.0, 1e-2
    torch.manual_seed(4321)
    m = ref_model.TFKGEModel("InterHT", E, R, H, gamma, double_entity_embedding=True,
                             triple_relation_embedding=True)
    opt = tf.keras.optimizers.Adam(learning_rate=lr)
    metric = tf.keras.metrics.Sum("training_loss")
    out = {"E": E, "R": R, "H": H, "gamma": gamma, "lr": lr,
           "entity_embedding_0": m.entity_embedding.numpy().copy(),
           "relation_embedding_0": m.relation_embedding.numpy().copy()}
    batches = []
    for step in range(3):
        pos, neg, w = _batch(E, R, B, N, 100 + step)
        mode = torch.full((B,), step % 2, dtype=torch.int64)  # 0 head-batch, 1 tail-batch
        batches.append((pos, neg, w, mode))
    trainer = ref_sup.Trainer(tf.distribute.get_strategy(), batches, m, opt, metric)
    it = iter(batches)
    for step in range(3):
        pos, neg, w, mode = batches[step]
        metric.reset_states()
        trainer.train_step(it)
        out["positive_sample_%d" % step] = pos.numpy()
        out["negative_sample_%d" % step] = neg.numpy()
        out["subsampling_weight_%d" % step] = w.numpy()
        out["mode_%d" % step] = mode.numpy()
        out["loss_%d" % step] = np.float32(metric.result())
        out["entity_embedding_%d" % (step + 1)] = m.entity_embedding.numpy().copy()
        out["relation_embedding_%d" % (step + 1)] = m.relation_embedding.numpy().copy()
    np.savez_compressed(os.path.join(GOLDEN, "ref_tf_trainstep.npz"), **out)
    return out


def golden_transparse(ref_model, ref_sup, tf):
    E, R, H, B, N, gamma, lr = 61, 5, 22, 6, 9, 12.0, 1e-2
    torch.manual_seed(2468)
    m = ref_model.TFKGEModel("TranSparse", E, R, H, gamma)
    pos, neg, w = _batch(E, R, B, N, 11)
    out = {"E": E, "R": R, "H": H, "gamma": gamma, "lr": lr,
           "entity_embedding": m.entity_embedding.numpy().copy(),
           "relation_embedding": m.relation_embedding.numpy().copy(),
           "W": m.W.numpy().copy(), "mask": m.mask.numpy().copy(),
           "positive_sample": pos.numpy(), "negative_sample": neg.numpy()}
    ent, rel = m.entity_embedding, m.relation_embedding
    h1 = tf.expand_dims(tf.gather(ent, pos[:, 0]), 1)
    r1 = tf.expand_dims(tf.gather(rel, pos[:, 1]), 1)
    t1 = tf.expand_dims(tf.gather(ent, pos[:, 2]), 1)
    hN = tf.reshape(tf.gather(ent, tf.reshape(neg, [-1])), [B, N, -1])
    tN = tf.reshape(tf.gather(ent, tf.reshape(neg, [-1])), [B, N, -1])
    bW, bM = tf.gather(m.W, pos[:, 1]), tf.gather(m.mask, pos[:, 1])
    out["score_single"] = m.TranSparse(h1, r1, t1, 3, bW, bM).numpy()
    out["score_head_batch"] = m.TranSparse(hN, r1, t1, 0, bW, bM).numpy()
    out["score_tail_batch"] = m.TranSparse(h1, r1, tN, 1, bW, bM).numpy()   # [B, 1]: model.py:229 never reads tN
    for mode in (0, 1, 3):
        out["call_mode%d" % mode] = m(((pos, neg), torch.tensor(mode))).numpy()
    # three train steps (supervisor.py:13-30) on the same model
    opt = tf.keras.optimizers.Adam(learning_rate=lr)
    metric = tf.keras.metrics.Sum("training_loss")
    batches = []
    for step in range(3):
        p, n, sw = _batch(E, R, B, N, 200 + step)
        batches.append((p, n, sw, torch.full((B,), (step + 1) % 2, dtype=torch.int64)))
    trainer = ref_sup.Trainer(tf.distribute.get_strategy(), batches, m, opt, metric)
    it = iter(batches)
    for step in range(3):
        p, n, sw, mode = batches[step]
        metric.reset_states()
        trainer.train_step(it)
        out["positive_sample_%d" % step] = p.numpy()
        out["negative_sample_%d" % step] = n.numpy()
        out["subsampling_weight_%d" % step] = sw.numpy()
        out["mode_%d" % step] = mode.numpy()
        out["loss_%d" % step] = np.float32(metric.result())
        out["entity_embedding_%d" % (step + 1)] = m.entity_embedding.numpy().copy()
        out["relation_embedding_%d" % (step + 1)] = m.relation_embedding.numpy().copy()
        out["W_%d" % (step + 1)] = m.W.numpy().copy()
    np.savez_compressed(os.path.join(GOLDEN, "ref_tf_transparse.npz"), **out)
    return out


class _ScoreTable:
    """Callable with the PyTorch ``model(sample, mode=...)`` signature the eager ``loss()`` expects
    (model.py:11,20), answering from recorded score matrices."""

    def __init__(self, tf, neg_score, pos_score, ent, rel):
        self._neg = tf.Variable(neg_score, trainable=False)
        self._pos = tf.Variable(pos_score, trainable=False)
        self.entity_embedding = tf.Variable(ent, trainable=False)
        self.relation_embedding = tf.Variable(rel, trainable=False)

    def __call__(self, sample, mode="single"):
        return self._neg if isinstance(sample, tuple) else self._pos


def golden_loss(ref_model, tf):
    B, N = 8, 13
    g = torch.Generator().manual_seed(99)
    neg_score = (torch.randn(B, N, generator=g) * 3).numpy().astype(np.float32)
    pos_score = (torch.randn(B, 1, generator=g) * 3).numpy().astype(np.float32)
    w = torch.sqrt(1.0 / torch.randint(8, 201, (B,), generator=g).float()).numpy()
    ent = (torch.rand(17, 6, generator=g) - 0.5).numpy().astype(np.float32)
    rel = (torch.rand(3, 6, generator=g) - 0.5).numpy().astype(np.float32)
    out = {"negative_score": neg_score, "positive_score": pos_score, "subsampling_weight": w,
           "entity_embedding": ent, "relation_embedding": rel}
    table = _ScoreTable(tf, neg_score, pos_score, ent, rel)
    pos_idx, neg_idx = torch.zeros(B, 3, dtype=torch.int64), torch.zeros(B, N, dtype=torch.int64)
    k = 0
    cases = []
    for adv in (True, False):
        for uni in (True, False):
            for reg in (0.0, 1e-3):
                for temp in ((0.5, 1.0) if adv else (1.0,)):
                    val = ref_model.loss(table, (pos_idx, neg_idx, w, "tail-batch"), reg, adv, uni, temp)
                    val = val.numpy() if hasattr(val, "numpy") else val
                    cases.append((float(adv), float(uni), reg, temp, float(np.asarray(val).reshape(-1)[0])))
                    k += 1
    out["cases"] = np.asarray(cases, dtype=np.float64)  # columns: adv, uni, reg, temperature, loss
    np.savez_compressed(os.path.join(GOLDEN, "ref_tf_loss.npz"), **out)
    return out


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(1)
    ref_model, ref_sup, tf = _import_reference()
    a = golden_interht(ref_model, tf)
    b = golden_trainstep(ref_model, ref_sup, tf)
    c = golden_loss(ref_model, tf)
    d = golden_transparse(ref_model, ref_sup, tf)
    print("ref_tf_transparse: call_mode0", d["call_mode0"].ravel()[:3], "losses", [float(d["loss_%d" % i]) for i in range(3)])
    print("ref_tf_interht: call_mode0", a["call_mode0"].ravel()[:3])
    print("ref_tf_trainstep: losses", [float(b["loss_%d" % i]) for i in range(3)])
    print("ref_tf_loss: cases", c["cases"].shape)


if __name__ == "__main__":
    main()
