"""TF-shaped twin of the reference's TensorFlow trainer, backed by the same sm_100a kernels.

Mirrors /root/reference/tensorflow_codes:

* ``TFKGEModel(model_name, nentity, nrelation, hidden_dim, gamma, double_entity_embedding,
  double_relation_embedding, triple_relation_embedding)`` -- constructor of ``model.py:49-52``;
  ``model((sample, mode))`` with the wire ints 0 = head-batch, 1 = tail-batch, 3 = single
  (``compress_data/main.py:65,69``; ``supervisor.py:18``) returns the fused per-positive terms ``[B,1]`` of
  ``model.py:114-205``: ``log_sigmoid(score)`` for mode 3, ``sum_n softmax(s)_n * log_sigmoid(-s_n)`` otherwise
  (temperature 1, gradient flowing through the softmax, ``model.py:168-171``).
* ``Trainer(strategy, dataloader, model, optimizer, metrics)`` with ``train_step(data_iter)`` and
  ``training(steps_per_tpu_call, epochs, steps_per_epoch)`` -- ``supervisor.py:5-58``.
* ``KerasAdam`` -- ``tf.keras.optimizers.Adam`` defaults (``run.py:111``): epsilon 1e-7 outside the bias
  correction; ``learning_rate`` may be a float, a callable schedule, and supports ``.assign`` (the Checker
  zeroes it: ``Checker/trainer_checker.py:10-14``).

No TensorFlow: tensors are torch CUDA tensors.  Deviation kept on purpose: the reference overwrites the
``-dr`` width in its ``-tr`` branch (``model.py:65-78``), which makes ComplEx unbuildable there; here ``-dr``
is honoured when ``-tr`` is off.
"""
from __future__ import annotations

import time
from collections import defaultdict

import torch

from . import _lib, ops
from .model import KGEModel
from .transparse import TranSparseModel

TF_MODE_HEAD_BATCH, TF_MODE_TAIL_BATCH, TF_MODE_SINGLE = 0, 1, 3


class _CallTerms(torch.autograd.Function):
    """The row reductions of ``TFKGEModel.call`` (model.py:145,168-171): kge_tf_call_terms and its backward."""

    @staticmethod
    def forward(ctx, score, single):
        score = score.contiguous().float()
        B, N = score.shape
        out = torch.empty(B, 1, dtype=torch.float32, device=score.device)
        with torch.cuda.device(score.device):
            _lib.check(_lib.load().kge_tf_call_terms(ops._ptr(score), B, N, int(single), ops._ptr(out), ops._stream()),
                       "kge_tf_call_terms")
        ctx.save_for_backward(score)
        ctx.single = single
        return out

    @staticmethod
    def backward(ctx, gout):
        (score,) = ctx.saved_tensors
        B, N = score.shape
        ds = torch.empty_like(score)
        gout = gout.contiguous().float()
        with torch.cuda.device(score.device):
            _lib.check(_lib.load().kge_tf_call_terms_backward(ops._ptr(score), ops._ptr(gout), B, N, int(ctx.single),
                                                              ops._ptr(ds), ops._stream()), "kge_tf_call_terms_backward")
        return ds, None


class TFKGEModel:
    def __init__(self, model_name, nentity, nrelation, hidden_dim, gamma,
                 double_entity_embedding=False, double_relation_embedding=False,
                 triple_relation_embedding=False, device="cuda", project_tail=False, **kwargs):
        if model_name == "TranSparse":
            # model.py:96-106: W [R, D, D] + the fixed 0/1 mask; :227 needs entity_dim == relation_dim.
            # project_tail=False keeps model.py:229 as written (the head is projected twice; see transparse.py)
            if double_entity_embedding or double_relation_embedding or triple_relation_embedding:
                raise ValueError("TranSparse multiplies [.., entity_dim] by [relation_dim, relation_dim] (model.py:227): "
                                 "no -de / -dr / -tr")
            self.kge = TranSparseModel(nentity, nrelation, hidden_dim, gamma, project_tail=project_tail).to(device)
        else:
            self.kge = KGEModel(model_name, nentity, nrelation, hidden_dim, gamma, double_entity_embedding,
                                double_relation_embedding, triple_relation_embedding).to(device)
        self.model_name, self.nentity, self.nrelation, self.hidden_dim = model_name, nentity, nrelation, hidden_dim
        self.epsilon = 2.0
        self.entity_dim, self.relation_dim = self.kge.entity_dim, self.kge.relation_dim
        if model_name == "InterHT":
            self.u = 1  # model.py:94-95

    # variables under the names the Checker maps (Checker/copy_weights.py:5-30)
    @property
    def entity_embedding(self):
        return self.kge.entity_embedding

    @property
    def relation_embedding(self):
        return self.kge.relation_embedding

    @property
    def W(self):
        return self.kge.W          # TranSparse only (model.py:105)

    @property
    def mask(self):
        return self.kge.mask       # TranSparse only (model.py:103)

    @property
    def gamma(self):
        return self.kge.gamma

    @property
    def embedding_range(self):
        return self.kge.embedding_range

    @property
    def trainable_variables(self):
        return [p for p in self.kge.parameters() if p.requires_grad]

    @property
    def weights(self):
        return list(self.kge.parameters())

    def state_dict(self):
        return self.kge.state_dict()

    def load_state_dict(self, sd, strict=True):
        return self.kge.load_state_dict(sd, strict=strict)

    @staticmethod
    def mode_name(mode) -> str:
        m = int(mode)
        if m == TF_MODE_SINGLE:
            return "single"
        return "head-batch" if m == TF_MODE_HEAD_BATCH else "tail-batch"  # model.py:124: anything but 0 is tail

    def call(self, sample, training=True, **kwargs):
        sample, mode = sample
        positive_sample, negative_sample = sample
        name = self.mode_name(mode)
        if name == "single":
            return _CallTerms.apply(self.kge(positive_sample, "single"), True)   # model.py:144-145
        s = self.kge((positive_sample, negative_sample), name)                   # model.py:148-199
        return _CallTerms.apply(s, False)                                        # model.py:168-171 / 195-198

    __call__ = call


class _Scalar:
    """``optimizer.learning_rate`` stand-in: float or schedule, with ``assign`` / ``numpy``."""

    def __init__(self, value):
        self.value = value

    def assign(self, value):
        self.value = value

    def numpy(self):
        return self.value

    def __call__(self, step):
        return float(self.value(step)) if callable(self.value) else float(self.value)


class KerasAdam:
    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        self.learning_rate = _Scalar(learning_rate)
        self.beta_1, self.beta_2, self.epsilon = beta_1, beta_2, epsilon
        self.iterations = 0
        self.state = defaultdict(dict)  # same per-parameter layout as torch.optim.Adam (step / exp_avg / exp_avg_sq)

    @property
    def param_groups(self):
        return [{"lr": self.learning_rate(self.iterations), "betas": (self.beta_1, self.beta_2), "eps": self.epsilon}]


class _DefaultStrategy:
    """``tf.distribute.get_strategy()`` of ``run.py:14``: one replica."""
    num_replicas_in_sync = 1

    def run(self, fn, args=()):
        return fn(*args)


class Trainer:
    def __init__(self, strategy, dataloader, model, optimizer, metrics):
        self.dataloader = dataloader
        self.model = model
        self.optimizer = optimizer
        self.metrics = metrics
        self.strategy = strategy if strategy is not None else _DefaultStrategy()

    def train_step(self, data_iter):
        """supervisor.py:13-30: two model calls, subsampling-weighted means, /2, gradient, Adam -- one C call
        (``kge_train_step`` with attached softmax weights, temperature 1, Keras epsilon placement)."""
        batch = data_iter if isinstance(data_iter, (tuple, list)) else next(data_iter)
        positive_sample, negative_sample, subsampling_weight, mode = batch
        kge = self.model.kge
        name = TFKGEModel.mode_name(mode.reshape(-1)[0] if hasattr(mode, "reshape") else mode)
        if name == "single":
            raise ValueError("mode %s not supported" % name)
        if self.model.model_name == "TranSparse":
            loss = kge.tf_train_step(self.optimizer, positive_sample, negative_sample, subsampling_weight, name)
            self.optimizer.iterations += 1
            if self.metrics is not None:
                self.metrics.update_state(loss * self.strategy.num_replicas_in_sync)
            return {"loss": loss}
        group = self.optimizer.param_groups[0]
        out = kge._fused_step(self.optimizer, group, positive_sample, negative_sample,
                              subsampling_weight.reshape(-1), name, True, 1.0, False, 0.0,
                              detach_weights=False, keras_adam=True)
        self.optimizer.iterations += 1
        loss = out[2] if isinstance(out, torch.Tensor) else torch.tensor(float(out[2]))
        if self.metrics is not None:
            self.metrics.update_state(loss * self.strategy.num_replicas_in_sync)
        return {"loss": loss}

    def training(self, steps_per_tpu_call, epochs, steps_per_epoch):
        """supervisor.py:32-58, including its step accounting (one real step per call, counted as
        ``steps_per_tpu_call``)."""
        step = 0
        epoch = 0
        epoch_start_time = time.time()
        iteration_data = iter(self.dataloader)
        while True:
            self.train_step(iteration_data)
            step += steps_per_tpu_call
            print("=", end="", flush=True)
            epoch_time = time.time() - epoch_start_time
            print("\nEPOCH {:d}/{:d}".format(epoch + 1, epochs))
            print("time: {:0.1f}s".format(epoch_time),
                  "loss: {:0.4f}".format(round(float(self.metrics.result()), 4)), flush=True)
            epoch = step // steps_per_epoch
            epoch_start_time = time.time()
            self.metrics.reset_states()
            if epoch >= epochs:
                break
        print("DONE")


class SumMetric:
    """``tf.keras.metrics.Sum`` (run.py:114)."""

    def __init__(self, name=None, dtype=None):
        self.total = 0.0

    def update_state(self, value):
        self.total += float(value)

    def result(self):
        return self.total

    def reset_states(self):
        self.total = 0.0
