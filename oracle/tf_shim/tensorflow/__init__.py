"""Torch-backed stand-in for the TensorFlow symbols the reference's hot-path files touch.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  TensorFlow is not installed in this
image and cannot be installed (no network).  To still pin the oracle against the
reference's *own source*, ``oracle/run_reference_tf.py`` puts this directory on
``sys.path`` and imports ``/root/reference/tensorflow_codes/model.py`` and
``supervisor.py`` unmodified; every ``tf.*`` call they make lands here.

Each function states the TensorFlow semantics it stands in for (public TF 2.x API docs);
all maths runs in torch fp32 on the CPU, autograd stands in for ``GradientTape``.
Only what those two files call is provided -- this is not a TensorFlow clone.
"""
from __future__ import annotations

import contextlib
import math as _math
import types

import numpy as np
import torch
import torch.nn.functional as _F

float32 = torch.float32
int64 = torch.int64


def _to_tensor(x, dtype=None):
    if isinstance(x, torch.Tensor):
        return x if dtype is None else x.to(dtype)
    if isinstance(x, (list, tuple)) and len(x) and isinstance(x[0], (np.ndarray, torch.Tensor)):
        x = np.stack([np.asarray(e.detach() if isinstance(e, torch.Tensor) else e) for e in x])
    t = torch.as_tensor(np.asarray(x))
    if dtype is not None:
        t = t.to(dtype)
    elif t.dtype == torch.float64:
        t = t.to(torch.float32)  # TF default float is float32
    return t.as_subclass(Tensor)


class Tensor(torch.Tensor):
    """EagerTensor / Variable stand-in: torch tensor with ``.numpy()`` that detaches (TF eager
    ``.numpy()`` never carries gradient) and ``.assign``."""

    def numpy(self):
        return torch.Tensor.numpy(self.detach().as_subclass(torch.Tensor))

    def assign(self, value):
        with torch.no_grad():
            self.copy_(_to_tensor(value, self.dtype))
        return self

    # ndarray (op) Tensor: numpy defers to us because torch sets __array_priority__
    def __rmul__(self, other):
        return torch.mul(self, _to_tensor(other, self.dtype))

    def __radd__(self, other):
        return torch.add(self, _to_tensor(other, self.dtype))

    def __rsub__(self, other):
        return torch.sub(_to_tensor(other, self.dtype), self)


def Variable(initial_value, trainable=True, dtype=None, name=None):
    """``tf.Variable``: owns a copy of ``initial_value``; ``trainable`` -> watched by the tape."""
    t = _to_tensor(initial_value, dtype).detach().clone().as_subclass(Tensor)
    if trainable and t.is_floating_point():
        t.requires_grad_(True)
    t._tf_trainable = bool(trainable)
    t._tf_name = name
    return t


def zeros(shape, dtype=float32):
    return torch.zeros(tuple(int(s) for s in shape), dtype=dtype).as_subclass(Tensor)


def ones_like(x):
    return torch.ones_like(x)


class random_uniform_initializer:
    """``tf.random_uniform_initializer(minval, maxval)``: U[minval, maxval) of the asked shape."""

    def __init__(self, minval=-0.05, maxval=0.05, seed=None):
        self.minval = float(np.asarray(minval).reshape(-1)[0])
        self.maxval = float(np.asarray(maxval).reshape(-1)[0])

    def __call__(self, shape, dtype=float32):
        shape = tuple(int(s) for s in shape)
        return (torch.rand(shape, dtype=dtype) * (self.maxval - self.minval) + self.minval).as_subclass(Tensor)


def where(cond, x, y):
    return torch.where(cond, torch.as_tensor(x, dtype=torch.float32), torch.as_tensor(y, dtype=torch.float32))


def stack(values, axis=0):
    return torch.stack(list(values), dim=axis)


def equal(x, y):
    return torch.as_tensor(x) == y


def cond(pred, true_fn, false_fn):
    """``tf.cond``: eager semantics -- evaluate exactly one branch."""
    return true_fn() if bool(pred) else false_fn()


def gather(params, indices, axis=0):
    """``tf.gather(params, indices)``: rows of ``params`` along axis 0, shape indices.shape + params.shape[1:]."""
    assert axis == 0
    idx = torch.as_tensor(indices).to(torch.int64)
    out = torch.index_select(params, 0, idx.reshape(-1))
    return out.reshape(tuple(idx.shape) + tuple(params.shape[1:]))


def expand_dims(x, axis):
    return torch.unsqueeze(x, axis)


def shape(x):
    return tuple(int(s) for s in x.shape)


def reshape(x, shp):
    return torch.reshape(x, tuple(int(s) for s in shp))


def reduce_sum(x, axis=None, keepdims=False):
    if axis is None:
        return torch.sum(x)
    return torch.sum(x, dim=axis, keepdim=keepdims)


def split(value, num_or_size_splits, axis=0):
    """``tf.split`` with an int: that many equal parts (must divide)."""
    n = int(num_or_size_splits)
    size = value.shape[axis]
    assert size % n == 0, "tf.split: dimension %d not divisible by %d" % (size, n)
    return list(torch.split(value, size // n, dim=axis))


def norm(tensor, ord="euclidean", axis=None, keepdims=False):
    """``tf.norm``: ord=1 -> sum|x|; ord=2/'euclidean' -> sqrt(sum x^2); ord=p -> (sum|x|^p)^(1/p)."""
    x = torch.as_tensor(tensor)
    kw = {} if axis is None else {"dim": axis, "keepdim": keepdims}
    if ord == 1:
        return torch.sum(torch.abs(x), **kw)
    if ord in (2, "euclidean"):
        return torch.sqrt(torch.sum(x * x, **kw))
    p = float(ord)
    return torch.pow(torch.sum(torch.pow(torch.abs(x), p), **kw), 1.0 / p)


def matmul(a, b):
    return torch.matmul(a, b)


def function(fn=None, **_kw):
    """``@tf.function``: tracing is an optimisation; eager execution has the same results."""
    if fn is None:
        return lambda f: f
    return fn


class GradientTape:
    """``tf.GradientTape``: records ops on trainable variables; torch autograd already does."""

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def gradient(self, target, sources):
        return list(torch.autograd.grad(target, list(sources), allow_unused=True))


# ---- tf.nn / tf.math / tf.linalg / tf.random ---------------------------------------------
nn = types.SimpleNamespace(
    softmax=lambda logits, axis=-1: torch.softmax(logits, dim=axis),
)

math_ns = types.SimpleNamespace(
    log_sigmoid=lambda x: _F.logsigmoid(x),  # tf.math.log_sigmoid(x) = -softplus(-x)
)


def _normalize(tensor, ord="euclidean", axis=None):
    """``tf.linalg.normalize``: (tensor / norm, norm) with norm = tf.norm(keepdims=True); no epsilon."""
    n = norm(tensor, ord=ord, axis=axis, keepdims=True)
    return tensor / n, n


linalg = types.SimpleNamespace(normalize=_normalize)


def _random_uniform(shape, minval=0, maxval=None, dtype=float32):
    shape = tuple(int(s) for s in shape)
    maxval = 1 if maxval is None else maxval
    return (torch.rand(shape) * (maxval - minval) + minval).as_subclass(Tensor)


random = types.SimpleNamespace(uniform=_random_uniform)


# ---- tf.distribute ---------------------------------------------------------------------------
class _DefaultStrategy:
    num_replicas_in_sync = 1

    def run(self, fn, args=(), kwargs=None):
        return fn(*args, **(kwargs or {}))

    def scope(self):
        return contextlib.nullcontext()


distribute = types.SimpleNamespace(get_strategy=lambda: _DefaultStrategy())


# ---- tf.keras ----------------------------------------------------------------------------------
class _Model:
    """``tf.keras.Model``: ``model(x, **kw)`` dispatches to ``call``; variables are tracked by attribute."""

    def __init__(self, **kwargs):
        pass

    def __call__(self, inputs, *args, **kwargs):
        return self.call(inputs, *args, **kwargs)

    @property
    def weights(self):
        return [v for v in self.__dict__.values() if isinstance(v, Tensor) and hasattr(v, "_tf_trainable")]

    @property
    def trainable_variables(self):
        return [v for v in self.weights if v._tf_trainable]


class _Adam:
    """``tf.keras.optimizers.Adam`` (TF 2.x defaults beta_1=.9, beta_2=.999, epsilon=1e-7):
    ``lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m=b1 m+(1-b1)g; v=b2 v+(1-b2)g^2; p -= lr_t*m/(sqrt(v)+eps)``."""

    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        self.learning_rate = learning_rate
        self.beta_1, self.beta_2, self.epsilon = beta_1, beta_2, epsilon
        self.iterations = 0
        self._state = {}

    def apply_gradients(self, grads_and_vars):
        self.iterations += 1
        t = self.iterations
        lr = self.learning_rate
        if callable(lr):
            lr = float(lr(t - 1))
        alpha = lr * _math.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t)
        with torch.no_grad():
            for g, p in grads_and_vars:
                if g is None:
                    continue
                st = self._state.setdefault(id(p), (torch.zeros_like(p), torch.zeros_like(p)))
                m, v = st
                m.mul_(self.beta_1).add_(g, alpha=1.0 - self.beta_1)
                v.mul_(self.beta_2).addcmul_(g, g, value=1.0 - self.beta_2)
                p.sub_(alpha * m / (v.sqrt() + self.epsilon))


class _Sum:
    """``tf.keras.metrics.Sum``."""

    def __init__(self, name=None, dtype=None):
        self.total = 0.0

    def update_state(self, value):
        self.total += float(torch.as_tensor(value).detach().sum())

    def result(self):
        return self.total

    def reset_states(self):
        self.total = 0.0


keras = types.SimpleNamespace(
    Model=_Model,
    optimizers=types.SimpleNamespace(Adam=_Adam),
    metrics=types.SimpleNamespace(Sum=_Sum),
)

math = math_ns  # ``tf.math`` (the stdlib module is imported as ``_math``)
