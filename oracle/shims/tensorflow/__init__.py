"""Minimal `tensorflow` (0.x API) stand-in for running /root/reference/python/{FM,LR,FNN,FNN_IP_L3,FNN_IP_L5,FNN_IP_L7,
FNN_OP_L3,tf_util,baseline,ipinyou}.py here -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

TensorFlow 0.x (unpinned, most likely 0.8-0.9) is not installed and cannot be.  This module implements the
~30 API points those files call, with the documented TF 0.x semantics, on the lazy graph of oracle/shims/_sym.py:
  * tf.concat(axis, values) is axis-FIRST; tf.slice(t, begin, size); SparseTensor(indices, values, shape=);
  * embedding_lookup_sparse(params, sp_ids, sp_weights, combiner='sum'): out[r] = sum over the entries of sparse row r of
    weight * params[id]; rows = last row index + 1; an id outside params raises (TF: InvalidArgument);
  * sigmoid_cross_entropy_with_logits(logits, targets) = max(x,0) - x*z + log(1 + exp(-|x|));  l2_loss = sum(t^2)/2;
  * nn.dropout(x, keep_prob) = x * (1/keep_prob) * floor(keep_prob + U[0,1)); TF leaves it unseeded, so the binary mask
    comes from DROPOUT_HOOK(shape, keep_prob, k) when set (k = index of the dropout op in graph order) and every mask
    is logged in LAST_RUN['dropout'][k];
  * optimizers: GradientDescent, Adam (lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m, v; theta -= lr_t*m/(sqrt(v)+eps)) and Ftrl
    (ApplyFtrl, lr_power -0.5, accumulators 0.1, l1 = l2 = 0) over EVERY variable of the graph, dense (every model adds
    a dense term or slices the concatenated table, so TF densifies the embedding gradients);
  * Session.run evaluates all tensor fetches from the pre-update variables, then applies the train op;
  * tf.random_uniform / random_normal initialisers: TF's Philox streams cannot be reproduced, numpy PCG64(seed) is used.
Arithmetic runs in float64 (set SHIM_DTYPE = torch.float32 before building a graph for TF's own precision).
"""
import contextlib

import numpy as np
import torch

import _sym
from _sym import Env, Sym, as_tensor, evaluate, gradients

SHIM_DTYPE = torch.float64
DROPOUT_HOOK = None
LAST_RUN = {}

float32, float64, int64, int32 = "float32", "float64", "int64", "int32"

_graph_stack = []
_session_stack = []


def _dt():
    _sym.DTYPE = SHIM_DTYPE
    return SHIM_DTYPE


# ---------------------------------------------------------------------------------------------- graph / session
class Graph(object):
    def __init__(self):
        self.variables = []
        self.dropouts = []

    @contextlib.contextmanager
    def as_default(self):
        _graph_stack.append(self)
        try:
            yield self
        finally:
            _graph_stack.pop()


_default_graph = Graph()


def get_default_graph():
    return _graph_stack[-1] if _graph_stack else _default_graph


class GPUOptions(object):
    def __init__(self, **kw):
        pass


class ConfigProto(object):
    def __init__(self, **kw):
        pass


class Operation(object):
    def __init__(self, fn, name="op"):
        self.fn, self.name = fn, name

    def run(self, feed_dict=None, session=None):
        (session or _session_stack[-1]).run(self, feed_dict)


class Session(object):
    def __init__(self, graph=None, config=None):
        self.graph = graph or get_default_graph()
        self._cm = None

    def __enter__(self):
        _session_stack.append(self)
        self._cm = self.graph.as_default()
        self._cm.__enter__()
        return self

    def __exit__(self, *a):
        self._cm.__exit__(*a)
        _session_stack.pop()
        return False

    def close(self):
        pass

    def run(self, fetches, feed_dict=None):
        single = not isinstance(fetches, (list, tuple))
        fl = [fetches] if single else list(fetches)
        env = Env()
        _dt()
        for ph, val in (feed_dict or {}).items():
            env.bind(ph, ph.convert(val))
        for v in self.graph.variables:
            if v.value is not None:
                env.bind(v, v.value)
        LAST_RUN.clear()
        LAST_RUN["dropout"] = {}
        env.dropout_log = LAST_RUN["dropout"]
        _ENV[0] = env
        try:
            with torch.no_grad():
                vals = [evaluate(f, env) if isinstance(f, Sym) else None for f in fl]
                for f in fl:
                    if isinstance(f, Operation):
                        f.fn(env)
        finally:
            _ENV[0] = None
        out = [None if v is None else _np(v) for v in vals]
        return out[0] if single else out


_ENV = [None]


def _np(t):
    a = t.detach().cpu().numpy()
    if a.dtype == np.float64 and SHIM_DTYPE == torch.float32:
        a = a.astype(np.float32)
    return a.copy()


def _eval(self, feed_dict=None, session=None):
    return (session or _session_stack[-1]).run(self, feed_dict)


Sym.eval = _eval


# ---------------------------------------------------------------------------------------------- leaves
class Placeholder(Sym):
    def __init__(self, dtype, shape=None, name=None):
        Sym.__init__(self, None, (), name or "placeholder")
        self.dtype, self.shape = dtype, (None if shape is None else list(shape))

    def convert(self, val):
        a = np.asarray(val)
        if self.shape is not None:
            want = [int(s) for s in self.shape]
            if list(a.shape) != want:
                raise ValueError("Cannot feed value of shape %r for Tensor %r, which has shape %r" % (a.shape, self.name, tuple(want)))
        if "int" in self.dtype:
            return torch.as_tensor(a.astype(np.int64))
        return torch.as_tensor(a.astype(np.float64)).to(_dt())


def placeholder(dtype, shape=None, name=None):
    return Placeholder(dtype, shape, name)


class _Init(object):
    """deferred initial value (tf.zeros / random_uniform / random_normal)"""

    def __init__(self, make):
        self.make = make


def zeros(shape, dtype=float32):
    return _Init(lambda: np.zeros(shape, dtype=np.float64))


def random_uniform(shape, minval=0.0, maxval=1.0, dtype=float32, seed=None):
    return _Init(lambda: np.random.Generator(np.random.PCG64(seed)).uniform(minval, maxval, size=shape).astype(np.float32))


def random_normal(shape, mean=0.0, stddev=1.0, dtype=float32, seed=None):
    return _Init(lambda: np.random.Generator(np.random.PCG64(seed)).normal(mean, stddev, size=shape).astype(np.float32))


class Variable(Sym):
    def __init__(self, initial_value, name=None, trainable=True):
        Sym.__init__(self, None, (), name or "Variable")
        self.initial, self.value, self.trainable = initial_value, None, trainable
        self.slots = {}
        get_default_graph().variables.append(self)

    def initialize(self):
        v = self.initial.make() if isinstance(self.initial, _Init) else np.asarray(self.initial)
        self.value = torch.as_tensor(np.asarray(v, dtype=np.float64)).to(_dt()).clone()


def initialize_all_variables():
    g = get_default_graph()

    def fn(env):
        for v in g.variables:
            v.initialize()
    return Operation(fn, "init")


# ---------------------------------------------------------------------------------------------- ops
def _op(fn, *args):
    return Sym(fn, args)


def concat(concat_dim, values, name=None):
    return _op(lambda vs: torch.cat(list(vs), dim=concat_dim), list(values))


def slice(input_, begin, size, name=None):          # noqa: A001  (the reference calls tf.slice)
    begin, size = [int(b) for b in begin], [int(s) for s in size]

    def fn(t):
        idx = tuple(np.s_[b:(t.shape[k] if s == -1 else b + s)] for k, (b, s) in enumerate(zip(begin, size)))
        for k, (b, s) in enumerate(zip(begin, size)):
            if b < 0 or (s != -1 and b + s > t.shape[k]):
                raise ValueError("slice [%d, %d) out of range for dimension %d of size %d" % (b, b + s, k, t.shape[k]))
        return t[idx]
    return _op(fn, input_)


def reshape(tensor, shape, name=None):
    shape = [int(s) for s in shape]
    return _op(lambda t: t.reshape(shape), tensor)


def matmul(a, b, name=None):
    def fn(x, y):
        if x.dim() != 2 or y.dim() != 2 or x.shape[1] != y.shape[0]:
            raise ValueError("Dimensions must be equal: matmul of shapes %r and %r" % (tuple(x.shape), tuple(y.shape)))
        return x @ y
    return _op(fn, a, b)


def transpose(a, perm=None, name=None):
    return _op(lambda t: t.t() if perm is None else t.permute(*perm), a)


def identity(a, name=None):
    return _op(lambda t: t, a)


def square(a, name=None):
    return _op(lambda t: t * t, a)


def sigmoid(a, name=None):
    return _op(torch.sigmoid, a)


def tanh(a, name=None):
    return _op(torch.tanh, a)


def reduce_sum(a, reduction_indices=None, keep_dims=False, name=None):
    if reduction_indices is None:
        return _op(lambda t: t.sum(), a)
    return _op(lambda t: t.sum(dim=reduction_indices, keepdim=keep_dims), a)


def reduce_mean(a, reduction_indices=None, keep_dims=False, name=None):
    if reduction_indices is None:
        return _op(lambda t: t.mean(), a)
    return _op(lambda t: t.mean(dim=reduction_indices, keepdim=keep_dims), a)


class SparseTensor(object):
    def __init__(self, indices, values, shape=None):
        self.indices = np.asarray(indices, dtype=np.int64).reshape(-1, 2) if not isinstance(indices, Sym) else indices
        self.values, self.shape = values, shape


class _NN(object):
    @staticmethod
    def relu(a, name=None):
        return _op(torch.relu, a)

    @staticmethod
    def l2_loss(t, name=None):
        return _op(lambda x: (x * x).sum() / 2, t)

    @staticmethod
    def sigmoid_cross_entropy_with_logits(logits, targets, name=None):
        def fn(x, z):
            z = z.to(x.dtype)
            return torch.clamp(x, min=0) - x * z + torch.log1p(torch.exp(-torch.abs(x)))
        return _op(fn, logits, targets)

    @staticmethod
    def dropout(x, keep_prob, noise_shape=None, seed=None, name=None):
        g = get_default_graph()
        k = len(g.dropouts)
        g.dropouts.append(k)
        rs = np.random.RandomState(0xD0 + k)

        def fn(t):
            masks = _ENV[0].dropout_log           # one draw per Session.run: the gradient pass re-evaluates the graph
            if k not in masks:
                if DROPOUT_HOOK is not None:
                    masks[k] = np.asarray(DROPOUT_HOOK(tuple(t.shape), keep_prob, k), dtype=np.float64)
                else:
                    masks[k] = np.floor(keep_prob + rs.uniform(size=tuple(t.shape)))
            return t * (1.0 / keep_prob) * torch.as_tensor(masks[k]).to(t.dtype)
        return _op(fn, x)

    @staticmethod
    def embedding_lookup_sparse(params, sp_ids, sp_weights, partition_strategy="mod", name=None, combiner="mean"):
        if combiner != "sum":
            raise NotImplementedError("the reference only uses combiner='sum'")
        seg = torch.as_tensor(sp_ids.indices[:, 0])
        if np.any(np.diff(sp_ids.indices[:, 0]) < 0):
            raise ValueError("segment ids are not sorted")
        rows = int(sp_ids.indices[:, 0].max()) + 1

        def fn(p, ids, w):
            ids = ids.reshape(-1).to(torch.int64)
            if ids.numel() and (int(ids.min()) < 0 or int(ids.max()) >= p.shape[0]):
                raise IndexError("InvalidArgument: indices = %d is not in [0, %d)" % (int(ids.max() if ids.max() >= p.shape[0] else ids.min()), p.shape[0]))
            e = p.index_select(0, ids) * w.reshape(-1, 1).to(p.dtype)
            out = torch.zeros(rows, p.shape[1], dtype=p.dtype)
            return out.index_add(0, seg, e)
        return _op(fn, params, sp_ids.values, sp_weights.values)


nn = _NN()


# ---------------------------------------------------------------------------------------------- optimizers
class _Optimizer(object):
    def minimize(self, loss, var_list=None):
        g = get_default_graph()
        opt = self

        def fn(env):
            vs = [v for v in (var_list or g.variables) if v.trainable]
            grads = gradients(loss, vs, env)
            opt.begin()
            for v, gr in zip(vs, grads):
                v.value = opt.apply(v, v.value, gr.to(v.value.dtype)).detach()
            LAST_RUN["grads"] = {v.name: gr for v, gr in zip(vs, grads)}
        return Operation(fn, "train")

    def begin(self):
        pass


class _Train(object):
    class GradientDescentOptimizer(_Optimizer):
        def __init__(self, learning_rate):
            self.lr = learning_rate

        def apply(self, var, theta, g):
            return theta - self.lr * g

    class AdamOptimizer(_Optimizer):
        def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8):
            self.lr, self.b1, self.b2, self.eps, self.t = learning_rate, beta1, beta2, epsilon, 0

        def begin(self):
            self.t += 1

        def apply(self, var, theta, g):
            m = var.slots.setdefault("m", torch.zeros_like(theta))
            v = var.slots.setdefault("v", torch.zeros_like(theta))
            lr_t = self.lr * np.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
            m.mul_(self.b1).add_((1 - self.b1) * g)
            v.mul_(self.b2).add_((1 - self.b2) * g * g)
            return theta - lr_t * m / (torch.sqrt(v) + self.eps)

    class FtrlOptimizer(_Optimizer):
        def __init__(self, learning_rate, learning_rate_power=-0.5, initial_accumulator_value=0.1,
                     l1_regularization_strength=0.0, l2_regularization_strength=0.0):
            self.lr, self.p, self.init = learning_rate, learning_rate_power, initial_accumulator_value
            self.l1, self.l2 = l1_regularization_strength, l2_regularization_strength

        def apply(self, var, theta, g):
            n = var.slots.setdefault("accum", torch.full_like(theta, self.init))
            z = var.slots.setdefault("linear", torch.zeros_like(theta))
            n_new = n + g * g
            z.add_(g - (n_new ** (-self.p) - n ** (-self.p)) / self.lr * theta)
            quad = n_new ** (-self.p) / self.lr + 2.0 * self.l2
            out = torch.where(z.abs() > self.l1, (self.l1 * torch.sign(z) - z) / quad, torch.zeros_like(z))
            n.copy_(n_new)
            return out


train = _Train()
