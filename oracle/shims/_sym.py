"""Core of the two engine shims (oracle/shims/theano, oracle/shims/tensorflow) -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

The reference builds a symbolic graph once and evaluates it many times.  A `Sym` is a lazy node (function + arguments);
`evaluate` walks the graph starting from what an `Env` binds (placeholders, variables, random draws), so a variable
that is bound cuts the graph above it -- exactly what `theano.function(inputs=[x, ...])` does with a non-root `x`
(FNN_wnzh.py:144-145, 177-178).  Values are torch tensors (float64 unless a shim says otherwise); gradients come from
torch.autograd on a re-evaluation in which the differentiation targets are leaves.
"""
import numpy as np
import torch

DTYPE = torch.float64


def as_tensor(v):
    if isinstance(v, torch.Tensor):
        return v
    a = np.asarray(v)
    if a.dtype.kind in "iub":
        return torch.as_tensor(a.astype(np.int64))
    return torch.as_tensor(a.astype(np.float64)).to(DTYPE)


def _promote(a, b):
    a, b = as_tensor(a), as_tensor(b)
    if a.dtype.is_floating_point and not b.dtype.is_floating_point:
        b = b.to(a.dtype)
    elif b.dtype.is_floating_point and not a.dtype.is_floating_point:
        a = a.to(b.dtype)
    elif a.dtype != b.dtype and a.dtype.is_floating_point:
        wide = torch.promote_types(a.dtype, b.dtype)
        a, b = a.to(wide), b.to(wide)
    return a, b


def _add(a, b):
    a, b = _promote(a, b)
    return a + b


def _sub(a, b):
    a, b = _promote(a, b)
    return a - b


def _mul(a, b):
    a, b = _promote(a, b)
    return a * b


def _div(a, b):
    a, b = _promote(a, b)
    if not a.dtype.is_floating_point:
        a, b = a.to(DTYPE), b.to(DTYPE)
    return a / b


def _pow(a, b):
    a, b = _promote(a, b)
    return a ** b


class Sym(object):
    __array_priority__ = 1000        # numpy scalars / arrays on the left defer to the reflected operators

    def __init__(self, fn=None, args=(), name=None):
        self.fn, self.args, self.name = fn, tuple(args), name

    def _bin(self, other, op, swap=False):
        a, b = (other, self) if swap else (self, other)
        return type(self)._wrap(Sym(op, (a, b)))

    @classmethod
    def _wrap(cls, s):
        return s

    def __add__(self, o): return self._bin(o, _add)
    def __radd__(self, o): return self._bin(o, _add, True)
    def __sub__(self, o): return self._bin(o, _sub)
    def __rsub__(self, o): return self._bin(o, _sub, True)
    def __mul__(self, o): return self._bin(o, _mul)
    def __rmul__(self, o): return self._bin(o, _mul, True)
    def __truediv__(self, o): return self._bin(o, _div)
    def __rtruediv__(self, o): return self._bin(o, _div, True)
    def __pow__(self, o): return self._bin(o, _pow)
    def __neg__(self): return Sym(lambda a: -a, (self,))
    def __getitem__(self, idx): return Sym(lambda a: a[idx], (self,))
    def __hash__(self): return id(self)
    def __eq__(self, o): return self is o
    def __iter__(self): raise TypeError("symbolic value is not iterable")

    def __repr__(self):
        return "<%s %s>" % (type(self).__name__, self.name or hex(id(self)))


def _syms_of(a):
    if isinstance(a, Sym):
        return (a,)
    if isinstance(a, (list, tuple)):
        return tuple(x for x in a if isinstance(x, Sym))
    return ()


class Env(object):
    """bound: id(Sym) -> tensor for the true inputs of one evaluation; cache: values computed from them."""

    def __init__(self, bound=None):
        self.bound = dict(bound or {})
        self.cache = {}

    def bind(self, sym, value):
        self.bound[id(sym)] = value

    def get(self, sym):
        k = id(sym)
        if k in self.bound:
            return self.bound[k]
        return self.cache.get(k)


def evaluate(node, env):
    """Value of `node` under `env` (iterative post-order walk; bound nodes are never expanded)."""
    if not isinstance(node, Sym):
        if isinstance(node, (list, tuple)) and _syms_of(node):
            return [evaluate(x, env) for x in node]
        return as_tensor(node)
    stack = [node]
    while stack:
        n = stack[-1]
        if env.get(n) is not None:
            stack.pop()
            continue
        if n.fn is None:
            raise RuntimeError("unbound symbolic input %r (missing from the feed / the function inputs)" % (n,))
        pending = [x for a in n.args for x in _syms_of(a) if env.get(x) is None]
        if pending:
            stack.extend(pending)
            continue
        vals = []
        for a in n.args:
            if isinstance(a, Sym):
                vals.append(env.get(a))
            elif isinstance(a, (list, tuple)) and _syms_of(a):
                vals.append([env.get(x) if isinstance(x, Sym) else x for x in a])
            else:
                vals.append(a)
        env.cache[id(n)] = n.fn(*vals)
        stack.pop()
    return env.get(node)


def reachable(nodes, stop=()):
    """every Sym reachable from `nodes`; the walk does not go above the nodes in `stop` (bound inputs)."""
    stop = set(id(s) for s in stop)
    seen, out, stack = set(), [], list(nodes)
    while stack:
        n = stack.pop()
        if not isinstance(n, Sym):
            if isinstance(n, (list, tuple)):
                stack.extend(n)
            continue
        if id(n) in seen:
            continue
        seen.add(id(n))
        out.append(n)
        if id(n) not in stop:
            stack.extend(n.args)
    return out


def gradients(cost, wrt, env):
    """d cost / d each of `wrt` at the values `env` binds.  The targets become leaves of a fresh evaluation, so a target
    that is itself computed from other nodes is differentiated as an input: the graph above it is cut."""
    leaves = [evaluate(w, env).detach().clone().to(DTYPE).requires_grad_(True) for w in wrt]
    sub = Env(env.bound)
    for w, v in zip(wrt, leaves):
        sub.bind(w, v)
    with torch.enable_grad():
        c = evaluate(cost, sub)
        gs = torch.autograd.grad(c, leaves, allow_unused=True)
    return [torch.zeros_like(l) if g is None else g for g, l in zip(gs, leaves)]
