"""theano.tensor stand-in: what FNN_wnzh.py:133-174 calls (see the package docstring).  TEST INFRASTRUCTURE."""
import torch

from _sym import Sym
import theano as _th


def matrix(name=None, dtype=None):
    return Sym(None, (), name)


def vector(name=None, dtype=None):
    return Sym(None, (), name)


scalar = vector


def dot(a, b):
    return Sym(lambda x, y: torch.matmul(x, y), (a, b))


def exp(a):
    return Sym(torch.exp, (a,))


def tanh(a):
    return Sym(torch.tanh, (a,))


def log(a):
    return Sym(torch.log, (a,))


def _sum(self, axis=None):
    return Sym((lambda x: x.sum()) if axis is None else (lambda x: x.sum(axis)), (self,))


Sym.sum = _sum


def grad(cost, wrt):
    single = not isinstance(wrt, (list, tuple))
    wl = [wrt] if single else list(wrt)
    gs = [_th.GradNode(cost, wl, i) for i in range(len(wl))]
    return gs[0] if single else gs


from . import shared_randomstreams  # noqa: E402,F401
