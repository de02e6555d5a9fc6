"""Minimal `theano` stand-in for running /root/reference/python/FNN_wnzh.py here -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Theano (0.7 / 0.8 era, unpinned) is not installed and cannot be.  This module provides only what
FNN_wnzh.py:7-11,133-183 and dl_utils.py use: shared variables, theano.function with `updates`, config.floatX, and
(theano.tensor) matrix / vector / dot / exp / tanh / log / grad, (shared_randomstreams) RandomStreams.binomial.
Semantics kept from Theano:
  * float64 everywhere (config.floatX default), integer inputs are upcast;
  * function(inputs=[v, ...]) may name a NON-ROOT variable: the graph above it is cut (FNN_wnzh.py:144-145 rebinds `x`
    to `x * r0[0]` before compiling, so the input-dropout mask r0 is not part of the compiled function);
  * outputs are computed from the PRE-update shared values; updates are applied after all outputs;
  * every RandomStreams op owns a numpy RandomState seeded from RandomState(seed).randint(2 ** 30) in creation order and
    draws once per function call when it is part of that function's graph (recollection of
    theano.tensor.shared_randomstreams; the draws are logged so a comparison never depends on this rule).
Every function call is appended to CALL_LOG: dict(fn=name, inputs=[...], random={op name: draw}, outputs=[...], t0=, t1=) with
t0 / t1 = time.perf_counter() at entry / exit of the call (bench.py times the script's own loop from them).  BUDGET_S (seconds,
default None) bounds a run: the clock starts at the script's FIRST function call (imports and the model-file parse are not
charged) and the first call after the budget raises BudgetExceeded, which the runner (oracle/ref_run.py) catches -- the
reference script has no way to stop its 100-epoch loop early.
"""
import time

import numpy as np
import torch

from _sym import DTYPE, Env, Sym, as_tensor, evaluate, gradients, reachable


class _Config(object):
    floatX = "float64"


config = _Config()
CALL_LOG = []
BUDGET_S = None
_FIRST_CALL = [None]


class BudgetExceeded(Exception):
    pass


class SharedVariable(Sym):
    def __init__(self, value, name=None):
        Sym.__init__(self, None, (), name)
        self.value = as_tensor(value).to(DTYPE).clone()

    def get_value(self, borrow=False):
        v = self.value.numpy().copy()
        return v if v.ndim else float(v)

    def set_value(self, v, borrow=False):
        self.value = as_tensor(v).to(DTYPE).clone()


def shared(value, name=None, **kw):
    if isinstance(value, np.random.RandomState):
        raise TypeError("random state variables are created by RandomStreams")
    return SharedVariable(value, name)


class GradNode(Sym):
    """one element of T.grad(cost, wrt): evaluated by the owning Function (needs the whole wrt list at once)"""

    def __init__(self, cost, wrt, index):
        Sym.__init__(self, None, (), "grad%d" % index)
        self.cost, self.wrt, self.index = cost, wrt, index


class RandomNode(Sym):
    def __init__(self, rstate, draw, name):
        Sym.__init__(self, None, (), name)
        self.rstate, self.draw = rstate, draw


class Function(object):
    _count = 0

    def __init__(self, inputs, outputs, updates=None, name=None):
        self.inputs = list(inputs)
        self.single = not isinstance(outputs, (list, tuple))
        self.outputs = [outputs] if self.single else list(outputs)
        self.updates = list(updates.items()) if isinstance(updates, dict) else list(updates or ())
        Function._count += 1
        self.name = name or "function%d" % Function._count
        roots = self.outputs + [e for _, e in self.updates]
        # grad nodes hide their cost graph: walk through them too (the cut at the inputs applies there as well)
        nodes, frontier = [], roots
        seen = set()
        while frontier:
            new = [n for n in reachable(frontier, stop=self.inputs) if id(n) not in seen]
            for n in new:
                seen.add(id(n))
            nodes += new
            frontier = [n.cost for n in new if isinstance(n, GradNode)]
        self.shared = [n for n in nodes if isinstance(n, SharedVariable)]
        self.randoms = [n for n in nodes if isinstance(n, RandomNode)]
        self.grads = [n for n in nodes if isinstance(n, GradNode)]

    def __call__(self, *args):
        if len(args) != len(self.inputs):
            raise TypeError("%s takes %d inputs" % (self.name, len(self.inputs)))
        t0 = time.perf_counter()
        if BUDGET_S is not None:
            if _FIRST_CALL[0] is None:
                _FIRST_CALL[0] = t0
            elif t0 - _FIRST_CALL[0] > BUDGET_S:
                raise BudgetExceeded(self.name)
        env = Env()
        ins = []
        for s, a in zip(self.inputs, args):
            v = as_tensor(np.asarray(a)).to(DTYPE)
            env.bind(s, v)
            ins.append(v.numpy().copy())
        for s in self.shared:
            env.bind(s, s.value)
        draws = {}
        for r in self.randoms:
            d = r.draw()
            draws[r.name] = np.array(d)
            env.bind(r, as_tensor(d))
        with torch.no_grad():
            done = {}
            for g in self.grads:
                key = (id(g.cost), tuple(id(w) for w in g.wrt))
                if key not in done:
                    done[key] = gradients(g.cost, g.wrt, env)
                env.bind(g, done[key][g.index])
            outs = [evaluate(o, env) for o in self.outputs]
            new = [(s, evaluate(e, env)) for s, e in self.updates]
        outs = [o.detach().clone() for o in outs]
        for s, v in new:
            s.value = v.detach().clone().to(DTYPE)
        res = [o.numpy() if o.ndim else np.float64(o.item()) for o in outs]
        CALL_LOG.append(dict(fn=self.name, inputs=ins, random=draws, outputs=[np.array(r) for r in res], t0=t0, t1=time.perf_counter()))
        return res[0] if self.single else res


def function(inputs, outputs=None, updates=None, name=None, **kw):
    return Function(inputs, outputs, updates, name)


from . import tensor  # noqa: E402
