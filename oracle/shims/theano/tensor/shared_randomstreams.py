"""theano.tensor.shared_randomstreams stand-in (FNN_wnzh.py:11,15,144,154,166).  TEST INFRASTRUCTURE."""
import numpy as np

import theano as _th


class RandomStreams(object):
    def __init__(self, seed=None):
        self.gen_seedgen = np.random.RandomState(seed)
        self.n = 0

    def binomial(self, size=None, n=1, p=0.5, ndim=None, dtype="int64"):
        rstate = np.random.RandomState(int(self.gen_seedgen.randint(2 ** 30)))
        node = _th.RandomNode(rstate, lambda: rstate.binomial(n, p, size=size), "binomial%d" % self.n)
        self.n += 1
        return node
