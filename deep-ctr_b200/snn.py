"""Fine-tuning step of the reference's SNN models (python/SNN_DAE.py, python/SNN_RBM.py) on the B200 hot path -- SURVEY.md
8(f) f4.  The pre-training that produces (ww0, bb0, ww1, bb1, ww2, bb2) (denoising auto-encoder / RBM) is out of scope; the
weights arrive as that pickle tuple (SNN_DAE.py:83-88) or are drawn like dl_utils.init_weight.

The reference builds x_dense = sigmoid(sum_feat ww0[feat] + bb0) per sample on the host (get_fi_h1_y, :234-260), runs the
Theano `train(x, y)` and applies `ww0[feat] -= lr*gx*x*(1-x)` in a Python double loop (:282-288).  Here the bag, the MLP and
the touched-row update run on the GPU (engine.BagEngine), so `train` takes the FEATURE IDS and returns (loss, p_drop).
"""
import numpy as np
import torch

from .engine import BagEngine
from .models import _Base
from .theano_fnn import MaskStreams


def init_weight(rng, fan_in, fan_out, acti_type):
    """dl_utils.init_weight (dl_utils.py:44-56): U(+-sqrt(6/(fan_in+fan_out))), x4 for sigmoid; zero bias."""
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    v = rng.uniform(low=-lim, high=lim, size=(fan_in, fan_out))
    w = v * 4 if acti_type == "sigmoid" else (v if acti_type == "tanh" else rng.uniform(-1, 1, size=(fan_in, fan_out)))
    return w, np.zeros(fan_out)


class SNN(_Base):
    def __init__(self, x_dim=133465, max_feats=16, batch_size=1000, lr=0.0005, dropout=0.99, hidden0=200, hidden1=300,
                 hidden2=100, lambda1=0.0, acti_type="tanh", weights=None, device="cuda"):
        """Hyper-parameters default to the advertiser-2997 branch (SNN_DAE.py:19-49).  weights: the pickle tuple
        (ww0 [x_dim, hidden0], bb0, ww1, bb1, ww2, bb2) of :83-88; w3 = 0, b3 = 0 (:91-94).  max_feats: slots per sample
        (samples with fewer active features are padded with a zero-weight slot)."""
        if weights is None:
            rng = np.random.RandomState(1234)                                       # :18-19
            ww0, bb0 = init_weight(rng, x_dim, hidden0, "sigmoid")                   # :65-67
            ww1, bb1 = init_weight(rng, hidden0, hidden1, "sigmoid")
            ww2, bb2 = init_weight(rng, hidden1, hidden2, "sigmoid")
        else:
            ww0, bb0, ww1, bb1, ww2, bb2 = weights
        self.lr, self.dropout, self.lambda1 = lr, dropout, lambda1
        act = {"tanh": "tanh", "sigmoid": "sigmoid", "linear": "none"}[acti_type]
        X = int(x_dim)
        eng = BagEngine(X + 1, max_feats, hidden0, [hidden1, hidden2], batch_size, act=act, device=device)   # row X: all-zero pad row
        eng.set_table(np.concatenate([np.asarray(ww0, np.float32), np.zeros((1, hidden0), np.float32)], axis=0))
        eng.bias().copy_(torch.as_tensor(np.asarray(bb0), dtype=torch.float32))
        eng.set_layer(0, ww1, bb1)
        eng.set_layer(1, ww2, bb2)
        eng.set_layer(2, np.zeros((hidden2, 1)), np.zeros(1))
        eng.wts_are_ones = False
        # the dropout branch computes hidden layer 2 with tanh whatever acti_type says (SNN_DAE.py:136-137); the prediction
        # branch uses acti_type (:127-133)
        self._acts_train, self._acts_eval = [None, act, "tanh"], [None, act, act]
        eng.acts = self._acts_train
        eng.skip_row = X            # 0 / 1 weights, 0 only on the pad row: no weight gather in the fold
        self.pad_row = X
        self.mask_streams = MaskStreams(hidden1, hidden2, dropout) if dropout < 1.0 else None
        extra = []
        if self.mask_streams is not None:
            extra = [("mask1", np.float32, (1, eng.ldi[1])), ("mask2", np.float32, (1, eng.ldi[2]))]
        self._setup_host(eng, batch_size, max_feats, extra_blocks=extra)
        if self.mask_streams is not None:
            eng.masks = [None, self._stage.dev["mask1"], self._stage.dev["mask2"]]
        self._next_masks = None

    def slots(self, feats):
        """[n, <= max_feats] feature ids (-1 = absent; a ragged list of lists is padded) -> (ids, wts) slot blocks."""
        if not isinstance(feats, np.ndarray):
            rows = [list(r) for r in feats]
            arr = np.full((len(rows), self.F), -1, dtype=np.int64)
            for i, r in enumerate(rows):
                if len(r) > self.F:
                    raise ValueError("sample %d has %d active features, the model was built for %d" % (i, len(r), self.F))
                arr[i, :len(r)] = r
            feats = arr
        if feats.shape[1] > self.F:
            raise ValueError("%d feature slots, the model was built for %d" % (feats.shape[1], self.F))
        n = feats.shape[0]
        ids = np.full((n, self.F), self.pad_row, dtype=np.int64)
        wts = np.zeros((n, self.F), dtype=np.float32)
        on = feats >= 0
        ids[:, :feats.shape[1]][on] = feats[on]
        wts[:, :feats.shape[1]][on] = 1.0
        return ids, wts

    def set_masks(self, r1, r2):
        """Inject the dropout masks of the NEXT train() call ([1,H] or [H] arrays of 0/1)."""
        self._next_masks = (np.asarray(r1, np.float32).reshape(1, -1), np.asarray(r2, np.float32).reshape(1, -1))

    def _pre_step(self):
        if self.mask_streams is None:
            return
        r1, r2 = self._next_masks if self._next_masks is not None else self.mask_streams.take()
        self._next_masks = None
        h = self._stage.host
        h["mask1"][:, :r1.shape[1]] = r1
        h["mask2"][:, :r2.shape[1]] = r2

    def _post_launch(self):
        if self.mask_streams is not None:
            self.mask_streams.prefetch()

    def _device_step(self):
        # cost adds lambda1 * (|w1|^2 + |w2|^2 + |w3|^2 + |b|^2) (SNN_DAE.py:151-153): decay on the MLP, not on ww0 / bb0
        self.engine.train_step(self.lr, 1.0, 1.0, train_bias=True, mlp_decay=1.0 - 2.0 * self.lr * self.lambda1)

    def train(self, feats, y):
        """One iteration of the loop body SNN_DAE.py:276-288: (sum of the cross-entropy terms, p_drop [N])."""
        ids, wts = self.slots(feats)
        return self._stage_and_step(ids, wts, y)

    def _swap_masks(self, masks):
        """masks = None: the prediction branch (no dropout, acti_type on both hidden layers); otherwise back to training"""
        old = self.engine.masks
        self.engine.masks = masks if masks is not None else [None] * self.engine.nl
        self.engine.acts = self._acts_eval if masks is None else self._acts_train
        return old

    def predict(self, feats):
        """predict(x) -> p_1, the branch without dropout (SNN_DAE.py:127-141,164)."""
        ids, wts = self.slots(feats)
        return self._predict_chunks(ids, wts)

    def get_weights(self):
        """(ww0, bb0, w1, b1, w2, b2, w3, b3) after the update."""
        e = self.engine
        (w1, b1), (w2, b2), (w3, b3) = e.get_layer(0), e.get_layer(1), e.get_layer(2)
        return e.get_table()[:self.pad_row], e.bias().cpu().numpy(), w1, b1, w2, b2, w3.reshape(-1), float(b3[0])
