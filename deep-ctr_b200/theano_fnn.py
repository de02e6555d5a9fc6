"""Theano-generation FNN (python/FNN_wnzh.py) on the B200 hot path.

The reference script compiles `train(x, y) -> [gx, w1, w2, w3, b1, b2, b3]` and `predict(x) -> p`
(FNN_wnzh.py:177-183), gathers x on the host from a Python dict (:87-96) and applies the touched-row SGD
in a Python triple loop (:300-306).  Here the gather, the MLP, the loss and the sequential decayed scatter
all run on the GPU, so `train` takes FEATURE IDS instead of the dense x and returns (loss, p_drop).
"""
import numpy as np
import torch

from . import _lib as L
from .engine import StepEngine
from .models import _Base

NAME_FIELD = {'weekday': 0, 'hour': 1, 'useragent': 2, 'IP': 3, 'region': 4, 'city': 5, 'adexchange': 6, 'domain': 7,
              'slotid': 8, 'slotwidth': 9, 'slotheight': 10, 'slotvisibility': 11, 'slotformat': 12, 'creative': 13,
              'advertiser': 14, 'slotprice': 15}   # FNN_wnzh.py:51-53


def init_mlp(xdim, hidden1, hidden2, acti_type="tanh", seed=1234):
    """FNN_wnzh.py:16-17,106-130: numpy legacy global RNG seeded 1234; U(+-sqrt(6/(fan_in+fan_out))), x4 for tanh;
    w3 = 0, biases 0."""
    rng = np.random.RandomState(seed)
    lim1 = np.sqrt(6.0 / (xdim + hidden1))
    w = rng.uniform(low=-lim1, high=lim1, size=(xdim, hidden1))
    ww1 = w if acti_type == "sigmoid" else (w * 4 if acti_type == "tanh" else rng.uniform(-1, 1, size=(xdim, hidden1)))
    lim2 = np.sqrt(6.0 / (hidden1 + hidden2))
    v = rng.uniform(low=-lim2, high=lim2, size=(hidden1, hidden2))
    ww2 = v if acti_type == "sigmoid" else (v * 4 if acti_type == "tanh" else rng.uniform(-1, 1, size=(hidden1, hidden2)))
    return [ww1, ww2, np.zeros((hidden2, 1))], [np.zeros(hidden1), np.zeros(hidden2), np.zeros(1)]


class MaskStreams(object):
    """Dropout masks r1, r2 = srng.binomial(size=(1,H), n=1, p=dropout) (FNN_wnzh.py:154,166) from
    RandomStreams(seed=234) (:15).  Theano seeds each op's numpy RandomState from
    RandomState(234).randint(2**30) in creation order r0, r1, r2 (recollection, cannot be checked
    without Theano) -- parity tests inject masks instead of relying on this."""

    def __init__(self, hidden1, hidden2, p, seed=234):
        g = np.random.RandomState(seed)
        seeds = [int(g.randint(2 ** 30)) for _ in range(3)]
        self.r1 = np.random.RandomState(seeds[1])
        self.r2 = np.random.RandomState(seeds[2])
        self.h1, self.h2, self.p = hidden1, hidden2, p

    def draw(self):
        return (self.r1.binomial(n=1, p=self.p, size=(1, self.h1)).astype(np.float32),
                self.r2.binomial(n=1, p=self.p, size=(1, self.h2)).astype(np.float32))

    # The two binomial draws cost ~20 us of host time, as much as an eighth of a whole step at batch 4096: the model draws
    # the NEXT step's pair right after it has launched the current step, while the GPU is busy (prefetch), and take() hands
    # it out.  The sequence of pairs is exactly the sequence draw() would have produced.
    _ahead = None

    def prefetch(self):
        if self._ahead is None:
            self._ahead = self.draw()

    def take(self):
        m, self._ahead = self._ahead, None
        return m if m is not None else self.draw()


class FNN_wnzh(_Base):
    def __init__(self, fm_model, batch_size=100, lr=0.001, dropout=0.5, hidden1=300, hidden2=100, lambda1=0.0,
                 lambda_fm=0.1, acti_type="tanh", n_fields=16, device="cuda", weights=None, max_feats=None):
        """fm_model: dict(w_0=float, table=[X, k+1] float array indexed by feature id, feat_field=[X] int)
        as loaded by data.load_fm_model (FNN_wnzh.py:62-84).  Hyper-parameters default to the advertiser-2997
        branch (:43-49), the only one that defines them all.  max_feats: longest feature list of a sample (default
        n_fields; larger when lines carry several features of one field)."""
        table = np.asarray(fm_model["table"], dtype=np.float32)
        X, d = table.shape
        self.k = d
        self.xdim = 1 + n_fields * d                                  # :78
        self.feat_field = np.asarray(fm_model["feat_field"], dtype=np.int64)
        self.lr, self.dropout, self.lambda1, self.lambda_fm = lr, dropout, lambda1, lambda_fm
        act = {"tanh": "tanh", "sigmoid": "sigmoid", "linear": "none"}[acti_type]
        # one extra all-zero row X: the slot of a field that a sample does not have (x stays 0 there, :91-95)
        eng = StepEngine(X + 1, n_fields, d - 1, [hidden1, hidden2], batch_size, cross="none", bias_first=True,
                         act0="none", act=act, scatter_mode="seq_decay", device=device)
        eng.set_table(np.concatenate([table, np.zeros((1, d), np.float32)], axis=0))
        eng.bias().fill_(float(fm_model["w_0"]))
        Ws, bs = weights if weights is not None else init_mlp(self.xdim, hidden1, hidden2, acti_type)
        for l in range(3):
            eng.set_layer(l, Ws[l], bs[l])
        eng.wts_are_ones = False
        eng.skip_row = X            # weights are 1 except on the pad row: the fold drops that row instead of gathering weights
        self.pad_row = X
        # the dropout branch computes hidden layer 2 with tanh whatever acti_type says (d2 = T.tanh(...), FNN_wnzh.py:163); only
        # the prediction branch uses acti_type there (:157-162)
        self._acts_train, self._acts_eval = [None, act, "tanh"], [None, act, act]
        eng.acts = self._acts_train
        N = batch_size
        self.J = int(max_feats or n_fields)
        eng.use_update_list(self.J)         # every feature of a sample is updated, overwritten ones too (FNN_wnzh.py:300-306)
        extra = [("feats", np.int32, (N, self.J))]
        self.mask_streams = MaskStreams(hidden1, hidden2, dropout) if dropout < 1.0 else None
        if self.mask_streams is not None:
            extra += [("mask1", np.float32, (1, eng.ldi[1])), ("mask2", np.float32, (1, eng.ldi[2]))]
        # feature ids, labels and the two broadcast dropout masks travel in ONE host->device copy; the slot-per-field
        # arrangement (FNN_wnzh.py:91-95) happens on the device (dctr_arrange_fields)
        self._setup_host(eng, batch_size, n_fields, extra_blocks=extra, with_ids=False)
        self.feats = self._stage.dev["feats"]
        self.feat_field_dev = torch.as_tensor(self.feat_field.astype(np.int32)).to(eng.dev)
        if self.mask_streams is not None:
            eng.masks = [None, self._stage.dev["mask1"], self._stage.dev["mask2"]]
        self._next_masks = None

    # ---- host side of feats_to_layer_one_array: slot of a feature = its field (FNN_wnzh.py:87-96) ----
    def arrange(self, feats):
        feats = np.asarray(feats, dtype=np.int64)
        n = feats.shape[0]
        ids = np.full((n, self.F), self.pad_row, dtype=np.int64)
        wts = np.zeros((n, self.F), dtype=np.float32)
        valid = feats >= 0
        rows = np.broadcast_to(np.arange(n)[:, None], feats.shape)[valid]
        slot = self.feat_field[feats[valid]]
        ids[rows, slot] = feats[valid]            # a later feature of the same field overwrites (:94-95)
        wts[rows, slot] = 1.0
        return ids, wts

    def set_masks(self, r1, r2):
        """Inject the dropout masks of the NEXT train() call ([1,H] or [H] arrays of 0/1)."""
        self._next_masks = (np.asarray(r1, np.float32).reshape(1, -1), np.asarray(r2, np.float32).reshape(1, -1))

    def _pre_step(self):
        """draw (or take the injected) masks of this step into the pinned staging block."""
        if self.mask_streams is None:
            return
        r1, r2 = self._next_masks if self._next_masks is not None else self.mask_streams.take()
        self._next_masks = None
        h = self._stage.host
        h["mask1"][:, :r1.shape[1]] = r1
        h["mask2"][:, :r2.shape[1]] = r2

    def _post_launch(self):
        """the step is running on the device: draw the next step's masks meanwhile"""
        if self.mask_streams is not None:
            self.mask_streams.prefetch()

    def _upload_masks(self):
        """masks only (bench.py's device-resident loop keeps feats / labels on the device)."""
        if self.mask_streams is not None:
            st = self._stage
            for k in ("mask1", "mask2"):
                st.dev[k].copy_(torch.from_numpy(st.host[k]), non_blocking=True)

    def _arrange_on_device(self):
        e = self.engine
        torch.ops.deepctr.arrange_update_list(self.feats, self.feat_field_dev, self.F, self.pad_row, self.pad_row, e.ids, e.wts,
                                              e.upd_rows.view(self.N, self.J), e.upd_pos.view(self.N, self.J))

    def _device_step(self):
        decay = 1.0 - 2.0 * self.lambda_fm * self.lr / getattr(self, "global_N", self.N)   # FNN_wnzh.py:305 (b_size)
        self._arrange_on_device()
        self.engine.train_step(self.lr, decay, 1.0, train_bias=False, last_decay=1.0 - 2.0 * self.lr * self.lambda1)

    def _stage_feats(self, feats, rows=None):
        h = self._stage.host["feats"]
        feats = np.asarray(feats)
        k, j = feats.shape
        if j > self.J:
            raise ValueError("a sample has %d features, the model was built with max_feats=%d" % (j, self.J))
        h[:k, :j] = feats
        if j < self.J:
            h[:k, j:] = -1
        if k < self.N:
            h[k:] = -1

    def train(self, feats, y):
        """One step of the loop body FNN_wnzh.py:296-306.  feats: [N, <=16] feature ids (-1 = absent)."""
        self._no_step_in_flight()
        self._stage_feats(feats)
        np.copyto(self._stage.host["y"], np.asarray(y).reshape(self.N), casting="unsafe")
        self._pre_step()
        return self._finish_step()

    def _stage_batch(self, feats, y):
        self._stage_feats(feats)
        np.copyto(self._stage.host["y"], np.asarray(y).reshape(self.N), casting="unsafe")

    def train_async(self, feats, y):
        """train(), two-deep pipelined (models._Base._train_async): returns (loss, p_drop) of the PREVIOUS call (None on the first);
        flush_async() returns the last one.  The mask stream is consumed in call order, as train() does."""
        return self._train_async(lambda: self._stage_batch(feats, y))

    def train_step_device(self, feats, y):
        """train() from DEVICE-resident blocks (ingest.DeviceText.parse_sparse with J = 16, pad_id = -1): int32 feature ids
        [N, 16], float32 labels [N]; only the two dropout masks go up and pred | loss come back."""
        self.feats.copy_(feats)
        self.engine.y.copy_(y)
        self._pre_step()
        self._upload_masks()
        self._run_step()
        self._post_launch()
        return self._fetch_out()

    def _swap_masks(self, masks):
        old = self.engine.masks
        self.engine.masks = masks if masks is not None else [None] * self.engine.nl
        return old

    def predict(self, feats):
        """predict(x) -> p_1: the no-dropout branch, no rescaling (FNN_wnzh.py:157-170,183)."""
        feats = np.asarray(feats)
        E = feats.shape[0]
        out = np.empty(E, dtype=np.float32)
        self._predict_into(feats, out, None)
        return out

    def _predict_into(self, feats, out_np, out_dev):
        E = feats.shape[0]
        self._no_step_in_flight()
        saved = self._swap_masks(None)
        self.engine.acts = self._acts_eval
        self._bind_status()
        try:
            for s in range(0, E, self.N):
                k = min(self.N, E - s)
                self._stage_feats(feats[s:s + k])
                self._stage.upload()
                self._arrange_on_device()
                self.engine.forward(False)
                if out_dev is not None:
                    out_dev[s:s + k].copy_(self.engine.pred[:k])
                self._out_pinned.copy_(self.engine.out, non_blocking=True)
                torch.cuda.current_stream().synchronize()
                self._check_ids()
                if out_np is not None:
                    out_np[s:s + k] = self._out_np[:k]
        finally:
            self.engine.acts = self._acts_train
            self._swap_masks(saved)

    def evaluate(self, feats, labels, nds_rate=None):
        """Predict E samples (feature lists [E, <= max_feats], -1 = absent) and score them ON THE DEVICE: dict(auc, log_loss,
        rmse, n_pos) -- print_err / get_err_bat of the script (FNN_wnzh.py:187-221) without the D2H of the predictions."""
        from .metrics import device_metrics
        feats = np.asarray(feats)
        preds = torch.empty(feats.shape[0], dtype=torch.float32, device=self.engine.dev)
        self._predict_into(feats, None, preds)
        return device_metrics(preds, np.asarray(labels, dtype=np.float32).reshape(-1), device=self.engine.dev, nds_rate=nds_rate)

    def set_weights(self, w):
        """(w1, w2, w3, b1, b2, b3) as get_weights() returns them -- w1.set_value(best_w1) ... of get_pred (FNN_wnzh.py:257-263)."""
        w1, w2, w3, b1, b2, b3 = w
        e = self.engine
        e.set_layer(0, w1, b1)
        e.set_layer(1, w2, b2)
        e.set_layer(2, np.asarray(w3).reshape(-1, 1), [b3])

    def fit(self, train_feats, train_y, test_feats, test_y, epochs=100, masks=None, log=None):
        """The epoch loop of the script (FNN_wnzh.py:286-343): per epoch n_batch = train_size // batch_size consecutive batches
        (the ragged tail is never trained on), then AUC / RMSE of the whole train and test files (print_err / get_err_bat,
        :187-221 -- scored on the device), then the patience rule: a new best test AUC snapshots the weights and earns credit
        (times_reduce up to 3), anything else costs one, below zero stops.  The snapshot is what the LAST train() call of the
        epoch returned, i.e. the weights BEFORE that step (:179-182, 330-335) -- kept.
        masks: optional iterator of (r1, r2) pairs injected per step (parity runs).  Returns dict(history, min_err,
        min_err_epoch, best) with best = the snapshot get_pred would load."""
        train_feats, test_feats = np.asarray(train_feats), np.asarray(test_feats)
        train_y, test_y = np.asarray(train_y), np.asarray(test_y)
        N = self.N
        n_batch = len(train_y) // N
        best, min_err, min_err_epoch, times_reduce, history = self.get_weights(), 0, 0, 0, []
        last = best
        for i in range(epochs):
            for j in range(n_batch):
                if j == n_batch - 1:
                    last = self.get_weights()
                if masks is not None:
                    self.set_masks(*next(masks))
                self.train(train_feats[j * N:(j + 1) * N], train_y[j * N:(j + 1) * N])
            tr = self.evaluate(train_feats, train_y)
            te = self.evaluate(test_feats, test_y)
            history.append(dict(epoch=i, train_auc=tr["auc"], train_rmse=tr["rmse"], test_auc=te["auc"], test_rmse=te["rmse"]))
            if log is not None:
                log("Test Err:%d\t%s\t%s" % (i, te["auc"], te["rmse"]))
            if te["auc"] > min_err:
                best, min_err, min_err_epoch = last, te["auc"], i
                if times_reduce < 3:
                    times_reduce += 1
            else:
                times_reduce -= 1
            if times_reduce < 0:
                break
        return dict(history=history, min_err=min_err, min_err_epoch=min_err_epoch, best=best)

    def get_weights(self):
        """(w1, w2, w3, b1, b2, b3) like the tail of train()'s outputs -- POST-update here."""
        e = self.engine
        (w1, b1), (w2, b2), (w3, b3) = e.get_layer(0), e.get_layer(1), e.get_layer(2)
        return w1, w2, w3.reshape(-1), b1, b2, float(b3[0])

    def get_table(self):
        return self.engine.get_table()[:self.pad_row]
