"""CPU oracle for the deep-ctr hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A numpy (float64 by default) restatement of the reference algorithm for the path of BASELINE.json's
north_star.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this file; the product package (deep-ctr_b200/) never does.

PARITY PINNED against the reference itself (round 2): the reference's own Python -- /root/reference/python/*.py after the
mechanical Py2->3 fixes of oracle/build_ref.py -- is executed on two small engine shims (oracle/shims/theano, oracle/shims/
tensorflow: lazy graphs evaluated by torch, autograd for T.grad / minimize), and tests/test_oracle_vs_ref.py asserts this
restatement equal to it: every train() / predict() call of the whole FNN_wnzh.py script (23 epochs, duplicate / missing /
repeated features included, 1e-8), 3 optimiser steps + eval + dump of FM, LR, FNN, FNN_IP_L3/L5/L7, FNN_OP_L3 under sgd / adam /
ftrl with logged dropout masks (1e-9), build_inds, get_fxy, get_batch_xy, early_stop, watch_train, the whole baseline.train()
loop, collect / stat / load_ipinyou_data, and every train() / predict() call of the whole SNN_DAE.py script against OracleSNN
(1e-9).  What the shims stand in for is the ENGINES (Theano, TensorFlow 0.x: absent from
/root/reference, unpinned versions): their op semantics are restated from the published behaviour (oracle/shims/*/__init__.py
list every point).  Golden vectors generated from those runs: tests/golden/ref_golden.pkl (tests/golden/make_golden_ref.py).
Independently of the reference, this file is also pinned by central finite differences of every hand-derived backward,
algebraic identities (FM pairwise identity, OP prefix form, IP index formula; FTRL = AdaGrad) and torch-autograd re-derivations
(tests/test_oracle.py).

Reference citations are file:line under /root/reference/python.
"""
import numpy as np

ACTS = ("none", "tanh", "relu", "sigmoid")


def activate(act, x):
    """tf_util.py:32-38 (tanh / relu / else sigmoid); 'none' is the Theano input layer (FNN_wnzh.py:147)."""
    if act == "none":
        return x
    if act == "tanh":
        return np.tanh(x)
    if act == "relu":
        return np.maximum(x, 0)
    return 1.0 / (1.0 + np.exp(-x))


def activate_grad_from_out(act, t):
    if act == "none":
        return np.ones_like(t)
    if act == "tanh":
        return 1 - t * t
    if act == "relu":
        return (t > 0).astype(t.dtype)
    return t * (1 - t)


def sigmoid(z):
    return 1.0 / (1.0 + np.exp(-z))


def xent_with_logits(z, y):
    """tf.nn.sigmoid_cross_entropy_with_logits (FM.py:34, FNN.py:54): max(z,0) - z*y + log1p(exp(-|z|)).
    Equal to the naive Theano form -y log p - (1-y) log(1-p) (FNN_wnzh.py:172) wherever that is finite."""
    return np.maximum(z, 0) - z * y + np.log1p(np.exp(-np.abs(z)))


def pair_index(i, j, F):
    """Position of pair (i<j) in the i-outer / j-inner enumeration of FNN_IP_L3.py:91-94."""
    return i * F - i * (i + 1) // 2 + (j - i - 1)


# --------------------------------------------------------------------------------------------------
# embedding layer: gather + concat (+ IP / OP)
# --------------------------------------------------------------------------------------------------
def gather_fields(table, ids, wts):
    """e[n,f,:] = wts[n,f] * table[ids[n,f], :].
    Theano: FNN_wnzh.py:91-96 (weights ignored = 1).  TF: numeric field i is value * row i (FNN.py:77),
    categorical field j is embedding_lookup_sparse with weight c_wts (FNN.py:78-82); with global ids
    `range(13) + offsets+local` (baseline.py:190-198) both are this one expression."""
    e = table[ids]
    if wts is not None:
        e = e * wts[..., None]
    return e


def cross_ip(e):
    """p[n, pi(i,j)] = <e_i, e_j>, i<j (FNN_IP_L3.py:91-94)."""
    N, F, d = e.shape
    iu, ju = np.triu_indices(F, 1)   # row-major upper triangle == i-outer / j-inner
    return np.einsum("npd,npd->np", e[:, iu, :], e[:, ju, :])


def cross_op(e):
    """q[n, a*d+b] = sum_{i<j} e_i[a] e_j[b] (FNN_OP_L3.py:93-97: outer(mbd_i, mbd_j) flattened row-major,
    summed over pairs)."""
    N, F, d = e.shape
    q = np.zeros((N, d, d), dtype=e.dtype)
    for i in range(F - 1):
        for j in range(i + 1, F):
            q += e[:, i, :, None] * e[:, j, None, :]
    return q.reshape(N, d * d)


def build_z(e, cross, bias, bias_first):
    """Theano: x = [w_0 | e...] (FNN_wnzh.py:92-95).  TF: z1 = [mbd | p_mbd | b_mbd] (FNN.py:83-85,
    FNN_IP_L3.py:95-97).  The OP width is the forward's true d*d, not the declared P."""
    N, F, d = e.shape
    parts = [e.reshape(N, F * d)]
    if cross == "ip":
        parts.append(cross_ip(e))
    elif cross == "op":
        parts.append(cross_op(e))
    b = np.full((N, 1), bias, dtype=e.dtype)
    parts = [b] + parts if bias_first else parts + [b]
    return np.concatenate(parts, axis=1)


def cross_backward(e, cross, dz_cross):
    """d loss / d e from the cross block (hand-derived; finite-difference checked in tests/test_oracle.py)."""
    N, F, d = e.shape
    de = np.zeros_like(e)
    if cross == "ip":
        iu, ju = np.triu_indices(F, 1)
        for t in range(len(iu)):
            i, j = iu[t], ju[t]
            de[:, i, :] += dz_cross[:, t, None] * e[:, j, :]
            de[:, j, :] += dz_cross[:, t, None] * e[:, i, :]
    elif cross == "op":
        dq = dz_cross.reshape(N, d, d)
        for i in range(F - 1):
            for j in range(i + 1, F):
                de[:, i, :] += np.einsum("nab,nb->na", dq, e[:, j, :])
                de[:, j, :] += np.einsum("nab,na->nb", dq, e[:, i, :])
    return de


# --------------------------------------------------------------------------------------------------
# the network: embedding layer + MLP
# --------------------------------------------------------------------------------------------------
class OracleNet:
    """in_1 = act0(z); pre_l = in_l W_l + b_l; in_{l+1} = mask_l * act(pre_l); logit = pre_L.
    TF (FNN.py:86-94): act0 = act, masks are tf.nn.dropout masks (0 or 1/keep) applied to EVERY layer input
    including act(z1); Theano (FNN_wnzh.py:147-170): act0 = none, masks r1, r2 in {0,1} broadcast over the
    batch, no rescale."""

    def __init__(self, table, bias, Ws, bs, F, d, cross="none", bias_first=False, act0="none", act="tanh",
                 dtype=np.float64):
        self.dtype = dtype
        self.table = np.array(table, dtype=dtype)      # [X, d]
        self.bias = dtype(bias)
        self.Ws = [np.array(w, dtype=dtype) for w in Ws]
        self.bs = [np.array(b, dtype=dtype).reshape(-1) for b in bs]
        self.F, self.d, self.cross, self.bias_first, self.act0, self.act = F, d, cross, bias_first, act0, act

    # masks: list of length L (mask on in_1 .. in_L) or None; each None or array broadcastable to the layer input
    def forward(self, ids, wts=None, masks=None):
        dt = self.dtype
        wts = None if wts is None else np.asarray(wts, dtype=dt)
        e = gather_fields(self.table, ids, wts)
        z = build_z(e, self.cross, self.bias, self.bias_first)
        L = len(self.Ws)
        masks = masks or [None] * L
        ins = []
        a = activate(self.act0, z)
        if masks[0] is not None:
            a = a * masks[0]
        ins.append(a)
        for l in range(L - 1):
            pre = ins[l] @ self.Ws[l] + self.bs[l]
            a = activate(self.act, pre)
            if masks[l + 1] is not None:
                a = a * masks[l + 1]
            ins.append(a)
        logit = (ins[L - 1] @ self.Ws[L - 1] + self.bs[L - 1]).reshape(-1)
        return logit, dict(e=e, z=z, ins=ins, masks=masks, ids=ids, wts=wts)

    def predict(self, ids, wts=None):
        return sigmoid(self.forward(ids, wts)[0])

    def backward(self, cache, dlogit):
        """Returns dWs, dbs, dbias, ge[N,F,d] (gradient w.r.t. the gathered table rows, weight applied) and
        gz (gradient w.r.t. z -- Theano's gx, FNN_wnzh.py:174)."""
        L = len(self.Ws)
        ins, masks = cache["ins"], cache["masks"]
        dWs, dbs = [None] * L, [None] * L
        dpre = dlogit.reshape(-1, 1)
        for l in range(L - 1, -1, -1):
            dWs[l] = ins[l].T @ dpre
            dbs[l] = dpre.sum(axis=0)
            din = dpre @ self.Ws[l].T
            act = self.act if l > 0 else self.act0
            m = masks[l]
            if m is not None:
                safe = np.where(m != 0, m, 1.0)
                t = np.where(m != 0, ins[l] / safe, 0.0)
                dpre = din * m * activate_grad_from_out(act, t)
            else:
                dpre = din * activate_grad_from_out(act, ins[l])
        gz = dpre                                   # d loss / d z  [N, D0]
        N = gz.shape[0]
        F, d = self.F, self.d
        off = 1 if self.bias_first else 0
        de = gz[:, off:off + F * d].reshape(N, F, d).copy()
        if self.cross != "none":
            cw = F * (F - 1) // 2 if self.cross == "ip" else d * d
            de += cross_backward(cache["e"], self.cross, gz[:, F * d:F * d + cw])
        dbias = gz[:, 0].sum() if self.bias_first else gz[:, -1].sum()
        ge = de if cache["wts"] is None else de * cache["wts"][..., None]
        return dWs, dbs, dbias, ge, gz

    # ---- TF semantics: loss = reduce(xent), plain SGD on every variable (tf_util.py:26-28) ------------
    def train_step_tf(self, ids, wts, y, lr, masks=None, reduce="sum"):
        y = np.asarray(y, dtype=self.dtype)
        logit, cache = self.forward(ids, wts, masks)
        N = len(y)
        scale = 1.0 if reduce == "sum" else 1.0 / N
        loss = xent_with_logits(logit, y).sum() * scale
        dlogit = (sigmoid(logit) - y) * scale
        dWs, dbs, dbias, ge, _ = self.backward(cache, dlogit)
        # embedding rows: the gradient of tf.slice/concat + embedding_lookup_sparse is the SUM over duplicates
        # (FNN.py:46-49); only touched rows have a non-zero gradient, so SGD moves only them.
        np.add.at(self.table, ids.reshape(-1), -lr * ge.reshape(-1, self.d))
        self.bias = self.bias - lr * dbias          # fm_b is a trained variable (FNN.py:30)
        for l in range(len(self.Ws)):
            self.Ws[l] -= lr * dWs[l]
            self.bs[l] -= lr * dbs[l]
        return loss, sigmoid(logit)

    def train_step_tf_adam(self, ids, wts, y, lr, eps=1e-8, masks=None, reduce="sum", beta1=0.9, beta2=0.999):
        """tf.train.AdamOptimizer(lr, epsilon=eps).minimize(loss) over ALL variables (tf_util.py:18-21; baseline.py:93 uses
        adam for every Criteo model).  TF's form:  lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t);  m = b1 m + (1-b1) g;
        v = b2 v + (1-b2) g^2;  theta -= lr_t * m / (sqrt(v) + eps).  The table gradient is densified (FNN.py:46-49), so
        the slots of EVERY row decay each step and an untouched row keeps moving while its m is non-zero."""
        y = np.asarray(y, dtype=self.dtype)
        logit, cache = self.forward(ids, wts, masks)
        N = len(y)
        scale = 1.0 if reduce == "sum" else 1.0 / N
        loss = xent_with_logits(logit, y).sum() * scale
        dlogit = (sigmoid(logit) - y) * scale
        dWs, dbs, dbias, ge, _ = self.backward(cache, dlogit)
        if not hasattr(self, "_adam"):
            z = lambda a: [np.zeros_like(np.asarray(a, dtype=self.dtype)), np.zeros_like(np.asarray(a, dtype=self.dtype))]
            self._adam = dict(t=0, table=z(self.table), bias=z(np.asarray(self.bias)), Ws=[z(w) for w in self.Ws], bs=[z(b) for b in self.bs])
        st = self._adam
        st["t"] += 1
        lr_t = lr * np.sqrt(1.0 - beta2 ** st["t"]) / (1.0 - beta1 ** st["t"])

        def upd(theta, g, mv):
            mv[0] = beta1 * mv[0] + (1 - beta1) * g
            mv[1] = beta2 * mv[1] + (1 - beta2) * g * g
            return theta - lr_t * mv[0] / (np.sqrt(mv[1]) + eps)
        G = np.zeros_like(self.table)
        np.add.at(G, ids.reshape(-1), ge.reshape(-1, self.d))
        self.table[...] = upd(self.table, G, st["table"])
        self.bias = float(upd(np.asarray(self.bias, dtype=self.dtype), np.asarray(dbias, dtype=self.dtype), st["bias"]))
        for l in range(len(self.Ws)):
            self.Ws[l] = upd(self.Ws[l], dWs[l], st["Ws"][l])
            self.bs[l] = upd(self.bs[l], dbs[l], st["bs"][l])
        return loss, sigmoid(logit)

    def train_step_tf_ftrl(self, ids, wts, y, lr, masks=None, reduce="sum", init_accum=0.1, lr_power=-0.5, l1=0.0, l2=0.0):
        """tf.train.FtrlOptimizer(lr).minimize(loss) over ALL variables (tf_util.py:22-24), dense like Adam (the table
        gradient is densified, FNN.py:46-49): see ftrl_apply."""
        y = np.asarray(y, dtype=self.dtype)
        logit, cache = self.forward(ids, wts, masks)
        N = len(y)
        scale = 1.0 if reduce == "sum" else 1.0 / N
        loss = xent_with_logits(logit, y).sum() * scale
        dlogit = (sigmoid(logit) - y) * scale
        dWs, dbs, dbias, ge, _ = self.backward(cache, dlogit)
        if not hasattr(self, "_ftrl"):
            z = lambda a: [np.full_like(np.asarray(a, dtype=self.dtype), init_accum), np.zeros_like(np.asarray(a, dtype=self.dtype))]
            self._ftrl = dict(table=z(self.table), bias=z(np.asarray(self.bias)), Ws=[z(w) for w in self.Ws], bs=[z(b) for b in self.bs])
        st = self._ftrl
        kw = dict(lr_power=lr_power, l1=l1, l2=l2)
        G = np.zeros_like(self.table)
        np.add.at(G, ids.reshape(-1), ge.reshape(-1, self.d))
        self.table[...] = ftrl_apply(self.table, G, st["table"], lr, **kw)
        self.bias = float(ftrl_apply(np.asarray(self.bias, dtype=self.dtype), np.asarray(dbias, dtype=self.dtype), st["bias"], lr, **kw))
        for l in range(len(self.Ws)):
            self.Ws[l] = ftrl_apply(self.Ws[l], dWs[l], st["Ws"][l], lr, **kw)
            self.bs[l] = ftrl_apply(self.bs[l], dbs[l], st["bs"][l], lr, **kw)
        return loss, sigmoid(logit)

    # ---- Theano semantics (FNN_wnzh.py:172-182, 298-306) --------------------------------------------
    def train_step_theano(self, ids, y, lr, lambda_fm, masks=None, lambda1=0.0, faithful=False, closed_form=False):
        y = np.asarray(y, dtype=self.dtype)
        logit, cache = self.forward(ids, None, masks)
        p = sigmoid(logit)
        # cost = xent.sum() + lambda1*(|w3|^2 + b3^2)   (FNN_wnzh.py:172-173)
        loss = xent_with_logits(logit, y).sum() + lambda1 * ((self.Ws[-1] ** 2).sum() + (self.bs[-1] ** 2).sum())
        dWs, dbs, _, ge, gz = self.backward(cache, p - y)
        dWs[-1] = dWs[-1] + 2 * lambda1 * self.Ws[-1]
        dbs[-1] = dbs[-1] + 2 * lambda1 * self.bs[-1]
        N, F = ids.shape
        decay = 1 - 2.0 * lambda_fm * lr / N
        if faithful:
            # the literal triple loop of FNN_wnzh.py:300-306 (sample-major, duplicates applied one after another)
            off = 1
            for t in range(N):
                for f in range(F):
                    feat = ids[t, f]
                    for l in range(self.d):
                        self.table[feat, l] = self.table[feat, l] * decay - lr * gz[t, off + f * self.d + l]
        elif closed_form:
            seq_decay_scatter_closed_form(self.table, ids, ge, lr, decay)
        else:
            seq_decay_scatter(self.table, ids, ge, lr, decay)
        for l in range(len(self.Ws)):               # w_0 (the bias column) is never updated (FNN_wnzh.py:93,300-306)
            self.Ws[l] -= lr * dWs[l]
            self.bs[l] -= lr * dbs[l]
        return loss, p


    def train_step_theano_feats(self, feats, feat_field, y, lr, lambda_fm, masks=None, lambda1=0.0):
        """The loop body FNN_wnzh.py:296-306 on per-sample FEATURE LISTS (ragged, as get_fxy returns them, :248-253), literal:
        x is built by feats_to_layer_one_array (:87-96: slot = field of the feature, a later feature of the same field
        overwrites the slot, a field without a feature stays 0), then EVERY feature of the sample -- the overwritten one
        too -- is updated from the gradient of the slot of its field (:300-306).  self.table must end with one all-zero row
        (the slot of an absent field); it is never updated."""
        y = np.asarray(y, dtype=self.dtype)
        N, F, d = len(feats), self.F, self.d
        pad = self.table.shape[0] - 1
        ids = np.full((N, F), pad, dtype=np.int64)
        for t, ft in enumerate(feats):
            for feat in ft:
                ids[t, feat_field[feat]] = feat
        logit, cache = self.forward(ids, None, masks)
        p = sigmoid(logit)
        loss = xent_with_logits(logit, y).sum() + lambda1 * ((self.Ws[-1] ** 2).sum() + (self.bs[-1] ** 2).sum())
        dWs, dbs, _, _, gz = self.backward(cache, p - y)
        dWs[-1] = dWs[-1] + 2 * lambda1 * self.Ws[-1]
        dbs[-1] = dbs[-1] + 2 * lambda1 * self.bs[-1]
        b_size = N
        for t in range(b_size):
            for feat in feats[t]:
                o = 1 + feat_field[feat] * d
                for l in range(d):
                    self.table[feat, l] = self.table[feat, l] * (1 - 2. * lambda_fm * lr / b_size) - lr * gz[t, o + l]
        for l in range(len(self.Ws)):
            self.Ws[l] -= lr * dWs[l]
            self.bs[l] -= lr * dbs[l]
        return loss, p, dict(x=cache["z"], gx=gz, ids=ids)


def seq_decay_scatter(table, ids, ge, lr, decay):
    """Vectorised equivalent of FNN_wnzh.py:300-306: per row, occurrences in (sample, field) order,
    w <- w*decay - lr*g.  Rounds of 'k-th occurrence of each row' keep the per-row order exact."""
    flat = ids.reshape(-1)
    g = ge.reshape(len(flat), -1)
    order = np.argsort(flat, kind="stable")
    srow = flat[order]
    first = np.r_[True, srow[1:] != srow[:-1]]
    start = np.maximum.accumulate(np.where(first, np.arange(len(srow)), 0))
    occ = np.arange(len(srow)) - start            # occurrence number within its segment
    for k in range(int(occ.max()) + 1 if len(occ) else 0):
        sel = order[occ == k]
        rows = flat[sel]
        table[rows] = table[rows] * decay - lr * g[sel]


def seq_decay_scatter_closed_form(table, ids, ge, lr, decay):
    """Same recurrence solved per row: w_m = w_0*c^m - lr * sum_j c^(m-1-j) g_j (j = occurrence number).
    One sort + one segmented reduction: the fair vectorised CPU cost of FNN_wnzh.py:300-306 (bench baseline)."""
    flat = ids.reshape(-1)
    g = ge.reshape(len(flat), -1)
    order = np.argsort(flat, kind="stable")
    srow = flat[order]
    first = np.r_[True, srow[1:] != srow[:-1]]
    starts = np.flatnonzero(first)
    lens = np.diff(np.r_[starts, len(srow)])
    occ = np.arange(len(srow)) - np.repeat(starts, lens)
    m = np.repeat(lens, lens)
    wgt = decay ** (m - 1 - occ)
    acc = np.add.reduceat(g[order] * wgt[:, None], starts, axis=0)
    rows = srow[starts]
    table[rows] = table[rows] * (decay ** lens)[:, None] - lr * acc


def ftrl_apply(theta, g, slots, lr, lr_power=-0.5, l1=0.0, l2=0.0):
    """One dense FTRL-proximal update in TensorFlow's ApplyFtrl form -- what tf.train.FtrlOptimizer(lr) (tf_util.py:22-24)
    runs: learning_rate_power -0.5, initial_accumulator_value 0.1, l1 = l2 = 0.  TensorFlow is not under /root/reference
    (unpinned 0.x dependency); this is its published kernel:
        n' = n + g^2;  z += g - (n'^(-p) - n^(-p)) / lr * theta;  theta = |z| > l1 ? (l1 sign(z) - z) / (n'^(-p)/lr + 2 l2) : 0
    slots = [n, z], updated in place; returns the new theta.  Elements with g = 0 are re-derived from z alone."""
    n, z = slots
    n_new = n + g * g
    z += g - (n_new ** (-lr_power) - n ** (-lr_power)) / lr * theta
    quad = n_new ** (-lr_power) / lr + 2.0 * l2
    out = np.where(np.abs(z) > l1, (l1 * np.sign(z) - z) / quad, 0.0)
    n[...] = n_new
    return out


# --------------------------------------------------------------------------------------------------
# FM (FM.py)
# --------------------------------------------------------------------------------------------------
class OracleFM:
    """table[:,0] = W, table[:,1:] = V (FM.py:15-20); b scalar."""

    def __init__(self, table, b, dtype=np.float64):
        self.dtype = dtype
        self.table = np.array(table, dtype=dtype)
        self.b = dtype(b)

    def logits(self, ids, wts):
        """FM.factorization, FM.py:55-64."""
        x = np.asarray(wts, dtype=self.dtype)
        rows = self.table[ids]                                  # [N,F,d]
        W, V = rows[..., 0], rows[..., 1:]
        yhat = (W * x).sum(1) + self.b                          # :57
        vx = (V * x[..., None]).sum(1)                          # :59
        v2x2 = ((V ** 2) * (x ** 2)[..., None]).sum(1)          # :61
        return yhat + 0.5 * ((vx ** 2).sum(1) - v2x2.sum(1)), vx   # :63

    def predict(self, ids, wts):
        return sigmoid(self.logits(ids, wts)[0])

    def train_step(self, ids, wts, y, lr, lam=0.0, reduce="sum", dense_l2=True):
        """loss = reduce(xent) + lam*(l2_loss(W)+l2_loss(V)+l2_loss(b)), l2_loss = sum(t^2)/2 (FM.py:33-39);
        plain SGD (tf_util.py:26-28).  dense_l2=False restricts the decay to the touched rows."""
        x = np.asarray(wts, dtype=self.dtype)
        y = np.asarray(y, dtype=self.dtype)
        N, F = ids.shape
        z, vx = self.logits(ids, x)
        scale = 1.0 if reduce == "sum" else 1.0 / N
        loss = xent_with_logits(z, y).sum() * scale
        if lam:
            loss += lam * 0.5 * ((self.table ** 2).sum() + self.b ** 2)
        g = (sigmoid(z) - y) * scale
        rows = self.table[ids]
        grad = np.empty_like(rows)
        grad[..., 0] = g[:, None] * x
        grad[..., 1:] = g[:, None, None] * (x[..., None] * vx[:, None, :] - (x ** 2)[..., None] * rows[..., 1:])
        dense = np.zeros_like(self.table)
        np.add.at(dense, ids.reshape(-1), grad.reshape(N * F, -1))
        if lam:
            if dense_l2:
                dense += lam * self.table
            else:
                touched = np.unique(ids)
                dense[touched] += lam * self.table[touched]
        self.table -= lr * dense
        self.b = self.b - lr * (g.sum() + lam * self.b)
        return loss, sigmoid(z)

    def train_step_adam(self, ids, wts, y, lr, eps=1e-8, lam=0.0, reduce="sum", beta1=0.9, beta2=0.999):
        """Same loss (FM.py:33-39, LR.py:30-37), minimised by tf.train.AdamOptimizer(lr, epsilon) (tf_util.py:18-21): dense
        over W, V and b, the L2 term's gradient lam*theta included in g."""
        x = np.asarray(wts, dtype=self.dtype)
        y = np.asarray(y, dtype=self.dtype)
        N, F = ids.shape
        z, vx = self.logits(ids, x)
        scale = 1.0 if reduce == "sum" else 1.0 / N
        loss = xent_with_logits(z, y).sum() * scale
        if lam:
            loss += lam * 0.5 * ((self.table ** 2).sum() + self.b ** 2)
        g = (sigmoid(z) - y) * scale
        rows = self.table[ids]
        grad = np.empty_like(rows)
        grad[..., 0] = g[:, None] * x
        grad[..., 1:] = g[:, None, None] * (x[..., None] * vx[:, None, :] - (x ** 2)[..., None] * rows[..., 1:])
        dense = np.zeros_like(self.table)
        np.add.at(dense, ids.reshape(-1), grad.reshape(N * F, -1))
        dense += lam * self.table
        gb = g.sum() + lam * self.b
        if not hasattr(self, "_adam"):
            self._adam = dict(t=0, m=np.zeros_like(self.table), v=np.zeros_like(self.table), mb=0.0, vb=0.0)
        st = self._adam
        st["t"] += 1
        lr_t = lr * np.sqrt(1.0 - beta2 ** st["t"]) / (1.0 - beta1 ** st["t"])
        st["m"] = beta1 * st["m"] + (1 - beta1) * dense
        st["v"] = beta2 * st["v"] + (1 - beta2) * dense * dense
        self.table -= lr_t * st["m"] / (np.sqrt(st["v"]) + eps)
        st["mb"] = beta1 * st["mb"] + (1 - beta1) * gb
        st["vb"] = beta2 * st["vb"] + (1 - beta2) * gb * gb
        self.b = self.b - lr_t * st["mb"] / (np.sqrt(st["vb"]) + eps)
        return loss, sigmoid(z)


    def train_step_ftrl(self, ids, wts, y, lr, lam=0.0, reduce="sum", init_accum=0.1, lr_power=-0.5, l1=0.0, l2=0.0):
        """Same loss (FM.py:33-39, LR.py:30-37), minimised by tf.train.FtrlOptimizer(lr) (tf_util.py:22-24): dense over W, V
        and b, the L2 term's gradient lam*theta included in g (see ftrl_apply)."""
        x = np.asarray(wts, dtype=self.dtype)
        y = np.asarray(y, dtype=self.dtype)
        N, F = ids.shape
        z, vx = self.logits(ids, x)
        scale = 1.0 if reduce == "sum" else 1.0 / N
        loss = xent_with_logits(z, y).sum() * scale
        if lam:
            loss += lam * 0.5 * ((self.table ** 2).sum() + self.b ** 2)
        g = (sigmoid(z) - y) * scale
        rows = self.table[ids]
        grad = np.empty_like(rows)
        grad[..., 0] = g[:, None] * x
        grad[..., 1:] = g[:, None, None] * (x[..., None] * vx[:, None, :] - (x ** 2)[..., None] * rows[..., 1:])
        dense = np.zeros_like(self.table)
        np.add.at(dense, ids.reshape(-1), grad.reshape(N * F, -1))
        dense += lam * self.table
        gb = np.asarray(g.sum() + lam * self.b, dtype=self.dtype)
        if not hasattr(self, "_ftrl"):
            self._ftrl = dict(t=[np.full_like(self.table, init_accum), np.zeros_like(self.table)],
                              b=[np.asarray(init_accum, dtype=self.dtype).copy(), np.asarray(0.0, dtype=self.dtype).copy()])
        st = self._ftrl
        kw = dict(lr_power=lr_power, l1=l1, l2=l2)
        self.table[...] = ftrl_apply(self.table, dense, st["t"], lr, **kw)
        self.b = self.dtype(ftrl_apply(np.asarray(self.b, dtype=self.dtype), gb, st["b"], lr, **kw))
        return loss, sigmoid(z)


# --------------------------------------------------------------------------------------------------
# SNN fine-tuning step (SNN_DAE.py:103-160, 234-288; SNN_RBM.py is the same graph) -- SURVEY 8(f) f4
# --------------------------------------------------------------------------------------------------
class OracleSNN:
    """x_dense = sigmoid(sum_{feat in sample} ww0[feat] + bb0) on the host (get_fi_h1_y, SNN_DAE.py:234-260), then the
    Theano graph z1 = x.w1 + b1, h1 = act(z1), d1 = h1*r1, d2 = tanh(d1.w2 + b2)*r2, p_drop = sigmoid(d2.w3 + b3),
    cost = xent.sum() + lambda1*(|w1|^2 + |w2|^2 + |w3|^2 + |b1|^2 + |b2|^2 + b3^2) (:113-153), SGD on w1..b3 (:157-163) and
    the host loop  bb0 -= lr*gx_t*x_t*(1-x_t);  ww0[feat] -= lr*gx_t*x_t*(1-x_t)  per sample and feature (:282-288)."""

    def __init__(self, ww0, bb0, Ws, bs, act="tanh", dtype=np.float64):
        self.dtype = dtype
        self.ww0, self.bb0 = np.array(ww0, dtype=dtype), np.array(bb0, dtype=dtype)
        self.Ws = [np.array(w, dtype=dtype).reshape(w.shape[0], -1) for w in map(np.asarray, Ws)]
        self.bs = [np.array(b, dtype=dtype).reshape(-1) for b in bs]
        self.act = act

    def dense(self, feats):
        """feats: list of feature-id lists (ragged) or an [N, J] array padded with -1."""
        x = np.zeros((len(feats), self.ww0.shape[1]), dtype=self.dtype)
        for t, ft in enumerate(feats):
            for feat in ft:
                if feat >= 0:
                    x[t] += self.ww0[feat]                        # :248
        return sigmoid(x + self.bb0)                              # :250-253

    def _mlp(self, x, r1, r2):
        h1 = activate(self.act, x @ self.Ws[0] + self.bs[0])
        d1 = h1 * r1
        d2 = np.tanh(d1 @ self.Ws[1] + self.bs[1]) * r2           # :137 -- tanh whatever acti_type says
        logit = (d2 @ self.Ws[2] + self.bs[2]).reshape(-1)
        return h1, d1, d2, logit

    def predict(self, feats):
        x = self.dense(feats)
        h1 = activate(self.act, x @ self.Ws[0] + self.bs[0])
        h2 = activate(self.act, h1 @ self.Ws[1] + self.bs[1])     # :127-133
        return sigmoid((h2 @ self.Ws[2] + self.bs[2]).reshape(-1))

    def loss_and_grads(self, x, y, r1, r2, lambda1=0.0):
        h1, d1, d2, logit = self._mlp(x, r1, r2)
        loss = xent_with_logits(logit, y).sum() + lambda1 * (sum((w ** 2).sum() for w in self.Ws) + sum((b ** 2).sum() for b in self.bs))
        dl = (sigmoid(logit) - y).reshape(-1, 1)
        gW3, gb3 = d2.T @ dl, dl.sum(0)
        dd2 = dl @ self.Ws[2].T
        dz2 = dd2 * r2 * (1 - np.where(r2 != 0, d2 / np.where(r2 != 0, r2, 1), 0.0) ** 2)
        gW2, gb2 = d1.T @ dz2, dz2.sum(0)
        dz1 = (dz2 @ self.Ws[1].T) * r1 * activate_grad_from_out(self.act, h1)
        gW1, gb1 = x.T @ dz1, dz1.sum(0)
        gx = dz1 @ self.Ws[0].T
        gWs = [gW1 + 2 * lambda1 * self.Ws[0], gW2 + 2 * lambda1 * self.Ws[1], gW3 + 2 * lambda1 * self.Ws[2]]
        gbs = [gb1 + 2 * lambda1 * self.bs[0], gb2 + 2 * lambda1 * self.bs[1], gb3 + 2 * lambda1 * self.bs[2]]
        return loss, sigmoid(logit), gx, gWs, gbs

    def train_step(self, feats, y, lr, r1=1.0, r2=1.0, lambda1=0.0, faithful=False):
        y = np.asarray(y, dtype=self.dtype)
        x = self.dense(feats)
        loss, p, gx, gWs, gbs = self.loss_and_grads(x, y, np.asarray(r1, dtype=self.dtype), np.asarray(r2, dtype=self.dtype), lambda1)
        gs = gx * x * (1 - x)
        if faithful:                                              # the literal loop of SNN_DAE.py:282-288
            for t in range(len(feats)):
                self.bb0 = self.bb0 - lr * gs[t]
                for feat in feats[t]:
                    if feat >= 0:
                        self.ww0[feat] = self.ww0[feat] - lr * gs[t]
        else:
            self.bb0 = self.bb0 - lr * gs.sum(0)
            for t, ft in enumerate(feats):
                ft = np.asarray([f for f in ft if f >= 0], dtype=np.int64)
                np.add.at(self.ww0, ft, -lr * gs[t])
        for l in range(3):
            self.Ws[l] = self.Ws[l] - lr * gWs[l]
            self.bs[l] = self.bs[l] - lr * gbs[l]
        return loss, p


# --------------------------------------------------------------------------------------------------
# evaluation-side host logic of the driver (baseline.py:262-281, 369-370)
# --------------------------------------------------------------------------------------------------
def early_stop_literal(step, errs, metric, least_step, skip_window, smooth_window, stop_window):
    """early_stop, baseline.py:262-281, statement by statement (prints dropped)."""
    if step > least_step:
        skip_metric = list(errs)[::skip_window]
        smooth_metric = np.array(skip_metric[smooth_window - 1:], dtype=np.float64)
        for i in range(smooth_window - 1):
            smooth_metric += skip_metric[i:(i - smooth_window + 1)]
        smooth_metric /= smooth_window
        if len(smooth_metric) < stop_window:
            return False
        smooth_error = smooth_metric[stop_window - 1:] - smooth_metric[:1 - stop_window]
        if metric == "rmse" and smooth_error[-1] > 0:
            return True
        elif metric == "auc" and smooth_error[-1] < 0:
            return True
        return False
    return False


def recalibrate_literal(eval_preds, nds_rate):
    """baseline.py:369-370: eval_preds /= eval_preds + (1 - eval_preds) / nds_rate."""
    eval_preds = np.array(eval_preds, dtype=np.float64)
    eval_preds /= eval_preds + (1 - eval_preds) / nds_rate
    return eval_preds


# --------------------------------------------------------------------------------------------------
# initialisation (FNN_wnzh.py:15-17,106-130) -- reproducible with numpy's legacy RandomState alone
# --------------------------------------------------------------------------------------------------
def theano_fnn_init(xdim, hidden1, hidden2, acti_type="tanh", seed=1234):
    rng = np.random.RandomState(seed)
    w = rng.uniform(low=-np.sqrt(6.0 / (xdim + hidden1)), high=np.sqrt(6.0 / (xdim + hidden1)), size=(xdim, hidden1))
    if acti_type == "sigmoid":
        ww1 = w
    elif acti_type == "tanh":
        ww1 = w * 4
    else:
        ww1 = rng.uniform(-1, 1, size=(xdim, hidden1))
    v = rng.uniform(low=-np.sqrt(6.0 / (hidden1 + hidden2)), high=np.sqrt(6.0 / (hidden1 + hidden2)), size=(hidden1, hidden2))
    if acti_type == "sigmoid":
        ww2 = v
    elif acti_type == "tanh":
        ww2 = v * 4
    else:
        ww2 = rng.uniform(-1, 1, size=(hidden1, hidden2))
    return [ww1, ww2, np.zeros((hidden2, 1))], [np.zeros(hidden1), np.zeros(hidden2), np.zeros(1)]
