"""Fixed seeded cases on which the REFERENCE'S OWN code is executed (oracle/ref_run.py: the converted /root/reference
sources on the theano / tensorflow shims).  Shared by tests/test_oracle_vs_ref.py (oracle == reference, CPU) and
tests/golden/make_golden_ref.py (vectors for the GPU parity tests, which cannot see /root/reference).
Test infrastructure only."""
import os
import pickle

import numpy as np

FIELD_SIZES = np.array([7, 24, 12, 40, 10, 20, 5, 30, 30, 8, 6, 4, 3, 20, 5, 12])      # 16 iPinYou-style fields, X = 236


def write_s1_demo(workdir, n_train=300, n_test=200, seed=1234, dup_field=False):
    """The S1 demo set of SURVEY 8(d) in the reference's text formats (fm.model.txt / train.fm.txt / test.fm.txt)."""
    from deepctr_b200 import data as D
    rng = np.random.default_rng(seed)
    F, d = 16, 11
    X = int(FIELD_SIZES.sum())
    feat_field = np.repeat(np.arange(F), FIELD_SIZES)
    table = rng.uniform(-0.05, 0.05, (X, d))
    os.makedirs(os.path.join(workdir, "data"), exist_ok=True)
    D.write_fm_model(os.path.join(workdir, "data", "fm.model.txt"), 0.1, table, feat_field)
    sets = {}
    for name, n in (("train", n_train), ("test", n_test)):
        ids = D.make_field_batch(rng, n, FIELD_SIZES)
        y = D.planted_labels(rng, ids, table * 20, pos_rate=0.3)
        if dup_field and name == "train":
            # adversarial lines: a sample with TWO features of one field (the later one overwrites the slot, both get updated,
            # FNN_wnzh.py:91-95, 300-306) and a sample that lacks a field altogether
            ids = [list(r) for r in ids]
            ids[3].append(int(np.flatnonzero(feat_field == 5)[2]))
            ids[7] = ids[7][:9] + ids[7][10:]
            ids[11].insert(0, int(np.flatnonzero(feat_field == 15)[1]))
        with open(os.path.join(workdir, "data", "%s.fm.txt" % name), "w") as f:
            for t in range(n):
                f.write("%d %s\n" % (y[t], " ".join("%d:1" % i for i in ids[t])))
        sets[name] = (ids, y)
    return dict(table=table, feat_field=feat_field, w_0=0.1, sets=sets, F=F, d=d, X=X)


# ---- TF-form cases -------------------------------------------------------------------------------------------------
def tf_case(name, seed=7):
    """Small Criteo-shaped configuration for one TF model class.  The reference hard-codes 13 numeric fields
    (FNN.py:42,47, FNN_IP_L3.py:86), so every FNN-family case has 13 numeric + len(cat_sizes) categorical fields."""
    rng = np.random.default_rng(seed + sum(map(ord, name)))
    c = dict(name=name, rank=3, act="tanh", N=6, E=4, lr=0.05, keep=1.0, reduce="sum", ptmzr="sgd", steps=3)
    if name in ("FM", "LR"):
        c.update(X_dim=40, X_feas=5, lam=0.01, rank=3 if name == "FM" else 0)
    else:
        # ... and 26 categorical ones: the placeholders are declared [batch, 13] / [batch, 26] (FNN.py:41-43)
        c.update(cat_sizes=np.array([9, 4, 17, 3, 5, 2, 6, 4, 3, 8, 2, 5, 3, 4, 7, 2, 3, 5, 4, 2, 6, 3, 4, 2, 5, 3]), N=4, E=3,
                 n_hidden={"FNN": 2, "FNN_IP_L3": 3, "FNN_OP_L3": 3, "FNN_IP_L5": 5, "FNN_IP_L7": 7}[name])
        c["hidden"] = [10, 8, 7, 6, 5, 4, 3][:c["n_hidden"]]
        c["cross"] = "ip" if "IP" in name else ("op" if "OP" in name else "none")
    c["rng"] = rng
    return c


def fnn_geometry(c):
    cat_sizes = c["cat_sizes"]
    offsets = [13 + int(cat_sizes[:i].sum()) for i in range(len(cat_sizes))]
    X_dim, X_feas, d = 13 + int(cat_sizes.sum()), 13 + len(cat_sizes), c["rank"] + 1
    cw = {"none": 0, "ip": X_feas * (X_feas - 1) // 2, "op": d * d}[c["cross"]]
    return cat_sizes, offsets, X_dim, X_feas, d, cw


def fnn_var_map(c, declared_width=None):
    """random float32 checkpoint for an FNN-family case; declared_width overrides the first layer's input width (the
    reference's FNN_OP_L3 declares it with P instead of d*d, FNN_OP_L3.py:18,24)."""
    rng = c["rng"]
    cat_sizes, offsets, X_dim, X_feas, d, cw = fnn_geometry(c)
    dims = [declared_width or (X_feas * d + cw + 1)] + list(c["hidden"]) + [1]
    vm = {"W": rng.uniform(-0.3, 0.3, (X_dim, 1)), "V": rng.uniform(-0.3, 0.3, (X_dim, c["rank"])), "b": np.array([0.1])}
    for i in range(len(dims) - 1):
        vm["h%d_w" % (i + 1)] = rng.uniform(-1, 1, (dims[i], dims[i + 1])) / np.sqrt(dims[i])
        vm["h%d_b" % (i + 1)] = rng.uniform(-0.1, 0.1, dims[i + 1])
    return {k: v.astype(np.float32) for k, v in vm.items()}, dims


def fnn_batch(c, n):
    rng = c["rng"]
    cat_sizes, offsets, X_dim, X_feas, d, cw = fnn_geometry(c)
    c_ids = np.stack([rng.integers(0, s, n) for s in cat_sizes], 1)
    c_ids[n // 2] = c_ids[0]                                   # duplicate rows inside the batch
    v_wts = rng.random((n, 13)).astype(np.float32)
    c_wts = np.ones((n, len(cat_sizes)), np.float32)
    y = rng.integers(0, 2, n)
    return v_wts, c_ids, c_wts, y


def dump_pickle(vm, path):
    with open(path, "wb") as f:
        pickle.dump(vm, f, protocol=2)
    return path
