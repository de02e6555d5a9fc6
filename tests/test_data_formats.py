"""Host-side data formats either side of the hot path: text writers / loaders round-trip
bit-exactly (ids, offsets, pad indices are integer work), and the loaders follow the reference's parsing rules."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepctr_b200 import data   # noqa: E402


def test_field_sizes_and_offsets():
    s = data.ipinyou_field_sizes(1_000_000)
    assert len(s) == 16 and int(s.sum()) == 1_000_000 and s.min() >= 2
    c = data.scaled_criteo_cat_sizes(100_000_013)
    assert len(c) == 26 and 13 + int(c.sum()) == 100_000_013
    off = data.offsets_of(data.CRITEO_CAT_SIZES, 13)                        # baseline.py:40
    assert off[0] == 13 and np.array_equal(np.diff(off), data.CRITEO_CAT_SIZES[:-1])
    assert 13 + int(data.CRITEO_CAT_SIZES.sum()) == 1_636_608               # the 1.6 M-row configuration of bench.py
    rng = np.random.default_rng(0)
    ids = data.make_field_batch(rng, 500, s, 1.1)
    lo = data.offsets_of(s)
    assert ids.shape == (500, 16) and np.all(ids >= lo) and np.all(ids < lo + s)      # every id inside its field's block


def test_fm_model_round_trip(tmp_path):
    """a13: `w_0 feat_num k` header, then `feat w v_1..v_k field:value` (FNN_wnzh.py:68-84)"""
    rng = np.random.default_rng(1)
    X, d = 64, 5
    table = rng.normal(0, 1, (X, d))
    feat_field = np.repeat(np.arange(16), X // 16)
    p = str(tmp_path / "fm.model.txt")
    data.write_fm_model(p, -0.25, table, feat_field)
    m = data.load_fm_model(p)
    assert m["w_0"] == -0.25 and m["k"] == d
    assert np.array_equal(m["table"], table)                               # %.17g round-trips a double exactly
    assert np.array_equal(m["feat_field"], feat_field)
    # a model file that skips feature ids: absent rows stay zero and are marked unknown (-1)
    with open(p) as f:
        lines = f.readlines()
    with open(p, "w") as f:
        f.writelines(lines[:1] + lines[1:4] + lines[10:])
    m2 = data.load_fm_model(p)
    assert np.all(m2["feat_field"][3:9] == -1) and np.all(m2["table"][3:9] == 0) and np.array_equal(m2["table"][9:], table[9:])


def test_fm_samples_round_trip(tmp_path):
    """a14: `y id:val id:val ...`, ids at the odd positions, values ignored (FNN_wnzh.py:240-253); ragged rows pad with -1"""
    p = str(tmp_path / "train.fm.txt")
    ids = [[3, 17, 40], [5, 9], [1, 2, 3]]
    data.write_fm_samples(p, ids, [1, 0, 1])
    feats, y = data.load_fm_samples(p)
    assert np.array_equal(y, [1, 0, 1])
    assert np.array_equal(feats, [[3, 17, 40], [5, 9, -1], [1, 2, 3]])
    with open(p, "a") as f:
        f.write("0 7:0.5 8:2\n\n")                                          # values other than 1 are ignored; blank line skipped
    feats, y = data.load_fm_samples(p)
    assert feats.shape == (4, 3) and np.array_equal(feats[3], [7, 8, -1]) and y[3] == 0


def test_yzx_format(tmp_path):
    """a15: `y z id:val ...`; X_dim = max_id + 2, rows padded to the max nnz with (id = X_dim - 1, val = 0), vals forced
    to 1 (ipinyou.py:23-65)"""
    p = str(tmp_path / "train.yzx.txt")
    ids = np.array([[2, 9, 30], [4, -1, -1], [0, 31, -1]])
    data.write_yzx(p, ids, [0, 1, 0], z=77)
    max_dim, max_fea = data.stat_yzx(p)
    assert (max_dim, max_fea) == (31, 3)
    X_dim = max_dim + 2
    with open(p) as f:
        ind, val, y = data.load_yzx(f.readlines(), X_dim - 1, max_fea)
    assert np.array_equal(y, [0, 1, 0])
    assert np.array_equal(ind, [[2, 9, 30], [4, 32, 32], [0, 31, 32]])
    assert np.array_equal(val, [[1, 1, 1], [1, 0, 0], [1, 1, 0]])


def test_criteo_tsv(tmp_path):
    """a16: `y \\t n_1..n_13 \\t c_1..c_26`; numeric / max_vals, categorical local index + offsets (baseline.py:190-198)"""
    rng = np.random.default_rng(2)
    sizes = data.CRITEO_CAT_SIZES
    offsets = data.offsets_of(sizes, 13)
    n = 20
    num_raw = np.stack([rng.integers(0, mv + 1, n) for mv in data.CRITEO_MAX_VALS], 1)
    cat = np.stack([rng.integers(0, s, n) for s in sizes], 1)
    y = rng.integers(0, 2, n)
    p = str(tmp_path / "day_0.tsv")
    data.write_criteo_tsv(p, y, num_raw, cat)
    with open(p) as f:
        for t, line in enumerate(f):
            yy, fidx, x = data.parse_criteo_line(line, offsets)
            assert yy == y[t]
            assert fidx[:13] == list(range(13)) and fidx[13:] == list(cat[t] + offsets)          # bit-exact index arithmetic
            np.testing.assert_allclose(x[:13], num_raw[t] / data.CRITEO_MAX_VALS, rtol=1e-15)
            assert x[13:] == [1] * 26
    ids, wts = data.make_criteo_batch(rng, 64, sizes)
    assert ids.shape == (64, 39) and np.array_equal(ids[:, :13], np.broadcast_to(np.arange(13), (64, 13)))
    assert np.all(ids[:, 13:] >= offsets) and np.all(ids[:, 13:] < offsets + sizes) and np.all(wts[:, 13:] == 1)


def test_planted_labels_rate():
    rng = np.random.default_rng(3)
    sizes = data.ipinyou_field_sizes(2048)
    ids = data.make_field_batch(rng, 4000, sizes, 1.1)
    table = rng.normal(0, 0.3, (2048, 11))
    y = data.planted_labels(rng, ids, table, pos_rate=0.2)
    assert set(np.unique(y)) <= {0, 1} and 0.1 < y.mean() < 0.3
