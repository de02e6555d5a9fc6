"""Synthetic data of the reference's shapes and readers / writers for its on-disk text formats.

The datasets the reference scripts open (`../data/...`) are not part of the snapshot, so the
formats are reconstructed from the parsers:
  fm.model.txt            FNN_wnzh.py:68-84     `w_0 feat_num k` then `feat w v_1..v_k field:value`
  train.fm.txt            FNN_wnzh.py:240-253   `y id:val id:val ...`
  *.yzx.txt               ipinyou.py:42-65      `y z id:val ...`, rows padded to X_feas with (X_dim-1, 0)
  Criteo indexed TSV      baseline.py:190-198   `y \t n_1..n_13 \t c_1..c_26` (field-local categorical ids)
"""
import numpy as np

from .theano_fnn import NAME_FIELD

FIELD_NAMES = sorted(NAME_FIELD, key=NAME_FIELD.get)

# relative sizes of the 16 iPinYou fields (a few huge: IP, domain, slotid; many tiny: weekday 7, hour 24)
_IPINYOU_SHAPE = [7, 24, 40, 500000, 35, 370, 5, 25000, 50000, 21, 14, 11, 4, 130, 9, 280]
# Criteo nds 2.5 %, frequency filter > 10 (baseline.py:31-32); +1 "other" bucket each (baseline.py:39)
CRITEO_CAT_SIZES = np.array([284414, 22289, 15249, 7022, 18956, 3, 6730, 1286, 50, 253234, 93011, 81371,
                             10, 2187, 8396, 61, 4, 932, 14, 286828, 190408, 274540, 79781, 9690, 70, 33]) + 1
CRITEO_MAX_VALS = np.array([65535, 8000, 715, 376613, 7995, 1430, 2191, 25462, 5337, 8, 221, 3023897, 8072])


def ipinyou_field_sizes(X=1_000_000):
    """16 field cardinalities summing to X, proportioned like the iPinYou fields (small fields kept exact)."""
    base = np.array(_IPINYOU_SHAPE, dtype=np.float64)
    small = base < 1000
    rest = X - int(base[small].sum())
    sizes = base.copy()
    sizes[~small] = np.floor(base[~small] / base[~small].sum() * rest)
    sizes = sizes.astype(np.int64)
    sizes[3] += X - sizes.sum()
    return sizes


def scaled_criteo_cat_sizes(X_total):
    """26 categorical cardinalities with the reference proportions scaled so that 13 + sum == X_total."""
    base = CRITEO_CAT_SIZES.astype(np.float64)
    sizes = np.maximum(2, np.floor(base / base.sum() * (X_total - 13))).astype(np.int64)
    sizes[np.argmax(sizes)] += (X_total - 13) - sizes.sum()
    return sizes


def offsets_of(sizes, first=0):
    """offsets[i] = first + sum(sizes[:i])  (baseline.py:40 with first = 13)."""
    return first + np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)


def zipf_local_ids(rng, n, size, a=1.1):
    """n draws from a truncated power law p(rank) ~ rank^-a over [0, size), rank spread by a bijective hash."""
    u = rng.random(n)
    if abs(a - 1.0) < 1e-9:
        r = np.exp(u * np.log(size + 1.0))
    else:
        r = (((size + 1.0) ** (1.0 - a) - 1.0) * u + 1.0) ** (1.0 / (1.0 - a))
    rank = np.minimum(np.floor(r).astype(np.int64) - 1, size - 1)
    mult = 2654435761 % size if size > 1 else 0
    while size > 1 and np.gcd(mult, size) != 1:
        mult += 1
    return (rank * mult) % size if size > 1 else rank


def make_field_batch(rng, N, sizes, a=1.1, first=0):
    """[N, len(sizes)] GLOBAL ids, one per field, Zipf(a) inside each field."""
    offs = offsets_of(sizes, first)
    cols = [zipf_local_ids(rng, N, int(s), a) + o for s, o in zip(sizes, offs)]
    return np.stack(cols, axis=1)


def make_criteo_batch(rng, N, cat_sizes=CRITEO_CAT_SIZES, a=1.05):
    """(cols [N,39] global ids, vals [N,39]) as get_fxy produces them (baseline.py:190-198): ids 0..12 with the
    scaled numeric value, then offsets[i] + local id with value 1."""
    cat = make_field_batch(rng, N, cat_sizes, a, first=13)
    cols = np.concatenate([np.broadcast_to(np.arange(13), (N, 13)), cat], axis=1)
    vals = np.concatenate([rng.random((N, 13)), np.ones((N, len(cat_sizes)))], axis=1).astype(np.float32)
    return cols, vals


def planted_labels(rng, ids, table, w0=0.0, pos_rate=0.2):
    """Labels from a planted FM over `table` ([X, d]); the intercept is shifted to hit ~pos_rate positives."""
    rows = table[ids]
    s = rows[..., 1:].sum(1)
    z = w0 + rows[..., 0].sum(1) + 0.5 * ((s ** 2).sum(1) - (rows[..., 1:] ** 2).sum((1, 2)))
    z = z - np.quantile(z, 1 - pos_rate)
    return (rng.random(len(z)) < 1 / (1 + np.exp(-4 * z))).astype(np.int64)


# ---- text formats --------------------------------------------------------------------------------
def write_fm_model(path, w0, table, feat_field):
    X, d = table.shape
    with open(path, "w") as f:
        f.write("%.17g %d %d\n" % (w0, X, d - 1))
        for feat in range(X):
            f.write("%d %s %s:%d\n" % (feat, " ".join("%.17g" % v for v in table[feat]), FIELD_NAMES[feat_field[feat]], feat))


def load_fm_model(path, n_fields=16):
    """FNN_wnzh.py:62-84 -> dict(w_0, table [max_feat+1, k+1], feat_field [max_feat+1] (-1 = unknown feature))."""
    feats, rows, fields = [], [], []
    with open(path) as f:
        s = f.readline().split()
        w0, k = float(s[0]), int(s[2]) + 1
        for line in f:
            s = line.strip().split()
            if not s:
                continue
            feats.append(int(s[0]))
            rows.append([float(s[1 + i]) for i in range(k)])
            fields.append(NAME_FIELD[s[1 + k][:s[1 + k].index(":")]])
    X = max(feats) + 1
    table = np.zeros((X, k))
    feat_field = np.full(X, -1, dtype=np.int64)
    table[feats] = rows
    feat_field[feats] = fields
    return dict(w_0=w0, table=table, feat_field=feat_field, k=k)


def write_fm_samples(path, ids, y):
    with open(path, "w") as f:
        for t in range(len(y)):
            f.write("%d %s\n" % (y[t], " ".join("%d:1" % i for i in ids[t])))


def load_fm_samples(path):
    """get_fxy of FNN_wnzh.py:248-253: y first, ids at the odd positions after ':' -> ' ', values ignored.
    Returns (feats [n, max_nnz] padded with -1, y [n])."""
    ys, fs = [], []
    with open(path) as f:
        for line in f:
            s = line.replace(":", " ").split()
            if not s:
                continue
            ys.append(int(s[0]))
            fs.append([int(s[j]) for j in range(1, len(s), 2)])
    m = max(len(r) for r in fs)
    feats = np.full((len(fs), m), -1, dtype=np.int64)
    for i, r in enumerate(fs):
        feats[i, :len(r)] = r
    return feats, np.array(ys, dtype=np.int64)


def write_yzx(path, ids, y, z=0):
    with open(path, "w") as f:
        for t in range(len(y)):
            f.write("%d %d %s\n" % (y[t], z, " ".join("%d:1" % i for i in ids[t] if i >= 0)))


def stat_yzx(path):
    """ipinyou.py:23-39: (max index, max nnz)."""
    max_fea = max_dim = 0
    with open(path) as f:
        for line in f:
            fields = line.strip().split()
            max_fea = max(max_fea, len(fields) - 2)
            max_dim = max(max_dim, max(int(x.split(":")[0]) for x in fields[2:]))
    return max_dim, max_fea


def load_yzx(lines, max_dim, max_fea):
    """load_ipinyou_data of ipinyou.py:42-65 without the shuffle: pad to max_fea with (max_dim, 0); values forced to 1."""
    X_ind, X_val, y = [], [], []
    for line in lines:
        fields = line.strip().split()
        y.append(int(fields[0]))
        ind = [int(x.split(":")[0]) for x in fields[2:]]
        X_val.append([1] * len(ind) + [0] * (max_fea - len(ind)))
        X_ind.append(ind + [max_dim] * (max_fea - len(ind)))
    return np.array(X_ind), np.array(X_val), np.array(y)


def write_criteo_tsv(path, y, num_raw, cat_local):
    with open(path, "w") as f:
        for t in range(len(y)):
            f.write("%d\t%s\t%s\n" % (y[t], "\t".join("%d" % v for v in num_raw[t]), "\t".join("%d" % v for v in cat_local[t])))


def parse_criteo_line(line, offsets, max_vals=CRITEO_MAX_VALS):
    """get_fxy of baseline.py:190-198."""
    fields = line.rstrip("\n").split("\t")
    y = int(fields[0])
    cats = fields[14:]
    f = list(range(13)) + [int(cats[i]) + int(offsets[i]) for i in range(len(offsets))]
    x = [float(fields[i]) / max_vals[i - 1] for i in range(1, 14)] + [1] * len(offsets)
    return y, f, x
