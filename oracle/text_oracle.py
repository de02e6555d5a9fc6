"""CPU restatement of the reference's TEXT PARSERS -- test infrastructure only (checker of the device-side ingestion,
deep-ctr_b200/csrc/text.cu); nothing in the product path imports it.

Parity: pinned against the reference's own parsers executed here (oracle/_ref): get_fxy / get_batch_xy of baseline.py and
collect / stat / load_ipinyou_data of ipinyou.py in tests/test_oracle_vs_ref.py, the FNN_wnzh.py line parser through the whole
script run.  The reference ships no data files: the files are written in the formats those parsers imply.
"""
import numpy as np


def nonblank_lines(text):
    """What the reference loops ever hand to a parser: `linecache.getline` + `if line.strip() != ''` (FNN_wnzh.py:229-230),
    `for line in buf` (ipinyou.py:50), `next(data_set)` + `if len(_line.strip())` (baseline.py:214-215)."""
    return [ln for ln in text.split("\n") if ln != ""]


def fm_get_fxy(line):
    """get_fxy, FNN_wnzh.py:248-252: s = line.replace(':', ' ').split(); y = int(s[0]); feats = s[1], s[3], ..."""
    s = line.replace(":", " ").split()
    y = int(s[0])
    feats = [int(s[j]) for j in range(1, len(s), 2)]
    return feats, y


def yzx_line(line):
    """load_ipinyou_data, ipinyou.py:51-56: fields = line.strip().split(); y = int(fields[0]);
    x_ind = [int(x.split(':')[0]) for x in fields[2:]]; x_val = [1] * len(x_ind)."""
    fields = line.strip().split()
    y = int(fields[0])
    x_ind = [int(x.split(":")[0]) for x in fields[2:]]
    return x_ind, [1] * len(x_ind), y


def sparse_block(lines, J, pad_id, skip):
    """[n, J] feature ids padded with (pad_id, 0) -- ipinyou.py:57-59 pads to max_fea with (max_dim, 0); skip = 0 parses
    train.fm.txt lines, skip = 1 yzx lines.  Returns feats int64, wts float32, y int64, nnz int64 (before truncation)."""
    n = len(lines)
    feats = np.full((n, J), pad_id, dtype=np.int64)
    wts = np.zeros((n, J), dtype=np.float32)
    y = np.zeros(n, dtype=np.int64)
    nnz = np.zeros(n, dtype=np.int64)
    for i, ln in enumerate(lines):
        if skip == 0:
            f, yy = fm_get_fxy(ln)
        else:
            f, _, yy = yzx_line(ln)
        y[i], nnz[i] = yy, len(f)
        k = min(len(f), J)
        feats[i, :k] = f[:k]
        wts[i, :k] = 1.0
    return feats, wts, y, nnz


def criteo_get_fxy(line, offsets, max_vals, n_num=13):
    """get_fxy, baseline.py:190-198: fields = _line.split('\\t'); _y = int(fields[0]); cats = fields[14:];
    _f = range(13) + [int(cats[i]) + offsets[i]]; _x = [float(fields[i]) / max_vals[i-1] for i in 1..13] + [1] * 26."""
    fields = line.split("\t")
    y = int(fields[0])
    cats = fields[1 + n_num:]
    f = list(range(n_num)) + [int(cats[i]) + int(offsets[i]) for i in range(len(offsets))]
    x = [float(fields[i]) / max_vals[i - 1] for i in range(1, 1 + n_num)] + [1] * len(offsets)
    return y, f, x


def criteo_block(lines, offsets, max_vals, n_num=13):
    ys, fs, xs = [], [], []
    for ln in lines:
        y, f, x = criteo_get_fxy(ln, offsets, max_vals, n_num)
        ys.append(y)
        fs.append(f)
        xs.append(x)
    # np.array(_vals) is float64; the placeholders are tf.float32 (FNN.py:41-43): the feed casts
    return np.array(fs, dtype=np.int64), np.array(xs, dtype=np.float64).astype(np.float32), np.array(ys, dtype=np.int64)
