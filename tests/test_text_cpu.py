"""CPU tests of the text-ingestion row: the reference parsers restated in oracle/text_oracle.py against the
fixture writers of data.py, and the device parser's algorithm (tests/text_emulation.py, a lane-by-lane restatement of
csrc/text.cu) against those parsers -- bit-exact."""
import numpy as np

import text_emulation as emu
import text_oracle as T
from deepctr_b200 import data as D


def test_reference_parsers_round_trip_the_writers(tmp_path):
    rng = np.random.default_rng(0)
    ids = rng.integers(0, 5000, (40, 16))
    y = rng.integers(0, 2, 40)
    p = str(tmp_path / "train.fm.txt")
    D.write_fm_samples(p, ids, y)
    lines = T.nonblank_lines(open(p).read())
    f, w, yy, nnz = T.sparse_block(lines, 16, -1, 0)
    assert np.array_equal(f, ids) and np.array_equal(yy, y) and (nnz == 16).all() and (w == 1).all()
    hf, hy = D.load_fm_samples(p)
    assert np.array_equal(hf, f) and np.array_equal(hy, yy)
    p2 = str(tmp_path / "train.yzx.txt")
    ragged = ids.copy()
    ragged[::3, 10:] = -1
    D.write_yzx(p2, ragged, y, z=123)
    lines = T.nonblank_lines(open(p2).read())
    f, w, yy, nnz = T.sparse_block(lines, 16, 4999, 1)
    xi, xv, xy = D.load_yzx(lines, 4999, 16)
    assert np.array_equal(f, xi) and np.array_equal(w, xv) and np.array_equal(yy, xy)
    assert np.array_equal(nnz, (ragged >= 0).sum(1))


def test_device_parser_rules_equal_reference_parsers():
    rng = np.random.default_rng(1)
    lines = []
    for i in range(150):
        k = 200 if i == 40 else int(rng.integers(1, 17))
        sep = "  " if i % 7 == 3 else " "
        lines.append("%d%s%s" % (rng.integers(0, 2), sep, sep.join("%d:1" % j for j in rng.integers(0, 2 ** 31 - 1, k))))
        if i % 9 == 0:
            lines.append("")
    for eol, tail in (("\n", "\n"), ("\r\n", ""), ("\n", "\n\n")):
        text = eol.join(lines) + tail
        t = text.encode()
        ref = T.nonblank_lines(text)
        ref = [ln for ln in ref if ln.strip() != ""]
        ls, total = emu.index_lines(t, len(lines))
        kept = [i for i in range(total) if t[ls[i]:ls[i + 1]].strip() != b""]
        assert len(kept) == len(ref) == 150
        f, w, y, st = emu.parse(t, ls, total, len(lines), False, 0, 16, -1)
        of, ow, oy, onnz = T.sparse_block(ref, 16, -1, 0)
        assert np.array_equal(f[kept], of) and np.array_equal(w[kept], ow) and np.array_equal(y[kept], oy)
        assert np.array_equal(st[kept], onnz) and (np.delete(st, kept) == -1).all()
    # yzx: one skipped token, (X_dim - 1, 0) padding; capacity below the line count keeps the head of the file
    lines = ["%d %d %s" % (rng.integers(0, 2), rng.integers(0, 300), " ".join("%d:1" % j for j in rng.integers(0, 10 ** 6, rng.integers(0, 20))))
             for _ in range(70)]
    t = ("\n".join(lines) + "\n").encode()
    ls, total = emu.index_lines(t, 70)
    f, w, y, st = emu.parse(t, ls, total, 70, False, 1, 12, 10 ** 6)
    of, ow, oy, onnz = T.sparse_block(lines, 12, 10 ** 6, 1)
    assert np.array_equal(f, of) and np.array_equal(w, ow) and np.array_equal(y, oy) and np.array_equal(st, onnz)
    ls, total = emu.index_lines(t, 10)
    assert total == 70 and np.array_equal(emu.parse(t, ls, total, 10, False, 1, 12, 10 ** 6)[0], of[:10])


def test_device_criteo_rules_equal_get_fxy():
    rng = np.random.default_rng(2)
    offsets = D.offsets_of(D.CRITEO_CAT_SIZES, first=13)
    lines = []
    for i in range(40):
        num = [int(rng.integers(0, m + 1)) for m in D.CRITEO_MAX_VALS]
        cat = [int(rng.integers(0, s)) for s in D.CRITEO_CAT_SIZES]
        lines.append("%d\t%s\t%s" % (rng.integers(0, 2), "\t".join(map(str, num)), "\t".join(map(str, cat))))
    t = ("\n".join(lines) + "\n").encode()
    ls, total = emu.index_lines(t, 40)
    ids, w, y, st = emu.parse(t, ls, total, 40, True, n_num=13, n_cat=26, offsets=offsets, max_vals=D.CRITEO_MAX_VALS.astype(float))
    of, ox, oy = T.criteo_block(lines, offsets, D.CRITEO_MAX_VALS.tolist())
    assert np.array_equal(ids, of) and np.array_equal(w, ox) and np.array_equal(y, oy) and (st == 40).all()
    for i in range(3):              # the same triple the host loader of data.py produces
        hy, hf, hx = D.parse_criteo_line(lines[i], offsets)
        assert hy == oy[i] and hf == of[i].tolist() and np.allclose(hx, ox[i], rtol=1e-7)
