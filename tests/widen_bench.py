"""Device timings of the SURVEY 8(f) widening kernels (text ingestion, embedding bag, FTRL sweep, SNN step) -- CUDA events
on the launching stream, 3 warm-ups, L2 flushed before every timed launch.  Prints one JSON object per line and writes them
to gpurun_out/r01_widen_bench.json.  Usage: python tests/widen_bench.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepctr_b200 import _lib as L   # noqa: E402
from deepctr_b200 import data as D   # noqa: E402

PEAK = 6457.4
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
_flush = None


ONCE = bool(os.environ.get("WIDEN_ONCE"))     # one launch of every kernel, no end-to-end loop: the run ncu attaches to


def timed(fn, reps=5, warm=3):
    global _flush
    if ONCE:
        reps, warm = 1, 0
    if _flush is None:
        _flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        _flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def line(out, name, ms, nbytes, **kw):
    r = dict(kernel=name, ms=round(ms, 4), algorithmic_MB=round(nbytes / 1e6, 2), GBps=round(nbytes / ms / 1e6, 1),
             frac_of_hbm_peak=round(nbytes / ms / 1e6 / PEAK, 3), **kw)
    print(json.dumps(r), flush=True)
    out.append(r)


def main():
    from deepctr_b200.ingest import DeviceText
    out = []
    rng = np.random.default_rng(0)
    # ---- text: ~64 MB of train.fm.txt-shaped lines (16 id:1 pairs), ~96 MB of Criteo TSV
    ids = D.make_field_batch(rng, 8192, D.ipinyou_field_sizes(1_000_000))
    block = "".join("%d %s\n" % (i & 1, " ".join("%d:1" % j for j in r)) for i, r in enumerate(ids)).encode()
    text = block * (64 * (1 << 20) // len(block))
    dt = DeviceText(text)
    n, nb = dt.n_lines, len(text)
    ms = timed(lambda: dt._index(dt.line_start, dt.max_lines))
    line(out, "text_index_lines (count + scan + fill)", ms, 2 * nb + 8 * n, lines=n, note="text read twice")
    ms = timed(lambda: dt.parse_sparse(16, 0, -1))
    line(out, "text_parse_kernel<sparse> J=16", ms, nb + n * (8 + 16 * 8 + 8), lines=n, lines_per_s=round(n / ms * 1e3))
    del dt
    offsets = D.offsets_of(D.CRITEO_CAT_SIZES, first=13)
    cols = [rng.integers(0, int(m) + 1, 4096) for m in D.CRITEO_MAX_VALS] + [rng.integers(0, int(s), 4096) for s in D.CRITEO_CAT_SIZES]
    block = "".join("%d\t%s\n" % (i & 1, "\t".join(str(int(c[i])) for c in cols)) for i in range(4096)).encode()
    text = block * (96 * (1 << 20) // len(block))
    dt = DeviceText(text)
    n, nb = dt.n_lines, len(text)
    ms = timed(lambda: dt.parse_criteo(offsets, D.CRITEO_MAX_VALS))
    line(out, "text_parse_kernel<criteo> 13+26", ms, nb + n * (8 + 39 * 8 + 8), lines=n, lines_per_s=round(n / ms * 1e3),
         note="includes two small H2D copies and a stream sync per call")
    del dt
    # ---- embedding bag: SNN shape, table >> L2
    N, F, H, X = 65536, 16, 200, 1_000_000
    table = torch.randn(X, H, device="cuda") * 0.1
    bids = torch.randint(0, X, (N, F), dtype=torch.int32, device="cuda")
    b0 = torch.zeros(H, device="cuda")
    o = torch.empty(N, H, device="cuda")
    ms = timed(lambda: L.call("dctr_bag_fwd", bids.data_ptr(), None, table.data_ptr(), X, H, N, F, H, b0.data_ptr(), 3, o.data_ptr(), H, None))
    line(out, "bag_fwd_kernel N=65536 F=16 H=200 X=1e6", ms, N * (F * 4 + F * H * 4 + H * 4))
    # its update: sort + SUM fold with a shared gradient row per sample (13 column slabs of 16)
    M = N * F
    rows, pos = torch.empty(M, dtype=torch.int32, device="cuda"), torch.empty(M, dtype=torch.int32, device="cuda")
    sws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    fws = torch.zeros(L.lib().dctr_scatter_workspace(M, H), dtype=torch.uint8, device="cuda")
    L.call("dctr_sort_rows", bids.data_ptr(), M, X, rows.data_ptr(), pos.data_ptr(), sws.data_ptr(), sws.numel(), None)
    g = torch.randn(N, H, device="cuda") * 1e-3
    ms = timed(lambda: L.call("dctr_scatter_sgd", 0, rows.data_ptr(), pos.data_ptr(), M, F, H, g.data_ptr(), H, 0, None, 0.01, 1.0,
                              table.data_ptr(), X, H, fws.data_ptr(), fws.numel(), None))
    line(out, "fold (bag update) M=1M d=200", ms, M * (8 + H * 4 + 2 * H * 4), launches=26)
    del table, g, o
    # ---- FTRL sweep over a Criteo-sized table (1.64 M rows x 12 floats)
    n4 = 1_636_608 * 12
    p, gg, nn, zz = (torch.full((n4,), v, device="cuda") for v in (0.01, 0.001, 0.1, 0.0))
    ms = timed(lambda: L.call("dctr_dense_ftrl", p.data_ptr(), gg.data_ptr(), nn.data_ptr(), zz.data_ptr(), n4, 0.01, -0.5, 0.0, 0.0, 0.0, 1, None))
    line(out, "dense_ftrl_kernel 19.6 M floats", ms, n4 * 4 * 8, note="4 reads + 4 writes per element")
    del p, gg, nn, zz
    if ONCE:
        return
    # ---- SNN fine-tuning step, reference shape (x_dim 133465, 200-300-100-1, batch 1000), graph replay, host feeds
    from deepctr_b200.snn import SNN
    import time
    m = SNN(batch_size=1000, max_feats=16)
    m.engine.set_layer(2, rng.uniform(-0.1, 0.1, (100, 1)), np.zeros(1))
    feats = rng.integers(0, 133465, (1000, 16))
    y = rng.integers(0, 2, 1000)
    for _ in range(5):
        m.train(feats, y)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        m.train(feats, y)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 50 * 1e3
    r = dict(kernel="SNN.train end to end (host ids in, pred | loss out)", ms=round(ms, 4), samples_per_s=round(1000 / ms * 1e3))
    print(json.dumps(r), flush=True)
    out.append(r)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "r01_widen_bench.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
