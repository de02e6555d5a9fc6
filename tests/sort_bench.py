"""Timing of dctr_sort_rows over key distributions (GPU box): python tests/sort_bench.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepctr_b200 import _lib as L, data   # noqa: E402


def run(name, ids, X):
    M = ids.size
    d = torch.as_tensor(ids.astype(np.int32).reshape(-1)).cuda()
    rows, pos = torch.empty(M, dtype=torch.int32, device="cuda"), torch.empty(M, dtype=torch.int32, device="cuda")
    ws = torch.empty(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        L.call("dctr_sort_rows", L.ptr(d), M, X, L.ptr(rows), L.ptr(pos), L.ptr(ws), ws.numel(), st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        L.call("dctr_sort_rows", L.ptr(d), M, X, L.ptr(rows), L.ptr(pos), L.ptr(ws), ws.numel(), st)
    e1.record()
    torch.cuda.synchronize()
    ok = bool((rows[1:] >= rows[:-1]).all())
    print("%-28s M=%8d X=%9d  %.1f us  sorted=%s" % (name, M, X, e0.elapsed_time(e1) * 100, ok), flush=True)


def main():
    rng = np.random.default_rng(0)
    N = 65536
    sizes = data.ipinyou_field_sizes(1_000_000)
    run("ipinyou zipf N=65536", data.make_field_batch(rng, N, sizes, 1.1), 1_000_000)
    run("uniform M=1M X=1M", rng.integers(0, 1_000_000, N * 16), 1_000_000)
    run("constant M=1M", np.full(N * 16, 12345), 1_000_000)
    run("16 values M=1M", rng.integers(0, 16, N * 16) * 60000, 1_000_000)
    run("ipinyou zipf N=4096", data.make_field_batch(rng, 4096, sizes, 1.1), 1_000_000)
    run("uniform M=16M X=20M", rng.integers(0, 20_000_000, 1 << 24), 20_000_000)


if __name__ == "__main__":
    main()
