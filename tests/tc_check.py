"""Standalone check of dctr_gemm_tc against fp64 numpy (run under `timeout` on a GPU box; a wrong barrier hangs)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from deepctr_b200 import _lib as L

L.lib()


def r4(n):
    return (n + 3) // 4 * 4


def run(layout, M, N, K, splits=1, fn="dctr_gemm_tc"):
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    A = rng.normal(0, 1, (M, K)).astype(np.float32)
    B = rng.normal(0, 1, (K, N)).astype(np.float32)
    want = A.astype(np.float64) @ B.astype(np.float64)

    def pad(a):
        r, c = a.shape
        o = np.zeros((r, r4(c) + 4), np.float32)
        o[:, :c] = a
        return torch.as_tensor(o).cuda()
    Ad = pad(A if layout != "tn" else A.T.copy())
    Bd = pad(B if layout != "nt" else B.T.copy())
    ldc = r4(N) + 4
    C = torch.full((splits, M, ldc), 9.0, device="cuda")
    lay = {"nn": L.GEMM_NN, "nt": L.GEMM_NT, "tn": L.GEMM_TN}[layout]
    L.call(fn, lay, M, N, K, Ad.data_ptr(), Ad.stride(0), Bd.data_ptr(), Bd.stride(0), C.data_ptr(), ldc,
           L.EPI_NONE, 0, None, None, 0, None, 0, splits, None)
    torch.cuda.synchronize()
    got = C.sum(0).cpu().numpy()[:, :N].astype(np.float64) if splits > 1 else C[0].cpu().numpy()[:, :N].astype(np.float64)
    err = np.abs(got - want).max() / (np.abs(want).max() + 1e-30)
    rel = np.abs(got - want) / (np.abs(want) + 1e-3 * np.sqrt(K))
    print("%s %s M=%d N=%d K=%d splits=%d  max|err|/max|want| = %.3e  max rel = %.3e" % (fn, layout, M, N, K, splits, err, rel.max()), flush=True)
    return err


if __name__ == "__main__" and len(sys.argv) == 1:
    worst = 0.0
    for layout, M, N, K, s in [("nt", 128, 64, 32, 1), ("nt", 128, 128, 64, 1), ("nn", 128, 64, 32, 1), ("tn", 128, 64, 32, 1),
                               ("nn", 4096, 300, 177, 1), ("nt", 4096, 177, 300, 1), ("tn", 177, 300, 4096, 16),
                               ("nn", 100, 300, 177, 1), ("nt", 33, 7, 5, 1), ("tn", 257, 129, 1171, 3), ("nn", 1, 1, 1, 1)]:
        worst = max(worst, run(layout, M, N, K, s))
        run(layout, M, N, K, s, fn="dctr_gemm")
    print("worst", worst)
    sys.exit(0 if worst < 5e-6 else 1)


def bench(layout, M, N, K, splits, fn, reps=50):
    Ad = torch.randn(max(M, K), r4(max(M, K, N)) + 4, device="cuda")
    Bd = torch.randn(max(N, K), r4(max(M, K, N)) + 4, device="cuda")
    ldc = r4(N)
    C = torch.empty(splits, M, ldc, device="cuda")
    lay = {"nn": L.GEMM_NN, "nt": L.GEMM_NT, "tn": L.GEMM_TN}[layout]
    args = (lay, M, N, K, Ad.data_ptr(), Ad.stride(0), Bd.data_ptr(), Bd.stride(0), C.data_ptr(), ldc, L.EPI_NONE, 0, None, None, 0, None, 0, splits, None)
    for _ in range(5):
        L.call(fn, *args)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        L.call(fn, *args)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    print("%-13s %s %6dx%5dx%5d s=%2d  %8.2f us  %7.2f TFLOP/s" % (fn, layout, M, N, K, splits, us, 2.0 * M * N * K / us / 1e6), flush=True)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "bench":
    for shape in [("nn", 4096, 300, 177, 1), ("nn", 4096, 100, 300, 1), ("nt", 4096, 300, 100, 1), ("nt", 4096, 177, 300, 1),
                  ("tn", 177, 300, 4096, 16), ("tn", 300, 100, 4096, 16),
                  ("nn", 65536, 1000, 297, 1), ("nn", 65536, 800, 1000, 1), ("nt", 65536, 1000, 800, 1), ("tn", 1000, 800, 65536, 8),
                  ("nn", 8192, 8192, 8192, 1)]:
        for fn in ("dctr_gemm", "dctr_gemm_tc"):
            bench(*shape, fn)
