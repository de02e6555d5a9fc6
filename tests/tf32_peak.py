"""Calibration: what cuBLAS reaches with TF32 / bf16 tensor cores on this box (8192^3), to place the 3xTF32 ceiling."""
import torch
a = torch.randn(8192, 8192, device="cuda"); b = torch.randn(8192, 8192, device="cuda")
for name, fn in [("tf32", lambda: torch.matmul(a, b)), ("bf16", None), ("fp32", None)]:
    if name == "tf32":
        torch.backends.cuda.matmul.allow_tf32 = True
    elif name == "bf16":
        x, y = a.bfloat16(), b.bfloat16(); fn = lambda: torch.matmul(x, y)
    else:
        torch.backends.cuda.matmul.allow_tf32 = False; fn = lambda: torch.matmul(a, b)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print("%s cuBLAS 8192^3: %.3f ms  %.1f TFLOP/s" % (name, ms, 2 * 8192 ** 3 / ms / 1e9), flush=True)
