"""Times dctr_mlp2tc_train alone (CUDA events, L2 flushed between launches) at the headline shape.  Usage: m2tc_time.py [N]"""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from deepctr_b200 import _lib as L  # noqa: E402
import deepctr_b200.ops  # noqa: E402,F401

O = torch.ops.deepctr
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
D0, H1, H2 = 177, 300, 100
f32 = dict(dtype=torch.float32, device="cuda")
g = torch.Generator(device="cuda").manual_seed(1)
x0 = torch.randn(N, 180, generator=g, **f32)
x0[:, D0:] = 0
W1, W2 = torch.randn(D0, 300, generator=g, **f32) * 0.07, torch.randn(H1, 100, generator=g, **f32) * 0.06
b1, b2, w3, b3 = torch.zeros(H1, **f32), torch.zeros(H2, **f32), torch.randn(H2, generator=g, **f32) * 0.1, torch.zeros(1, **f32)
m1, m2 = (torch.rand(1, 300, generator=g, **f32) < 0.5).float(), (torch.rand(1, 100, generator=g, **f32) < 0.5).float()
y = (torch.rand(N, generator=g, **f32) < 0.3).float()
packed = torch.zeros(L.lib().dctr_mlp2tc_packed_bytes(D0, H1, H2), dtype=torch.uint8, device="cuda")
ws = torch.zeros(L.lib().dctr_mlp2tc_workspace(N, H2), dtype=torch.uint8, device="cuda")
a1, dz1, dz2, dz0 = torch.zeros(N, 300, **f32), torch.zeros(N, 304, **f32), torch.zeros(N, 304, **f32), torch.zeros(N, 180, **f32)
sc = [torch.zeros(N, **f32) for _ in range(4)]
loss, gw3, gb3 = torch.zeros(1, **f32), torch.zeros(H2, **f32), torch.zeros(1, **f32)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
x = x0.clone()


def run():
    O.mlp2tc_train(D0, H1, H2, x, None, 0, packed, b1, 1, m1, b2, 1, m2, w3, b3, y, 1.0, 1.0, a1, dz1, dz2, dz0, sc[0], sc[1], sc[2], sc[3],
                   loss, gw3, gb3, False, 0.0, 1.0, ws)


for cold in (False, True):
    ts = []
    for i in range(25):
        if cold:
            flush.zero_()
        O.mlp2tc_pack(D0, H1, H2, W1, W2, packed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts = sorted(ts[5:])
    print("N=%d %s: median %.1f us, min %.1f us" % (N, "after an L2 flush + repack" if cold else "warm", ts[len(ts) // 2], ts[0]))
