"""csrc/mlp2.cu through the C-ABI against a float64 restatement of the same two-hidden-layer step (FNN_wnzh.py:147-182 /
FNN.py:86-94 and their autodiff): odd shapes (batch not a multiple of the 32-sample tile, widths that are not multiples of
4 or 8), every activation, broadcast and per-sample masks, gradient output and fused SGD.  Tolerance 1e-5 (fp32 FFMA)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def r4(n):
    return (n + 3) // 4 * 4


ACT = {"none": (lambda x: x, lambda t: np.ones_like(t)), "tanh": (np.tanh, lambda t: 1 - t * t),
       "relu": (lambda x: np.maximum(x, 0), lambda t: (t > 0).astype(np.float64)),
       "sigmoid": (lambda x: 1 / (1 + np.exp(-x)), lambda t: t * (1 - t))}


def reference(x, ms, Ws, bs, y, acts, gscale):
    """x = act0(z) (unmasked).  Returns loss_n, pred, grads of W1 b1 W2 b2 w3 b3, dz0."""
    m0, m1, m2 = ms
    a0 = x * m0
    a1 = ACT[acts[1]][0](a0 @ Ws[0] + bs[0]) * m1
    a2 = ACT[acts[2]][0](a1 @ Ws[1] + bs[1]) * m2
    z = a2 @ Ws[2].reshape(-1) + bs[2][0]
    p = 1 / (1 + np.exp(-z))
    loss_n = np.maximum(z, 0) - z * y + np.log1p(np.exp(-np.abs(z)))
    dl = (p - y) * gscale

    def dact(act, a, m):
        safe = np.where(m != 0, m, 1.0)
        return np.where(m != 0, m * ACT[act][1](a / safe), 0.0)
    gw3, gb3 = a2.T @ dl, dl.sum()
    dz2 = np.outer(dl, Ws[2].reshape(-1)) * dact(acts[2], a2, np.broadcast_to(m2, a2.shape))
    gW2, gb2 = a1.T @ dz2, dz2.sum(0)
    dz1 = (dz2 @ Ws[1].T) * dact(acts[1], a1, np.broadcast_to(m1, a1.shape))
    gW1, gb1 = a0.T @ dz1, dz1.sum(0)
    dz0 = (dz1 @ Ws[0].T) * dact(acts[0], a0, np.broadcast_to(m0, a0.shape))
    return loss_n, p, (gW1, gb1, gW2, gb2, gw3, gb3), dz0


CASES = [
    # N, D0, H1, H2, acts, masks (0 none / 1 broadcast row / 2 per sample), sgd
    (4096, 177, 300, 100, ("none", "tanh", "tanh"), (0, 1, 1), 1),
    (100, 177, 300, 100, ("none", "tanh", "tanh"), (0, 1, 1), 0),
    (77, 45, 30, 7, ("tanh", "relu", "sigmoid"), (2, 2, 2), 0),
    (33, 9, 5, 3, ("sigmoid", "sigmoid", "tanh"), (2, 0, 1), 1),
    (1, 300, 384, 129, ("relu", "tanh", "relu"), (0, 0, 0), 0),
    (257, 130, 257, 64, ("none", "relu", "tanh"), (1, 2, 0), 1),
    (1000, 201, 300, 100, ("sigmoid", "tanh", "tanh"), (0, 1, 1), 1),
]


@pytest.mark.parametrize("N,D0,H1,H2,acts,mk,sgd", CASES)
def test_mlp2_step(N, D0, H1, H2, acts, mk, sgd):
    from deepctr_b200 import _lib as L
    assert L.lib().dctr_mlp2_supported(D0, H1, H2)
    rng = np.random.default_rng(N + D0)
    z0 = rng.standard_normal((N, D0)) * 0.5
    x = ACT[acts[0]][0](z0)
    Ws = [rng.uniform(-1, 1, (D0, H1)) / np.sqrt(D0), rng.uniform(-1, 1, (H1, H2)) / np.sqrt(H1), rng.uniform(-1, 1, (H2, 1)) / np.sqrt(H2)]
    bs = [rng.uniform(-0.1, 0.1, H1), rng.uniform(-0.1, 0.1, H2), rng.uniform(-0.1, 0.1, 1)]
    y = rng.integers(0, 2, N).astype(np.float64)
    widths = (D0, H1, H2)
    ms = []
    for k, w in zip(mk, widths):
        if k == 0:
            ms.append(np.ones((1, w)))
        else:
            ms.append(rng.binomial(1, 0.5, (1 if k == 1 else N, w)) * (2.0 if k == 2 else 1.0))
    gscale, lr, dh, dout = 0.7, 0.05, 0.999, 0.99
    loss_n, p, grads, dz0 = reference(x, ms, Ws, bs, y, acts, gscale)
    f32 = dict(dtype=torch.float32, device="cuda")
    ld0, ld1, ld2, ldd = r4(D0), r4(H1), r4(H2), r4(max(H1, H2)) + 4

    def padded(a, ld):
        t = torch.zeros(a.shape[0], ld, **f32)
        t[:, :a.shape[1]] = torch.as_tensor(a, dtype=torch.float32).cuda()
        return t
    xd = padded(x, ld0)
    W1, W2 = padded(Ws[0], ld1), padded(Ws[1], ld2)
    b1, b2 = torch.as_tensor(bs[0], **f32), torch.as_tensor(bs[1], **f32)
    w3, b3 = torch.as_tensor(Ws[2].reshape(-1), **f32).contiguous(), torch.as_tensor(bs[2], **f32)
    yd = torch.as_tensor(y, **f32)
    md = [None if k == 0 else padded(m, ld) for k, m, ld in zip(mk, ms, (ld0, ld1, ld2))]
    mp = [(0, 0) if m is None else (m.data_ptr(), 0 if m.shape[0] == 1 else m.stride(0)) for m in md]
    a1, dz1, dz2 = torch.zeros(N, ld1, **f32), torch.full((N, ldd), 7.0, **f32), torch.full((N, ldd), 7.0, **f32)
    dz0d = torch.zeros(N, ld0, **f32)
    logit, pred, lossn, dlogit = (torch.zeros(N, **f32) for _ in range(4))
    loss = torch.zeros(1, **f32)
    gW1, gW2 = torch.zeros(D0, ld1, **f32), torch.zeros(H1, ld2, **f32)
    gb1, gb2, gw3, gb3 = torch.zeros(H1, **f32), torch.zeros(H2, **f32), torch.zeros(H2, **f32), torch.zeros(1, **f32)
    ws = torch.zeros(L.lib().dctr_mlp2_workspace(N, D0, H1, H2), dtype=torch.uint8, device="cuda")
    W1t, W2t = torch.full((H1, ld0), 9.0, **f32), torch.full((H2, ld1), 9.0, **f32)
    L.call("dctr_mlp2_transpose", D0, H1, H2, W1.data_ptr(), ld1, W2.data_ptr(), ld2, W1t.data_ptr(), ld0, W2t.data_ptr(), ld1, None)
    torch.cuda.synchronize()
    assert torch.equal(W1t[:, :D0], W1[:, :H1].t()) and torch.equal(W2t[:, :H1], W2[:, :H2].t())
    assert not W1t[:, D0:].any() and not W2t[:, H1:].any()
    A = L.ACT
    for rep in range(2 if not sgd else 1):        # twice without SGD: the workspace (tickets) is left reusable
        if rep:
            xd.copy_(padded(x, ld0))
        L.call("dctr_mlp2_train", N, D0, H1, H2, xd.data_ptr(), ld0, mp[0][0] or None, mp[0][1], A[acts[0]],
               W1.data_ptr(), ld1, b1.data_ptr(), A[acts[1]], mp[1][0] or None, mp[1][1],
               W2.data_ptr(), ld2, b2.data_ptr(), A[acts[2]], mp[2][0] or None, mp[2][1],
               W1t.data_ptr(), W2t.data_ptr(), w3.data_ptr(), b3.data_ptr(), yd.data_ptr(), gscale, 0.5,
               a1.data_ptr(), ld1, dz1.data_ptr(), dz2.data_ptr(), ldd, dz0d.data_ptr(),
               logit.data_ptr(), pred.data_ptr(), lossn.data_ptr(), dlogit.data_ptr(), loss.data_ptr(),
               gw3.data_ptr(), gb3.data_ptr(), sgd, lr, dout, ws.data_ptr(), ws.numel(), None)
        L.call("dctr_mlp2_wgrad", N, D0, H1, H2, xd.data_ptr(), ld0, a1.data_ptr(), ld1, dz1.data_ptr(), dz2.data_ptr(), ldd,
               W1.data_ptr(), ld1, b1.data_ptr(), W2.data_ptr(), ld2, b2.data_ptr(), gW1.data_ptr(), gb1.data_ptr(),
               gW2.data_ptr(), gb2.data_ptr(), sgd, lr, dh, ws.data_ptr(), ws.numel(), None)
    torch.cuda.synchronize()

    def close(got, want, scale=None):
        got, want = np.asarray(got.cpu().numpy(), np.float64), np.asarray(want, np.float64)
        tol = 1e-5 * (np.abs(want) + (scale if scale is not None else max(1e-3, float(np.abs(want).max()) * 0.1)))
        assert np.all(np.abs(got - want) <= tol), float(np.max(np.abs(got - want) / tol))
    close(pred, p)
    close(lossn, loss_n)
    close(loss, [0.5 * loss_n.sum()])
    close(dz0d[:, :D0], dz0)
    assert not dz0d[:, D0:].any()
    gW1r, gb1r, gW2r, gb2r, gw3r, gb3r = grads
    if sgd:
        close(W1[:, :H1], Ws[0] * dh - lr * gW1r)
        close(b1, bs[0] * dh - lr * gb1r)
        close(W2[:, :H2], Ws[1] * dh - lr * gW2r)
        close(b2, bs[1] * dh - lr * gb2r)
        close(w3, Ws[2].reshape(-1) * dout - lr * gw3r)
        close(b3, bs[2] * dout - lr * gb3r)
        assert not W1[:, H1:].any() and not W2[:, H2:].any()
    else:
        close(gW1[:, :H1], gW1r)
        close(gb1, gb1r)
        close(gW2[:, :H2], gW2r)
        close(gb2, gb2r)
        close(gw3, gw3r)
        close(gb3, [gb3r])
        close(W1[:, :H1], Ws[0], 0.0)            # untouched
