"""csrc/mlp2tc.cu (the two-hidden-layer step chained through tcgen05 inside one kernel) through torch.ops.deepctr against the
float64 restatement of tests/test_mlp2_gpu.py: batches that are not multiples of the 128-sample tile, widths that are not
multiples of 4 / 16 / 32 (edge columns of the MMA groups, zero-padded weight tiles), every activation, broadcast and per-sample
masks, gradient output and the fused SGD of the output unit.  Tolerance: 2e-5 (3xTF32 with one accumulator, ex2 / rcp
activations), relative to |value| + 0.1 max|tensor| like the other GEMM-path tests."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from test_mlp2_gpu import ACT, r4, reference  # noqa: E402


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


CASES = [
    # N, D0, H1, H2, acts, masks (0 none / 1 broadcast row / 2 per sample), sgd
    (4096, 177, 300, 100, ("none", "tanh", "tanh"), (0, 1, 1), 1),
    (100, 177, 300, 100, ("none", "tanh", "tanh"), (0, 1, 1), 0),
    (77, 45, 30, 7, ("tanh", "relu", "sigmoid"), (2, 2, 2), 0),
    (33, 9, 5, 3, ("sigmoid", "sigmoid", "tanh"), (2, 0, 1), 1),
    (1, 192, 320, 128, ("relu", "tanh", "relu"), (0, 0, 0), 0),
    (257, 130, 257, 64, ("none", "relu", "tanh"), (1, 2, 0), 1),
    (1000, 161, 161, 97, ("sigmoid", "tanh", "tanh"), (0, 1, 1), 1),
    (129, 32, 160, 33, ("none", "tanh", "none"), (0, 0, 2), 0),
]


def split_ref(w):
    """hi = tf32 round-to-nearest (ties away) of w, lo = the same of w - hi, on the bit pattern like the kernel."""
    b = w.view(torch.int32)
    hi = ((b + 0x1000) & ~0x1FFF).view(torch.float32)
    r = (w - hi).view(torch.int32)
    return hi, ((r + 0x1000) & ~0x1FFF).view(torch.float32)


@pytest.mark.parametrize("N,D0,H1,H2,acts,mk,sgd", CASES)
def test_mlp2tc_step(N, D0, H1, H2, acts, mk, sgd):
    from deepctr_b200 import _lib as L
    import deepctr_b200.ops  # noqa: F401
    O = torch.ops.deepctr
    assert L.lib().dctr_mlp2tc_supported(D0, H1, H2)
    rng = np.random.default_rng(N + D0)
    z0 = rng.standard_normal((N, D0)) * 0.5
    x = ACT[acts[0]][0](z0)
    Ws = [rng.uniform(-1, 1, (D0, H1)) / np.sqrt(D0), rng.uniform(-1, 1, (H1, H2)) / np.sqrt(H1), rng.uniform(-1, 1, (H2, 1)) / np.sqrt(H2)]
    bs = [rng.uniform(-0.1, 0.1, H1), rng.uniform(-0.1, 0.1, H2), rng.uniform(-0.1, 0.1, 1)]
    y = rng.integers(0, 2, N).astype(np.float64)
    ms = []
    for k, w in zip(mk, (D0, H1, H2)):
        ms.append(np.ones((1, w)) if k == 0 else rng.binomial(1, 0.5, (1 if k == 1 else N, w)) * (2.0 if k == 2 else 1.0))
    gscale, lr, dout = 0.7, 0.05, 0.99
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
    # the float64 reference works on exactly the float32 inputs the kernel sees
    x64 = xd[:, :D0].double().cpu().numpy()
    Ws64 = [W1[:, :H1].double().cpu().numpy(), W2[:, :H2].double().cpu().numpy(), w3.double().cpu().numpy().reshape(-1, 1)]
    bs64 = [b1.double().cpu().numpy(), b2.double().cpu().numpy(), b3.double().cpu().numpy()]
    loss_n, p, grads, dz0 = reference(x64, ms, Ws64, bs64, y, acts, gscale)
    packed = torch.full((L.lib().dctr_mlp2tc_packed_bytes(D0, H1, H2),), 0xA5, dtype=torch.uint8, device="cuda")
    O.mlp2tc_pack(D0, H1, H2, W1, W2, packed)
    torch.cuda.synchronize()
    # every byte of the packed image is written, and it holds each matrix's hi + lo parts exactly twice (zeros elsewhere)
    pf = packed.view(torch.float32)
    assert bool(torch.isfinite(pf).all())
    for w in (W1[:, :H1], W2[:, :H2]):
        h, l = split_ref(w.contiguous())
        assert float((w - h - l).abs().max()) <= float(w.abs().max()) * 2.0 ** -21
    tot = float(W1[:, :H1].double().sum() + W2[:, :H2].double().sum())
    assert abs(float(pf.double().sum()) - 2 * tot) <= 1e-6 * float(W1.abs().sum() + W2.abs().sum())
    a1, dz1, dz2 = torch.full((N, ld1), 7.0, **f32), torch.full((N, ldd), 7.0, **f32), torch.full((N, ldd), 7.0, **f32)
    dz0d = torch.full((N, ld0), 7.0, **f32)
    logit, pred, lossn, dlogit = (torch.zeros(N, **f32) for _ in range(4))
    loss = torch.zeros(1, **f32)
    gw3, gb3 = torch.zeros(H2, **f32), torch.zeros(1, **f32)
    ws = torch.zeros(L.lib().dctr_mlp2tc_workspace(N, H2), dtype=torch.uint8, device="cuda")
    A = L.ACT
    for rep in range(2 if not sgd else 1):        # twice without SGD: the workspace (ticket) is left reusable
        if rep:
            xd.copy_(padded(x, ld0))
        O.mlp2tc_train(D0, H1, H2, xd, md[0], A[acts[0]], packed, b1, A[acts[1]], md[1], b2, A[acts[2]], md[2], w3, b3, yd,
                       gscale, 0.5, a1, dz1, dz2, dz0d, logit, pred, lossn, dlogit, loss, gw3, gb3, bool(sgd), lr, dout, ws)
    torch.cuda.synchronize()

    def close(got, want, rel=2e-5):
        got = got.cpu().numpy() if hasattr(got, "cpu") else got
        got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
        tol = rel * (np.abs(want) + max(1e-3, float(np.abs(want).max()) * 0.1))
        assert np.all(np.abs(got - want) <= tol), float(np.max(np.abs(got - want) / tol))
    m0, m1, m2 = ms
    a0r = x64 * m0
    a1r = ACT[acts[1]][0](a0r @ Ws64[0] + bs64[0]) * m1
    close(xd[:, :D0], a0r)                                   # x *= m0 in place
    close(a1[:, :H1], a1r)
    assert not a1[:, H1:].any()
    close(pred, p)
    close(lossn, loss_n)
    close(loss, [0.5 * loss_n.sum()])
    close(dlogit, (p - y) * gscale)
    close(dz0d[:, :D0], dz0)
    assert not dz0d[:, D0:].any()
    gW1r, gb1r, gW2r, gb2r, gw3r, gb3r = grads
    # dz1 / dz2 are what the weight gradients read: check them through the gradients they imply
    close((a0r.T @ dz1[:, :H1].double().cpu().numpy()), gW1r)
    close((a1r.T @ dz2[:, :H2].double().cpu().numpy()), gW2r)
    close(dz1[:, :H1].double().sum(0), gb1r)
    close(dz2[:, :H2].double().sum(0), gb2r)
    assert not dz1[:, H1:r4(H1)].any() and not dz2[:, H2:r4(H2)].any()
    if sgd:
        close(w3, Ws64[2].reshape(-1) * dout - lr * gw3r)
        close(b3, bs64[2] * dout - lr * gb3r)
    else:
        close(gw3, gw3r)
        close(gb3, [gb3r])


def test_mlp2tc_is_deterministic():
    """two runs on the same inputs are bit-identical (fixed reduction orders, no float atomics)."""
    from deepctr_b200 import _lib as L
    import deepctr_b200.ops  # noqa: F401
    O = torch.ops.deepctr
    N, D0, H1, H2 = 1000, 177, 300, 100
    g = torch.Generator(device="cuda").manual_seed(5)
    f32 = dict(dtype=torch.float32, device="cuda")
    x0 = torch.randn(N, r4(D0), generator=g, **f32)
    x0[:, D0:] = 0
    W1, W2 = torch.randn(D0, r4(H1), generator=g, **f32) * 0.07, torch.randn(H1, r4(H2), generator=g, **f32) * 0.06
    b1, b2, w3, b3 = torch.zeros(H1, **f32), torch.zeros(H2, **f32), torch.randn(H2, generator=g, **f32) * 0.1, torch.zeros(1, **f32)
    y = (torch.rand(N, generator=g, **f32) < 0.3).float()
    packed = torch.zeros(L.lib().dctr_mlp2tc_packed_bytes(D0, H1, H2), dtype=torch.uint8, device="cuda")
    O.mlp2tc_pack(D0, H1, H2, W1, W2, packed)
    ws = torch.zeros(L.lib().dctr_mlp2tc_workspace(N, H2), dtype=torch.uint8, device="cuda")
    outs = []
    for _ in range(2):
        x = x0.clone()
        a1, dz1, dz2, dz0 = torch.zeros(N, r4(H1), **f32), torch.zeros(N, 304, **f32), torch.zeros(N, 304, **f32), torch.zeros(N, r4(D0), **f32)
        sc = [torch.zeros(N, **f32) for _ in range(4)]
        loss, gw3, gb3 = torch.zeros(1, **f32), torch.zeros(H2, **f32), torch.zeros(1, **f32)
        O.mlp2tc_train(D0, H1, H2, x, None, 0, packed, b1, 1, None, b2, 1, None, w3, b3, y, 1.0, 1.0,
                       a1, dz1, dz2, dz0, sc[0], sc[1], sc[2], sc[3], loss, gw3, gb3, False, 0.0, 1.0, ws)
        torch.cuda.synchronize()
        outs.append((a1, dz1, dz2, dz0, sc[1], loss, gw3, gb3))
    for u, v in zip(*outs):
        assert torch.equal(u, v)


def test_sgd_fused_reductions():
    """dctr_splitk_reduce_sgd / dctr_col_wsum_sgd: the fold of the split-K slabs and the bias column sums with the SGD step applied
    in place -- equal to the gradient-producing entry point followed by dctr_dense_sgd, bit for bit."""
    import deepctr_b200.ops  # noqa: F401
    from deepctr_b200 import _lib as L
    O = torch.ops.deepctr
    g = torch.Generator(device="cuda").manual_seed(9)
    f32 = dict(dtype=torch.float32, device="cuda")
    for splits, M, N in ((8, 177, 300), (3, 5, 7), (1, 64, 64)):
        ld = r4(N)
        slabs = torch.randn(splits, M, ld, generator=g, **f32)
        W = torch.randn(M, ld, generator=g, **f32)
        gW, W2 = torch.zeros(M, ld, **f32), W.clone()
        O.splitk_reduce(slabs, splits, M, N, ld, gW)
        O.dense_sgd(W2.view(-1), gW.view(-1), M * ld, 0.05, 0.999)
        W1 = W.clone()
        O.splitk_reduce_sgd(slabs, splits, M, N, ld, W1, 0.05, 0.999)
        torch.cuda.synchronize()
        assert torch.equal(W1[:, :N], W2[:, :N]) and torch.equal(W1[:, N:], W[:, N:])
    for N, H in ((4096, 300), (100, 7), (65536, 100)):
        A = torch.randn(N, r4(H) + 4, generator=g, **f32)
        b = torch.randn(H, generator=g, **f32)
        ws = torch.zeros(L.lib().dctr_col_wsum_workspace(N, H), dtype=torch.uint8, device="cuda")
        gb, b2 = torch.zeros(H, **f32), b.clone()
        O.col_wsum(A, H, None, gb, ws)
        O.dense_sgd(b2, gb, H, 0.05, 0.999)
        b1 = b.clone()
        O.col_wsum_sgd(A, H, None, b1, 0.05, 0.999, ws)
        torch.cuda.synchronize()
        assert torch.equal(b1, b2)
