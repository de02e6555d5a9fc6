"""dctr_col_wsum_static: the column reduction on the library's built-in scratch array, for callers that cannot allocate.
Called through the C-ABI (ctypes) and compared with dctr_col_wsum (caller-provided workspace) on two calls in a row with
different inputs -- the array keeps the previous call's contents, so the second call is the one that proves the clearing.
Same kernel, same summation order: the results are required to be bit-identical (the tolerance of the other col_wsum tests,
1e-5 relative, is asserted as well).  Collected last on purpose."""
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


def _both(L, A, lda, N, H, wv):
    f32 = dict(dtype=torch.float32, device="cuda")
    want, got = torch.full((H,), 7.0, **f32), torch.full((H,), -7.0, **f32)
    ws = torch.zeros(L.lib().dctr_col_wsum_workspace(N, H), dtype=torch.uint8, device="cuda")
    L.call("dctr_col_wsum", A.data_ptr(), lda, N, H, wv.data_ptr() if wv is not None else None, want.data_ptr(), ws.data_ptr(), ws.numel(), None)
    L.call("dctr_col_wsum_static", A.data_ptr(), lda, N, H, wv.data_ptr() if wv is not None else None, got.data_ptr(), None)
    torch.cuda.synchronize()
    return got, want


def test_col_wsum_static_matches_col_wsum_twice_in_a_row():
    from deepctr_b200 import _lib as L
    g = torch.Generator(device="cuda").manual_seed(11)
    f32 = dict(dtype=torch.float32, device="cuda")
    assert L.lib().dctr_col_wsum_static_max_h() == 1024
    # two calls in a row on different shapes (wide then narrow: the second one runs over what the first left in the scratch),
    # then the first shape again with new data, then the widest matrix the built-in scratch takes, unweighted and not 16-byte aligned
    cases = [(4096, 300, True, 0), (1000, 7, True, 0), (4096, 300, True, 0), (65536, 100, False, 0), (513, 1024, True, 0), (777, 33, False, 1)]
    for N, H, weighted, off in cases:
        lda = r4(H) + 4
        A = torch.randn(N * lda + 8, generator=g, **f32)
        Av = A[off:off + N * lda].view(N, lda)
        wv = torch.randn(N, generator=g, **f32) if weighted else None
        got, want = _both(L, Av, lda, N, H, wv)
        ref = ((Av[:, :H].double() * wv.double()[:, None]) if weighted else Av[:, :H].double()).sum(0)
        scale = ((Av[:, :H].double().abs() * (wv.double().abs()[:, None] if weighted else 1.0)).sum(0)).cpu().numpy()
        assert np.all(np.abs(want.double().cpu().numpy() - ref.cpu().numpy()) <= 1e-5 * scale + 1e-6)     # the existing entry point is right ...
        assert np.all(np.abs(got.double().cpu().numpy() - want.double().cpu().numpy()) <= 1e-5 * np.abs(want.double().cpu().numpy()) + 1e-6)
        assert torch.equal(got, want), (N, H)                                                        # ... and the new one is bit-identical


def test_col_wsum_static_rejects_what_does_not_fit():
    from deepctr_b200 import _lib as L
    f32 = dict(dtype=torch.float32, device="cuda")
    A = torch.zeros(8, 1028, **f32)
    out = torch.zeros(1028, **f32)
    with pytest.raises(Exception) as ei:
        L.call("dctr_col_wsum_static", A.data_ptr(), 1028, 8, 1025, None, out.data_ptr(), None)
    assert "1024" in str(ei.value) and "col_wsum_static" in str(ei.value)
    # the failed call left the entry point usable
    got, want = _both(L, A[:, :64].contiguous(), 64, 8, 64, None)
    assert torch.equal(got, want)
