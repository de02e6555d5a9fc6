"""torch.ops.deepctr.* (deep-ctr_b200/ops.py): schema checks (torch.library.opcheck: the declared mutations are the real ones),
boundary validation (dtype / device / layout / alignment) and equality with the raw C-ABI call."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def _emb_case():
    from deepctr_b200 import _lib as L, ops  # noqa: F401
    g = torch.Generator(device="cuda").manual_seed(1)
    X, F, d, ld, N = 500, 6, 5, 8, 37
    table = torch.zeros(X, ld, device="cuda")
    table[:, :d].uniform_(-1, 1, generator=g)
    ids = torch.randint(0, X, (N, F), device="cuda", dtype=torch.int32, generator=g)
    ids[3] = ids[2]
    wts = torch.rand(N, F, device="cuda", generator=g)
    bias = torch.tensor([0.3], device="cuda")
    return X, F, d, ld, N, table, ids, wts, bias


def test_opcheck_schemas():
    from deepctr_b200 import _lib as L
    from torch.library import opcheck
    O = torch.ops.deepctr
    X, F, d, ld, N, table, ids, wts, bias = _emb_case()
    D0 = F * d + 1
    out = torch.zeros(N, 32, device="cuda")
    opcheck(O.gather_concat_fwd.default, (ids, wts, table, d, 0, False, bias, 1, out), test_utils=("test_schema",))
    M = N * F
    rows, pos = torch.zeros(M, dtype=torch.int32, device="cuda"), torch.zeros(M, dtype=torch.int32, device="cuda")
    ws = torch.zeros(L.lib().dctr_sort_rows_workspace(M), dtype=torch.uint8, device="cuda")
    opcheck(O.sort_pairs.default, (ids, None, M, X, rows, pos, ws), test_utils=("test_schema",))
    O.sort_pairs(ids, None, M, X, rows, pos, ws)
    fws = torch.zeros(L.lib().dctr_scatter_workspace(M, ld), dtype=torch.uint8, device="cuda")
    g = torch.randn(N, 32, device="cuda")
    opcheck(O.scatter_sgd.default, (0, rows, pos, M, F, d, g, d, wts, -1, 0.1, 1.0, table.clone(), fws), test_utils=("test_schema",))
    p, gr = torch.ones(64, device="cuda"), torch.ones(64, device="cuda")
    opcheck(O.dense_sgd.default, (p, gr, 64, 0.5, 1.0), test_utils=("test_schema",))
    opcheck(O.vec_sum.default, (gr, 64, 1.0, torch.zeros(1, device="cuda")), test_utils=("test_schema",))


def test_ops_equal_raw_abi_and_validate():
    from deepctr_b200 import _lib as L
    O = torch.ops.deepctr
    X, F, d, ld, N, table, ids, wts, bias = _emb_case()
    a, b = torch.zeros(N, 32, device="cuda"), torch.zeros(N, 32, device="cuda")
    O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, a)
    L.call("dctr_gather_concat_fwd", ids.data_ptr(), wts.data_ptr(), table.data_ptr(), X, ld, N, F, d, 0, 0, bias.data_ptr(), 1,
           b.data_ptr(), 32, None)
    torch.cuda.synchronize()
    assert torch.equal(a, b) and a.abs().sum() > 0
    # a column-offset view carries its own pointer and leading dimension
    big = torch.zeros(N, 40, device="cuda")
    O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, big[:, 4:36])
    assert torch.equal(big[:, 4:36], a) and not big[:, :4].any() and not big[:, 36:].any()
    with pytest.raises(TypeError):
        O.gather_concat_fwd(ids.long(), wts, table, d, 0, False, bias, 1, a)
    with pytest.raises(ValueError):
        O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, big[:, 1:33])        # not 16-byte aligned
    with pytest.raises(ValueError):
        O.gather_concat_fwd(ids.t().contiguous().t(), wts, table, d, 0, False, bias, 1, a)    # innermost dimension not contiguous
    with pytest.raises(NotImplementedError):
        O.gather_concat_fwd(ids.cpu(), wts.cpu(), table.cpu(), d, 0, False, bias.cpu(), 1, a.cpu())
    with pytest.raises(RuntimeError, match="gather_concat_fwd"):                        # the C-ABI's own EINVAL surfaces as RuntimeError
        O.gather_concat_fwd(ids, wts, table, d, 7, False, bias, 1, a)
    # the op launches on the CURRENT stream of the tensors' device
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        c = torch.zeros(N, 32, device="cuda")
        O.gather_concat_fwd(ids, wts, table, d, 0, False, bias, 1, c)
    s.synchronize()
    assert torch.equal(c, a)
