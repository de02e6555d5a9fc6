"""CUDA path vs the committed golden fixtures (tests/golden/oracle_golden.npz; oracle-generated, see
make_golden.py).  Nothing here reads /root/reference or calls the oracle."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "oracle_golden.npz")


def close(a, b, rtol=3e-5, atol=2e-6):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


@pytest.fixture(scope="module")
def g():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return np.load(GOLD)


def test_theano_fnn_trajectory(g):
    from deepctr_b200.theano_fnn import FNN_wnzh, init_mlp
    ids, y, table = g["ids"], g["y"], g["table0"]
    X, d = table.shape
    F, N = ids.shape[1], ids.shape[0]
    Ws, bs = init_mlp(1 + F * d, 20, 10)
    Ws[2] = g["th_w3"]
    m = FNN_wnzh(dict(w_0=-0.1, table=table, feat_field=np.repeat(np.arange(F), X // F)), batch_size=N, lr=0.01, dropout=0.5,
                 hidden1=20, hidden2=10, lambda_fm=0.1, weights=(Ws, bs))
    losses = []
    for _ in range(3):
        m.set_masks(g["th_r1"], g["th_r2"])
        l, p = m.train(ids, y)
        losses.append(l)
    close(losses, g["th_losses"])
    close(p, g["th_p"])
    close(m.get_table(), g["th_table"])
    close(m.get_weights()[0], g["th_w1"])


def test_fm_trajectory(g):
    from deepctr_b200.models import FM
    import pickle, tempfile
    ids, wts, y, table = g["ids"], g["wts"], g["y"], g["table0"]
    X, d = table.shape
    N, F = ids.shape
    with tempfile.NamedTemporaryFile(suffix=".pickle", delete=False) as f:
        pickle.dump({"W": table[:, :1].astype(np.float32), "V": table[:, 1:].astype(np.float32), "b": np.array([0.05], np.float32)}, f)
    fm = FM(N, [X, F, d - 1], ["uniform", -1, 1, [1, 2], f.name], ["sgd", 0.05, "sum"], [0.01], "train", N)
    for _ in range(2):
        l, p = fm.train_step({fm.sp_id_hldr: ids.reshape(-1), fm.sp_wt_hldr: wts.reshape(-1), fm.lbl_hldr: y})
    close(p, g["fm_p"])
    close(fm.engine.get_table(), g["fm_table"])
    close(fm.engine.bias().cpu(), g["fm_b"])
    os.unlink(f.name)


@pytest.mark.parametrize("cross", ["none", "ip", "op"])
def test_embedding_layers(g, cross):
    from deepctr_b200 import _lib as L
    ids, wts, table = g["ids"], g["wts"], g["table0"]
    X, d = table.shape
    N, F = ids.shape
    want = g["z_" + cross]
    ldz = (want.shape[1] + 3) // 4 * 4
    t = torch.zeros(X, 12, device="cuda"); t[:, :d] = torch.as_tensor(table, dtype=torch.float32).cuda()
    idd, wd = torch.as_tensor(ids, dtype=torch.int32).cuda(), torch.as_tensor(wts, dtype=torch.float32).cuda()
    bias, out = torch.tensor([0.25], device="cuda"), torch.zeros(N, ldz, device="cuda")
    L.call("dctr_gather_concat_fwd", idd.data_ptr(), wd.data_ptr(), t.data_ptr(), X, 12, N, F, d, L.CROSS[cross], 0,
           bias.data_ptr(), 0, out.data_ptr(), ldz, None)
    torch.cuda.synchronize()
    close(out.cpu().numpy()[:, :want.shape[1]], want, 1e-5, 1e-6)
