"""Evaluation metrics on the GPU: what `watch_train` (baseline.py:246-259) and the Theano script
(FNN_wnzh.py:187-221) compute with sklearn on the host -- ROC AUC, RMSE, log-loss -- through `dctr_eval_metrics`, so the
predictions of an evaluation pass can stay on the device.  sklearn remains the referee in the tests."""
import numpy as np
import torch

from . import _lib as L


def recalibrate(preds, nds_rate):
    """Undo negative down-sampling on the predictions: p / (p + (1 - p) / nds_rate) (baseline.py:369-370, `re_calibration`).
    Elementwise; works on numpy arrays and on device tensors (where it stays on the device)."""
    return preds / (preds + (1.0 - preds) / float(nds_rate))


def early_stop(step, errs, metric="auc", least_step=0, skip_window=1, smooth_window=10, stop_window=10):
    """The stopping rule of baseline.py:262-281 on the record of evaluation metrics `errs`: take every skip_window-th record,
    smooth with a moving average of smooth_window records, and stop when the smoothed metric is worse than it was stop_window - 1
    smoothed records earlier (auc: lower, rmse: higher).  Host logic -- the records are a handful of floats."""
    if step <= least_step:
        return False
    picked = np.asarray(errs, dtype=np.float64)[::skip_window]
    if len(picked) < smooth_window:
        return False
    smooth = np.convolve(picked, np.ones(smooth_window) / smooth_window, mode="valid")
    if len(smooth) < stop_window:
        return False
    change = smooth[-1] - smooth[-stop_window]
    return bool(change > 0) if metric == "rmse" else (bool(change < 0) if metric == "auc" else False)


def device_metrics(preds, labels, device="cuda", nds_rate=None):
    """preds, labels: numpy arrays or torch tensors of equal length.  Returns dict(auc, log_loss, rmse, n_pos);
    auc is -1 when only one class is present (the reference's `except ValueError` branch).  nds_rate: recalibrate the
    predictions first (baseline.py:369-370)."""
    dev = torch.device(device)
    p = torch.as_tensor(np.asarray(preds, dtype=np.float32) if not torch.is_tensor(preds) else preds).to(dev, torch.float32).contiguous().view(-1)
    y = torch.as_tensor(np.asarray(labels, dtype=np.float32) if not torch.is_tensor(labels) else labels).to(dev, torch.float32).contiguous().view(-1)
    if p.numel() != y.numel() or p.numel() == 0:
        raise ValueError("preds and labels must be non-empty and of equal length")
    if nds_rate is not None:
        p = recalibrate(p, nds_rate).contiguous()
    n = p.numel()
    ws = torch.empty(L.lib().dctr_eval_metrics_workspace(n), dtype=torch.uint8, device=dev)
    out = torch.zeros(4, dtype=torch.float64, device=dev)
    L.call("dctr_eval_metrics", L.ptr(p), L.ptr(y), n, L.ptr(out), L.ptr(ws), ws.numel(), torch.cuda.current_stream(dev).cuda_stream)
    auc, ll, rmse, n_pos = out.cpu().tolist()
    return dict(auc=auc, log_loss=ll, rmse=rmse, n_pos=int(n_pos))
