"""Evaluation metrics on the GPU: what `watch_train` (baseline.py:246-259) and the Theano script
(FNN_wnzh.py:187-221) compute with sklearn on the host -- ROC AUC, RMSE, log-loss -- through `dctr_eval_metrics`, so the
predictions of an evaluation pass can stay on the device.  sklearn remains the referee in the tests."""
import numpy as np
import torch

from . import _lib as L


def device_metrics(preds, labels, device="cuda"):
    """preds, labels: numpy arrays or torch tensors of equal length.  Returns dict(auc, log_loss, rmse, n_pos);
    auc is -1 when only one class is present (the reference's `except ValueError` branch)."""
    dev = torch.device(device)
    p = torch.as_tensor(np.asarray(preds, dtype=np.float32) if not torch.is_tensor(preds) else preds).to(dev, torch.float32).contiguous().view(-1)
    y = torch.as_tensor(np.asarray(labels, dtype=np.float32) if not torch.is_tensor(labels) else labels).to(dev, torch.float32).contiguous().view(-1)
    if p.numel() != y.numel() or p.numel() == 0:
        raise ValueError("preds and labels must be non-empty and of equal length")
    n = p.numel()
    ws = torch.empty(L.lib().dctr_eval_metrics_workspace(n), dtype=torch.uint8, device=dev)
    out = torch.zeros(4, dtype=torch.float64, device=dev)
    L.call("dctr_eval_metrics", L.ptr(p), L.ptr(y), n, L.ptr(out), L.ptr(ws), ws.numel(), torch.cuda.current_stream(dev).cuda_stream)
    auc, ll, rmse, n_pos = out.cpu().tolist()
    return dict(auc=auc, log_loss=ll, rmse=rmse, n_pos=int(n_pos))
