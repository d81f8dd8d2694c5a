"""Evaluation step on the GPU -- drop-in for the reference's `utils/eval.py:6-50`
`flow_error_dense(flow_gt, flow_pred, event_img, is_car=False, is_dense=False)`, called per
evaluation window by `utils/testing.py:84-92` right after `get_count_image`.

numpy or CUDA-tensor inputs ((H, W, 2) flows, (H, W) event-count image); returns
`(AEE, percent_AEE, n_points)` as Python numbers like the reference (one D2H of 24 bytes).
`flow_error_dense_device` returns the three raw sums as a CUDA tensor without synchronising."""
import numpy as np
import torch

from . import _lib

_ws = {}


def _cuda(a, name):
    t = torch.as_tensor(np.ascontiguousarray(a)) if isinstance(a, np.ndarray) else a
    if not isinstance(t, torch.Tensor):
        raise TypeError(f'{name}: numpy array or torch tensor expected, got {type(a).__name__}')
    return (t if t.is_cuda else t.cuda()).contiguous()


def flow_error_dense_device(flow_gt, flow_pred, event_img, is_car=False, is_dense=False):
    """[sum of endpoint errors, number of points, number of points with EE < 3] as a (3,) fp64
    CUDA tensor (stream-ordered, no host sync)."""
    gt, pred = _cuda(flow_gt, 'flow_gt'), _cuda(flow_pred, 'flow_pred')
    for t, name in ((gt, 'flow_gt'), (pred, 'flow_pred')):
        if t.dtype not in (torch.float32, torch.float64):
            raise TypeError(f'{name}: float32/float64 expected, got {t.dtype}')
    if gt.dim() != 3 or gt.shape[2] != 2 or pred.shape != gt.shape:
        raise ValueError(f'flow_gt {tuple(gt.shape)} / flow_pred {tuple(pred.shape)}: (H, W, 2) expected')
    H, W = gt.shape[:2]
    # eval.py:16-22: `max_row = flow_gt.shape[1]` (the WIDTH, as written there) or 190, applied
    # as a row slice -- all rows for landscape frames, kept as is for drop-in behaviour
    rows = min(H, 190 if is_car else W)
    dev = gt.device
    cnt = None
    if not is_dense:
        cnt = _cuda(event_img, 'event_img')
        cnt = cnt.reshape([s for s in cnt.shape if s != 1] or [1])        # np.squeeze, eval.py:20
        if cnt.dim() != 2 or cnt.shape[0] < rows or cnt.shape[1] != W:
            raise ValueError(f'event_img {tuple(cnt.shape)} does not cover ({rows}, {W})')
        if cnt.dtype in (torch.float32, torch.float64):
            cnt = (cnt > 0).to(torch.int64)                               # only `> 0` is used
        elif cnt.dtype != torch.int64:
            cnt = cnt.to(torch.int64)
        cnt = cnt.contiguous()
    out = torch.empty(3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        key = (dev, torch.cuda.current_stream(dev).cuda_stream)     # the ticket word is per stream
        ws = _ws.get(key)
        if ws is None:
            ws = _ws[key] = torch.zeros(_lib.load().dvsof_flow_error_workspace_bytes(), dtype=torch.uint8,
                                        device=dev)
        _lib.call('dvsof_flow_error_dense', _lib.ptr(gt), gt.element_size(), _lib.ptr(pred),
                  pred.element_size(), _lib.ptr(cnt) if cnt is not None else None, rows, W,
                  1 if is_dense else 0, _lib.ptr(out), _lib.ptr(ws), _lib.stream_ptr(dev))
    return out


def flow_error_dense(flow_gt, flow_pred, event_img, is_car=False, is_dense=False):
    """utils/eval.py:6-50.  Returns (AEE, percent_AEE, n_points)."""
    s, n, below = flow_error_dense_device(flow_gt, flow_pred, event_img, is_car, is_dense).tolist()
    n_points = int(n)
    aee = s / n_points if n_points else float('nan')                      # np.mean of an empty array
    return aee, float(below) / float(n_points + 1e-5), n_points
