"""Event-image encoder -- drop-in for `EV_FlowNet/net.py::compute_event_image` (:88-150) and
`Model.quantize` / `_extend_size` (:194-205) of the reference, running the hand-written
sm_100a kernel of csrc/encode.cu through the C ABI (include/dvsof.h)."""
import numpy as np
import torch

from . import _lib

_EVENT_KEYS = ('x', 'y', 'timestamp', 'polarity', 'sample_index')


def extend_size(imsize):
    """EV_FlowNet/net.py:194-195 `_extend_size`: each side rounded up to a multiple of 16."""
    return tuple(map(lambda x: ((x - 1) // 16 + 1) * 16, imsize))


def compute_event_image(events, timestamps, sample_idx, imsize, device='cuda',
                        dtype=torch.float32):
    """Same signature and result as the reference (EV_FlowNet/net.py:88-93): (B,4,H,W) fp32
    [count(p=+1), count(p=-1), max t^(p=+1), max t^(p=-1)].  Tensors must live on a CUDA
    device (numpy input is uploaded to `device`, as the reference's numpy branch :105-109)."""
    assert dtype == torch.float32, 'the event image is computed in fp32'
    assert len(imsize) == 2
    if not isinstance(timestamps, torch.Tensor):                       # net.py:105-109
        dev = torch.device(device)
        events = {k: torch.as_tensor(np.asarray(v), device=dev) for k, v in events.items()
                  if k != 'size'}
        for k in ('x', 'y', 'polarity', 'sample_index'):
            events[k] = events[k].to(torch.int64)
        events['timestamp'] = events['timestamp'].to(torch.float32)
        timestamps = torch.as_tensor(np.asarray(timestamps), dtype=torch.float32, device=dev)
        sample_idx = torch.as_tensor(np.asarray(sample_idx), dtype=torch.int64, device=dev)
    assert timestamps.ndim == 1                                        # net.py:99
    assert sample_idx.ndim == 1                                        # net.py:100
    timestamps = _lib.require_cuda(timestamps, 'timestamps', torch.float32)
    dev = timestamps.device
    bs = timestamps.numel() // 2
    assert bs * 2 == timestamps.numel() and sample_idx.numel() == timestamps.numel()
    if _lib.Status.mode == 'sync':                                     # net.py:101 (reads a value)
        assert (int(sample_idx[-1]) + 1) * 2 == timestamps.numel()
    H, W = int(imsize[0]), int(imsize[1])
    out = torch.empty((bs, 4, H, W), dtype=torch.float32, device=dev)
    ev = {}
    for k in _EVENT_KEYS:
        ev[k] = _lib.require_cuda(events[k], f"events['{k}']",
                                  torch.float32 if k == 'timestamp' else torch.int64)
    n = ev['x'].numel()
    for k in _EVENT_KEYS:
        assert ev[k].numel() == n, f"events['{k}'] has {ev[k].numel()} entries, expected {n}"
    with _lib.on_device(dev):
        st = _lib.status_for(dev)
        st.flush()
        ws = torch.empty(_lib.load().dvsof_encode_workspace_bytes(n, bs, H, W), dtype=torch.uint8, device=dev)
        _lib.call('dvsof_encode_event_image', _lib.ptr(ev['x']), _lib.ptr(ev['y']),
                  _lib.ptr(ev['timestamp']), _lib.ptr(ev['polarity']), _lib.ptr(ev['sample_index']),
                  n, _lib.ptr(timestamps), bs, H, W, _lib.ptr(out), _lib.ptr(ws), ws.numel(),
                  _lib.ptr(st.word), _lib.stream_ptr(dev))
        if n:
            st.after_call()
    return out


def compute_event_image_compact(x, y, timestamp, polarity, events_per_sample, timestamps, imsize):
    """Storage-form entry: x,y int16, polarity bool/uint8 (1 = positive),
    timestamp fp32, as written by utils/dataset.py:286-289 `encode_batch`; `events_per_sample`
    int64 (B,).  Equals compute_event_image(decode_batch(...)) without the int64 blow-up."""
    timestamps = _lib.require_cuda(timestamps, 'timestamps', torch.float32)
    dev = timestamps.device
    bs = timestamps.numel() // 2
    x = _lib.require_cuda(x, 'x', torch.int16)
    y = _lib.require_cuda(y, 'y', torch.int16)
    t = _lib.require_cuda(timestamp, 'timestamp', torch.float32)
    p = _lib.require_cuda(polarity, 'polarity')
    if p.dtype == torch.bool:
        p = p.view(torch.uint8)
    assert p.dtype == torch.uint8
    eps = _lib.require_cuda(events_per_sample, 'events_per_sample', torch.int64)
    assert eps.numel() == bs
    begin = torch.zeros(bs + 1, dtype=torch.int64, device=dev)
    torch.cumsum(eps, 0, out=begin[1:])
    n = x.numel()
    H, W = int(imsize[0]), int(imsize[1])
    out = torch.empty((bs, 4, H, W), dtype=torch.float32, device=dev)
    with _lib.on_device(dev):
        st = _lib.status_for(dev)
        st.flush()
        ws = torch.empty(_lib.load().dvsof_encode_workspace_bytes(n, bs, H, W), dtype=torch.uint8, device=dev)
        _lib.call('dvsof_encode_event_image_compact', _lib.ptr(x), _lib.ptr(y), _lib.ptr(t),
                  _lib.ptr(p), _lib.ptr(begin), n, _lib.ptr(timestamps), bs, H, W, _lib.ptr(out),
                  _lib.ptr(ws), ws.numel(), _lib.ptr(st.word), _lib.stream_ptr(dev))
        if n:
            st.after_call()
    return out


def quantize(events, timestamps, sample_idx, imsize):
    """EV_FlowNet/net.py:197-205 `Model.quantize`: encode at the x16-extended size, no grad."""
    with torch.no_grad():
        return compute_event_image(events, timestamps, sample_idx, extend_size(imsize))


def compute_voxel_grid(events, timestamps, sample_idx, imsize, n_bins):
    """EXTENSION -- not in the reference (its network input is `compute_event_image`): the
    temporal-bilinear voxel grid (B, n_bins, H, W) fp32 that BASELINE.json's "bins" configs name
    (Zhu et al., CVPR 2019): `t* = (n_bins - 1) (t - start) / dt` with the event image's start / dt
    (EV_FlowNet/net.py:136-141), `out[b, k, y, x] = sum p * max(0, 1 - |t* - k|)`.  Same event
    dictionary, timestamps and sample_idx as `compute_event_image`; CUDA tensors only."""
    assert len(imsize) == 2 and timestamps.ndim == 1 and sample_idx.ndim == 1
    timestamps = _lib.require_cuda(timestamps, 'timestamps', torch.float32)
    dev = timestamps.device
    bs = timestamps.numel() // 2
    assert bs * 2 == timestamps.numel() and sample_idx.numel() == timestamps.numel()
    H, W, C = int(imsize[0]), int(imsize[1]), int(n_bins)
    ev = {k: _lib.require_cuda(events[k], f"events['{k}']", torch.float32 if k == 'timestamp' else torch.int64)
          for k in _EVENT_KEYS}
    n = ev['x'].numel()
    for k in _EVENT_KEYS:
        assert ev[k].numel() == n, f"events['{k}'] has {ev[k].numel()} entries, expected {n}"
    out = torch.empty((bs, C, H, W), dtype=torch.float32, device=dev)
    with _lib.on_device(dev):
        st = _lib.status_for(dev)
        st.flush()
        ws = torch.empty(_lib.load().dvsof_voxel_workspace_bytes(n, bs, H, W, C), dtype=torch.uint8, device=dev)
        _lib.call('dvsof_encode_voxel_grid', _lib.ptr(ev['x']), _lib.ptr(ev['y']), _lib.ptr(ev['timestamp']),
                  _lib.ptr(ev['polarity']), _lib.ptr(ev['sample_index']), n, _lib.ptr(timestamps), bs, C, H, W,
                  _lib.ptr(out), _lib.ptr(ws), ws.numel(), _lib.ptr(st.word), _lib.stream_ptr(dev))
        if n:
            st.after_call()
    return out
