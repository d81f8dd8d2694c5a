"""Event transforms of the reference's `utils/data.py` on the GPU: `get_count_image` (:120-136),
`EventCrop` (:24-42), `RandomRotation` (:155-220), `frame_generator` (:139-152).  numpy in -> numpy out (copies inside);
CUDA tensors in -> CUDA tensors out."""
import numpy as np
import torch

from . import _lib
from .transformation import RotationLUT


def _to_cuda(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a)
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda() if not t.is_cuda else t


def get_count_image(events, imsize):
    """utils/data.py:120-136: events = [x, y, ...] arrays -> (H, W) event counts.
    numpy input returns np.uint64 like the reference; CUDA input returns an int64 CUDA tensor
    (torch has no uint64 arithmetic; same bit pattern)."""
    is_np = not isinstance(events[0], torch.Tensor)
    x = _to_cuda(np.asarray(events[0]).astype(int) if is_np else events[0], torch.int64).contiguous()
    y = _to_cuda(np.asarray(events[1]).astype(int) if is_np else events[1], torch.int64).contiguous()
    H, W = int(imsize[0]), int(imsize[1])
    dev = x.device
    out = torch.empty((H, W), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        st = _lib.status_for(dev)
        st.flush()
        _lib.call('dvsof_count_image', _lib.ptr(x), _lib.ptr(y), x.numel(), H, W, _lib.ptr(out),
                  _lib.ptr(st.word), _lib.stream_ptr(dev))
        if is_np:
            st.check(ValueError)     # np.ravel_multi_index raises ValueError on out-of-range
        else:
            st.after_call(ValueError)
    return out.cpu().numpy().view(np.uint64) if is_np else out


class EventCrop:
    """utils/data.py:24-42: keep events inside box = [y0, x0, h, w] and shift them to its origin;
    order preserved.  events: (N, >=2) float32/float64, columns x, y, ..."""

    def __init__(self, box):
        self.box = box

    def __call__(self, events, box=None):
        if box is None:
            box = self.box
        is_np = isinstance(events, np.ndarray)
        ev = _to_cuda(events)
        if ev.dtype not in (torch.float32, torch.float64):
            raise TypeError(f'EventCrop: float32/float64 events expected, got {ev.dtype}')
        ev = ev.contiguous()
        n, E = ev.shape
        dev = ev.device
        out = torch.empty_like(ev)
        with torch.cuda.device(dev):
            lib = _lib.load()
            ws = torch.empty(lib.dvsof_compact_workspace_bytes(n), dtype=torch.uint8, device=dev)
            n_out = torch.empty(1, dtype=torch.int64, device=dev)
            _lib.call('dvsof_event_crop', _lib.ptr(ev), n, E, ev.element_size(), float(box[0]),
                      float(box[1]), float(box[2]), float(box[3]), _lib.ptr(out), _lib.ptr(n_out),
                      _lib.ptr(ws), _lib.stream_ptr(dev))
            out = out[:int(n_out.item())]
        return out.cpu().numpy() if is_np else out


def RandomRotation(interval, shape):
    """utils/data.py:155-220.  Returns rotation(images, events, angle=None) ->
    (rotated images, remapped events, angle).  The nearest-neighbour LUT is built on the host
    exactly as the reference does (float64 rotation + rint); the image gather and the event
    remap run on the GPU."""
    H, W = int(shape[0]), int(shape[1])
    xs, ys = np.meshgrid(range(W), range(H))
    xs, ys = xs.ravel(), ys.ravel()
    idx = ys * W + xs
    centred = np.vstack(((xs.astype(float) - W / 2)[None], (ys.astype(float) - H / 2)[None]))
    if not hasattr(interval, '__len__'):
        interval = (-abs(interval), abs(interval))
    assert len(interval) == 2, "You should specify at most two anges"
    assert interval[0] <= interval[1], "The first interval should be lower or equal to the second"

    def rotation(images, events, angle=None):
        if angle is None:
            angle = np.random.rand() * (interval[1] - interval[0]) + interval[0]
        rad = angle * np.pi / 180
        mat = np.array([[np.cos(rad), -np.sin(rad)], [np.sin(rad), np.cos(rad)]])
        r = mat.dot(centred)
        x1 = np.rint(r[0] + W / 2).astype(int)
        y1 = np.rint(r[1] + H / 2).astype(int)
        m = (x1 >= 0) & (x1 < W) & (y1 >= 0) & (y1 < H)
        cur_idx = idx[m]
        cur_ridx = y1[m] * W + x1[m]
        is_np = isinstance(images, np.ndarray)
        dev = 'cuda' if is_np else images.device
        d_idx = torch.as_tensor(cur_idx, dtype=torch.int64, device=dev)
        d_ridx = torch.as_tensor(cur_ridx, dtype=torch.int64, device=dev)
        img = _to_cuda(images)
        flat = img.reshape(img.shape[0], -1)
        rimg = torch.zeros_like(flat)
        rimg[:, d_idx] = flat[:, d_ridx]                     # image[y, x] = image[y1, x1]
        rimg = rimg.reshape(img.shape)
        ev = _to_cuda(events, torch.float32).contiguous()
        revents = RotationLUT((H, W), d_ridx, d_idx, device=ev.device).apply(ev)
        if is_np:
            return rimg.cpu().numpy(), revents.cpu().numpy(), angle
        return rimg, revents, angle

    return rotation


def rotation_lut(shape, angle, device='cuda'):
    """The nearest-neighbour LUT of utils/data.py:186-200 for `angle` degrees as a device-side
    `RotationLUT` (reusable over many event arrays; feeds `augment_events`)."""
    H, W = int(shape[0]), int(shape[1])
    xs, ys = np.meshgrid(range(W), range(H))
    xs, ys = xs.ravel(), ys.ravel()
    rad = angle * np.pi / 180
    mat = np.array([[np.cos(rad), -np.sin(rad)], [np.sin(rad), np.cos(rad)]])
    r = mat.dot(np.vstack(((xs.astype(float) - W / 2)[None], (ys.astype(float) - H / 2)[None])))
    x1 = np.rint(r[0] + W / 2).astype(int)
    y1 = np.rint(r[1] + H / 2).astype(int)
    m = (x1 >= 0) & (x1 < W) & (y1 >= 0) & (y1 < H)
    return RotationLUT((H, W), torch.as_tensor(y1[m] * W + x1[m], dtype=torch.int64, device=device),
                       torch.as_tensor((ys * W + xs)[m], dtype=torch.int64, device=device), device=device)


def augment_events(events, shape, is_flip=False, lut=None, box=None):
    """The event side of utils/dataset.py:753-769 in one device pass: horizontal flip
    (`x = W - x - 1`), rotation remap through `lut` (a `RotationLUT`, e.g. `rotation_lut(shape,
    angle)`; None = no rotation) and `EventCrop(box)` with box = [y0, x0, h, w] (None = no crop).
    Same rows in the same order as the three reference steps applied one after the other.
    events: (N, E>=2) float32, numpy or CUDA tensor; returns the same kind."""
    is_np = isinstance(events, np.ndarray)
    ev = _lib.require_cuda(_to_cuda(events), 'events', torch.float32)
    n, E = ev.shape
    H, W = int(shape[0]), int(shape[1])
    dev = ev.device
    cbox = (_lib.C.c_double * 4)(*[float(v) for v in box[:4]]) if box is not None else None
    lut_ptr = _lib.ptr(lut.ws) if lut is not None else None
    if lut is not None:
        assert (lut.H, lut.W) == (H, W), 'LUT built for another frame size'
    with torch.cuda.device(dev):
        ws = torch.empty(_lib.load().dvsof_compact_workspace_bytes(n), dtype=torch.uint8, device=dev)
        n_out = torch.empty(1, dtype=torch.int64, device=dev)
        _lib.call('dvsof_augment_events_count', _lib.ptr(ev), n, E, H, W, int(bool(is_flip)), lut_ptr, cbox,
                  _lib.ptr(ws), _lib.ptr(n_out), _lib.stream_ptr(dev))
        out = torch.empty((int(n_out.item()), E), dtype=torch.float32, device=dev)   # data-dependent size
        _lib.call('dvsof_augment_events_write', _lib.ptr(ev), n, E, H, W, int(bool(is_flip)), lut_ptr, cbox,
                  _lib.ptr(ws), _lib.ptr(out), _lib.stream_ptr(dev))
    return out.cpu().numpy() if is_np else out


def frame_offsets(t, frames):
    """Per-window event ranges: `np.searchsorted(t, frames.ravel(), side='right').reshape(-1, 2)`
    (utils/data.py:150) computed on the device.  t: sorted timestamps (numpy or CUDA tensor,
    fp32/fp64); frames: [(start_ts, stop_ts), ...].  Returns an int64 CUDA tensor (F, 2)."""
    tt = _to_cuda(t)
    if tt.dtype not in (torch.float32, torch.float64):
        tt = tt.to(torch.float64)
    tt = tt.contiguous()
    dev = tt.device
    bounds = torch.as_tensor(np.asarray(frames, dtype=np.float64).ravel()).to(dev)
    out = torch.empty(bounds.numel(), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call('dvsof_window_offsets', _lib.ptr(tt), tt.element_size(), tt.numel(), _lib.ptr(bounds),
                  bounds.numel(), _lib.ptr(out), _lib.stream_ptr(dev))
    return out.view(-1, 2)


def frame_generator(events, frames):
    """utils/data.py:139-152: yields ([component[i_start:i_stop] for component in events], start,
    stop) for every (start, stop) window; events = [x, y, t, p] sorted by t.  The window table is
    built by one kernel launch; slices are views of the caller's arrays (numpy or CUDA)."""
    frames = np.array(frames)
    idx = frame_offsets(events[2], frames).cpu().numpy()
    for (start, stop), (i_start, i_stop) in zip(frames, idx):
        yield [p[i_start:i_stop] for p in events], start, stop
