"""`utils.transformation.map` of the reference (pybind11 module built from
utils/modules_to_build/transformations/src/{transformation.cpp:35-85, module.cpp:18-40}) on the
GPU: count -> scan -> write LUT remap (csrc/transforms.cu).

    map(events, shape, src, dst) -> events'

numpy in -> numpy out (host<->device copies inside, like the reference's array-in/array-out
call); CUDA tensors in -> CUDA tensor out.  dtype/layout rules of the pybind signature are kept:
events float32 (N,E) C-contiguous, src/dst uint64 (M,), otherwise TypeError; size mismatch ->
ValueError (std::invalid_argument)."""
import numpy as np
import torch

from . import _lib


class RotationLUT:
    """A LUT uploaded and indexed once (per-pixel CSR lists), reusable over many event arrays."""

    def __init__(self, shape, src, dst, device='cuda'):
        self.H, self.W = int(shape[0]), int(shape[1])
        src = torch.as_tensor(src, device=device).contiguous()
        dst = torch.as_tensor(dst, device=device).contiguous()
        if src.shape[0] != dst.shape[0]:
            raise ValueError('src and dst should have the same size!')   # transformation.cpp:41-42
        self.device = src.device
        M = src.shape[0]
        with torch.cuda.device(self.device):
            lib = _lib.load()
            self.ws = torch.empty(lib.dvsof_map_lut_workspace_bytes(self.H, self.W, M),
                                  dtype=torch.uint8, device=self.device)
            st = _lib.status_for(self.device)
            st.flush(ValueError)
            _lib.call('dvsof_map_events_build', _lib.ptr(src), _lib.ptr(dst), M, self.H, self.W,
                      _lib.ptr(self.ws), _lib.ptr(st.word), _lib.stream_ptr(self.device))
            st.after_call(ValueError)

    def apply(self, events):
        """events: CUDA fp32 (N,E) -> CUDA fp32 (N_out,E)."""
        events = _lib.require_cuda(events, 'events', torch.float32)
        n, E = events.shape
        dev = events.device
        with torch.cuda.device(dev):
            lib = _lib.load()
            ev_ws = torch.empty(lib.dvsof_compact_workspace_bytes(n), dtype=torch.uint8, device=dev)
            n_out = torch.empty(1, dtype=torch.int64, device=dev)
            _lib.call('dvsof_map_events_count', _lib.ptr(events), n, E, self.H, self.W,
                      _lib.ptr(self.ws), _lib.ptr(ev_ws), _lib.ptr(n_out), _lib.stream_ptr(dev))
            total = int(n_out.item())          # output size is data dependent (one sync)
            out = torch.empty((total, E), dtype=torch.float32, device=dev)
            _lib.call('dvsof_map_events_write', _lib.ptr(events), n, E, self.H, self.W,
                      _lib.ptr(self.ws), _lib.ptr(ev_ws), _lib.ptr(out), _lib.stream_ptr(dev))
        return out


def map(events, shape, src, dst):  # noqa: A001 -- the reference's exported name
    is_np = isinstance(events, np.ndarray)
    if is_np:
        if events.dtype != np.float32 or events.ndim != 2 or not events.flags['C_CONTIGUOUS']:
            raise TypeError('map(): events must be a C-contiguous float32 (N, E) array')
        for name, a in (('src', src), ('dst', dst)):
            if not isinstance(a, np.ndarray) or a.dtype != np.uint64 or a.ndim != 1:
                raise TypeError(f'map(): {name} must be a 1-D uint64 array')
        ev = torch.from_numpy(events).cuda()
        # uint64 -> reinterpret as int64 for the copy (values < 2^63)
        src_t = torch.from_numpy(src.view(np.int64)).cuda()
        dst_t = torch.from_numpy(dst.view(np.int64)).cuda()
    else:
        ev, src_t, dst_t = events, src, dst
    lut = RotationLUT(shape, src_t, dst_t, device=ev.device)
    out = lut.apply(ev)
    return out.cpu().numpy() if is_np else out
