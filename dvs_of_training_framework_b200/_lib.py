"""ctypes binding of libdvsof.so (include/dvsof.h).  No CPU fallback: if the CUDA library is
missing or a tensor is not on a CUDA device, the call fails loudly."""
import ctypes as C
import os
from pathlib import Path

import torch

# DVSOF_LIB selects another build of the same library (kernel tuning sweeps); default: in-tree
_LIB_PATH = Path(os.environ['DVSOF_LIB']).resolve() if os.environ.get('DVSOF_LIB') else \
    Path(__file__).resolve().parent / 'libdvsof.so'
_lib = None

MAX_SCALES = 8
ST_BAD_POLARITY, ST_NOT_GROUPED, ST_BAD_SAMPLE, ST_BAD_PIXEL, ST_NO_MATCH, ST_BAD_LUT, ST_BAD_COUNTS = \
    0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40
_ST_TEXT = {ST_BAD_POLARITY: 'polarity must be +1/-1 (EV_FlowNet/net.py:124)',
            ST_NOT_GROUPED: 'events must be grouped by ascending sample_index (net.py:131-134)',
            ST_BAD_SAMPLE: 'sample_index outside [0, batch)',
            ST_BAD_PIXEL: 'event coordinate outside the image',
            ST_NO_MATCH: 'each prediction needs exactly one image per timestamp (utils/loss.py:193)',
            ST_BAD_LUT: 'map: src/dst index outside the image',
            ST_BAD_COUNTS: 'batch codec: run lengths do not cover the events, or an element index outside the batch'}


class LossScales(C.Structure):
    _fields_ = [('n_scales', C.c_int32), ('n_pred', C.c_int32), ('n_images', C.c_int32), ('reserved_', C.c_int32),
                ('H', C.c_int32 * MAX_SCALES), ('W', C.c_int32 * MAX_SCALES),
                ('flow', C.c_void_p * MAX_SCALES), ('images', C.c_void_p * MAX_SCALES),
                ('grad_flow', C.c_void_p * MAX_SCALES)]


class LossFrames(C.Structure):
    _fields_ = [('images', C.c_void_p), ('n_images', C.c_int32), ('H0', C.c_int32), ('W0', C.c_int32),
                ('flow_ts', C.c_void_p), ('flow_sample_idx', C.c_void_p),
                ('timestamps', C.c_void_p), ('sample_idx', C.c_void_p)]


_SIGNATURES = {
    'dvsof_last_error': (C.c_char_p, []),
    'dvsof_version': (C.c_int, []),
    'dvsof_encode_workspace_bytes': (C.c_size_t, [C.c_int64, C.c_int64, C.c_int32, C.c_int32]),
    'dvsof_encode_event_image': (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_void_p, C.c_int64,
                                            C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t,
                                            C.c_void_p, C.c_void_p]),
    'dvsof_encode_prepare': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                       C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_encode_scatter': (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_int64, C.c_int32, C.c_int32]
                             + [C.c_void_p] * 4),
    'dvsof_encode_path': (C.c_int, [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_size_t]),
    'dvsof_encode_binned_pass': (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                           C.c_void_p, C.c_size_t, C.c_void_p, C.c_int32, C.c_void_p]),
    'dvsof_encode_event_image_compact': (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_void_p,
                                                    C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                                    C.c_size_t, C.c_void_p, C.c_void_p]),
    'dvsof_voxel_workspace_bytes': (C.c_size_t, [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32]),
    'dvsof_voxel_path': (C.c_int, [C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_size_t]),
    'dvsof_encode_voxel_grid': (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_void_p, C.c_int64, C.c_int32,
                                           C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t,
                                           C.c_void_p, C.c_void_p]),
    'dvsof_encode_static_max_samples': (C.c_int64, []),
    'dvsof_encode_event_image_static': (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_void_p, C.c_int64,
                                                   C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_count_image': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                    C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_compact_workspace_bytes': (C.c_size_t, [C.c_int64]),
    'dvsof_event_crop': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_double,
                                   C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p]),
    'dvsof_map_lut_workspace_bytes': (C.c_size_t, [C.c_int32, C.c_int32, C.c_int64]),
    'dvsof_map_events_build': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_map_events_count': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_map_events_write': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_augment_events_count': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                             C.c_int32, C.c_void_p, C.POINTER(C.c_double), C.c_void_p,
                                             C.c_void_p, C.c_void_p]),
    'dvsof_augment_events_write': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                             C.c_int32, C.c_void_p, C.POINTER(C.c_double), C.c_void_p,
                                             C.c_void_p, C.c_void_p]),
    'dvsof_codec_workspace_bytes': (C.c_size_t, [C.c_int64, C.c_int64]),
    'dvsof_decode_events': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_encode_events': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_upsample2x_reflect_pad_forward': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                                       C.c_void_p, C.c_int32, C.c_void_p]),
    'dvsof_upsample2x_reflect_pad_backward': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                                        C.c_void_p, C.c_int32, C.c_void_p]),
    'dvsof_upsample2x_reflect_pad_cat_forward': (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.c_int32,
                                                           C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                                           C.c_void_p]),
    'dvsof_upsample2x_reflect_pad_cat_backward': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int32),
                                                            C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                                            C.c_void_p]),
    'dvsof_relu_bn_workspace_bytes': (C.c_size_t, [C.c_int64, C.c_int32, C.c_int64, C.c_int32]),
    'dvsof_relu_bn_forward': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    'dvsof_relu_bn_backward': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    'dvsof_window_offsets': (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int64,
                                       C.c_void_p, C.c_void_p]),
    'dvsof_flow_error_workspace_bytes': (C.c_size_t, []),
    'dvsof_flow_error_dense': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p,
                                         C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                         C.c_void_p]),
    'dvsof_match_timestamps': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                         C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    'dvsof_resize_bilinear': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                        C.c_int32, C.c_int32, C.c_void_p]),
    'dvsof_loss_workspace_bytes': (C.c_size_t, [C.POINTER(LossScales)]),
    'dvsof_warp_loss_forward': (C.c_int, [C.POINTER(LossScales), C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
    'dvsof_multiscale_loss_forward': (C.c_int, [C.POINTER(LossScales), C.POINTER(LossFrames)]
                                      + [C.c_void_p] * 6),
    'dvsof_warp_loss_backward': (C.c_int, [C.POINTER(LossScales), C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p]),
    'dvsof_multiscale_loss_fused': (C.c_int, [C.POINTER(LossScales), C.POINTER(LossFrames),
                                              C.POINTER(C.c_float)] + [C.c_void_p] * 7),
    'dvsof_warp_loss_fused': (C.c_int, [C.POINTER(LossScales), C.c_void_p, C.POINTER(C.c_float),
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                        C.c_void_p]),
    'dvsof_scale_gradients': (C.c_int, [C.POINTER(LossScales), C.c_void_p, C.c_void_p]),
}


def lib_path():
    return _LIB_PATH


def load():
    """Load libdvsof.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not _LIB_PATH.exists():
            raise RuntimeError(f'{_LIB_PATH} is missing: run `python -c "import __graft_entry__ as g; '
                               'g.build()"` (nvcc, sm_100a). There is no CPU fallback.')
        lib = C.CDLL(str(_LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def call(name, *args):
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.dvsof_last_error().decode()
        raise (ValueError if rc == -1 else RuntimeError)(f'{name}: {msg} (code {rc})')


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() else C.c_void_p(0)


def _index(device):
    i = device.index if isinstance(device, torch.device) else torch.device(device).index
    return torch.cuda.current_device() if i is None else i


def raw_stream(device):
    """The current stream of `device` as an integer handle (torch.cuda.current_stream(...).cuda_stream
    without building the Stream object: ~0.3 us instead of ~10 us on the hot path)."""
    return torch._C._cuda_getCurrentRawStream(_index(device))


def stream_ptr(device):
    return C.c_void_p(raw_stream(device))


class on_device:
    """`with torch.cuda.device(dev)` that does nothing when `dev` already is the current device."""
    __slots__ = ('idx', 'ctx')

    def __init__(self, device):
        self.idx, self.ctx = _index(device), None

    def __enter__(self):
        if torch.cuda.current_device() != self.idx:
            self.ctx = torch.cuda.device(self.idx)
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)


def require_cuda(t, name, dtype=None):
    if isinstance(t, torch.Tensor) and t.is_cuda and (dtype is None or t.dtype is dtype) and t.is_contiguous():
        return t                                                  # the hot path
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f'{name}: expected a CUDA tensor (libdvsof has no CPU path), got '
                           f'{type(t).__name__}{"" if not isinstance(t, torch.Tensor) else " on " + str(t.device)}')
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f'{name}: expected {dtype}, got {t.dtype}')
    return t.contiguous()


class Status:
    """Status word the kernels OR error bits into.  It lives in pinned HOST memory that the device
    addresses directly (unified addressing: the device pointer of a pinned allocation is its host
    pointer), so looking at it is a plain memory read -- no copy, no event, no CUDA call -- and a
    kernel only touches it when it has an error to report.

    mode 'sync'     -- check right after the call (reference behaviour: AssertionError raised
                       by the call itself; one stream synchronisation per call);
    mode 'deferred' -- nothing happens after a call; the word is looked at before the next call
                       (`flush`: whatever finished kernels have reported by then) and at
                       `flush(wait=True)` / `synchronize()`: no sync on the hot path, errors
                       surface one or two calls late.
    """
    mode = 'sync'

    def __init__(self, device):
        self.device = torch.device(device)
        self.host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.word = self.host                        # what the kernels are handed (`_lib.ptr(st.word)`)
        self._np = self.host.numpy()                 # plain memory view for the hot path

    def _raise(self, bits, exc):
        raise exc('; '.join(t for b, t in _ST_TEXT.items() if bits & b))

    def _look(self, exc):
        bits = int(self._np[0]) & 0xffffffff
        if bits:
            torch.cuda.current_stream(self.device).synchronize()    # let every reporter finish first
            bits = int(self._np[0]) & 0xffffffff
            self._np[0] = 0
            self._raise(bits, exc)

    def check(self, exc=AssertionError):
        torch.cuda.current_stream(self.device).synchronize()
        self._look(exc)

    def after_call(self, exc=AssertionError):
        if Status.mode == 'sync':
            self.check(exc)

    def flush(self, exc=AssertionError, wait=False):
        if wait:
            self.check(exc)
        else:
            self._look(exc)

    def synchronize(self, exc=AssertionError):
        self.check(exc)


_status = {}


def status_for(device):
    device = torch.device(device)
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    st = _status.get(key)
    if st is None:
        st = _status[key] = Status(device)
    return st
