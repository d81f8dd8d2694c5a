"""TEST INFRASTRUCTURE: ctypes access to the oracle's two native builds of `map_events`
(oracle/Makefile):

  ref_map(...)   the REFERENCE's own utils/modules_to_build/transformations/src/transformation.cpp
                 (:35-85), compiled from /root/reference against a container stand-in for the
                 Eigen headers (oracle/ref_shim/transformation.h) -> oracle/_ref/libmap_events_ref.so
  port_map(...)  oracle/map_events_ref.cpp, the plain-C++ restatement -> oracle/libmap_events_port.so

Both take what `transformation.map` takes (src/module.cpp:18-40): events float32 (N, E) row-major,
shape (H, W), src / dst uint64 (M,), and return a fresh float32 (N_out, E) array; a size mismatch
raises ValueError as the pybind binding does for std::invalid_argument.
"""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF_SO = HERE / '_ref' / 'libmap_events_ref.so'
PORT_SO = HERE / 'libmap_events_port.so'
_libs = {}


def build():
    subprocess.run(['make', '-s', '-C', str(HERE)], check=True)


def available(kind):
    return (REF_SO if kind == 'ref' else PORT_SO).exists()


def _fn(kind):
    if kind not in _libs:
        path = REF_SO if kind == 'ref' else PORT_SO
        if not path.exists() and kind == 'port':
            build()
        lib = C.CDLL(str(path))
        fn = getattr(lib, f'{kind}_map_events')
        fn.restype = C.c_int
        fn.argtypes = [C.c_void_p, C.c_long, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_long, C.c_void_p,
                       C.c_long, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_long)]
        free = getattr(lib, f'{kind}_map_events_free')
        free.argtypes = [C.POINTER(C.c_float)]
        free.restype = None
        _libs[kind] = (fn, free)
    return _libs[kind]


def _call(kind, events, shape, src, dst):
    events = np.ascontiguousarray(events, dtype=np.float32)
    src = np.ascontiguousarray(src, dtype=np.uint64)
    dst = np.ascontiguousarray(dst, dtype=np.uint64)
    assert events.ndim == 2
    fn, free = _fn(kind)
    out, n_out = C.POINTER(C.c_float)(), C.c_long(0)
    rc = fn(events.ctypes.data, events.shape[0], events.shape[1], int(shape[0]), int(shape[1]),
            src.ctypes.data, src.shape[0], dst.ctypes.data, dst.shape[0], C.byref(out), C.byref(n_out))
    if rc == 1:
        raise ValueError('src and dst should have the same size!')
    assert rc == 0
    try:
        n = n_out.value
        res = np.ctypeslib.as_array(out, shape=(max(n, 1), events.shape[1]))[:n].copy() if n else \
            np.zeros((0, events.shape[1]), np.float32)
    finally:
        free(out)
    return res


def ref_map(events, shape, src, dst):
    return _call('ref', events, shape, src, dst)


def port_map(events, shape, src, dst):
    return _call('port', events, shape, src, dst)
