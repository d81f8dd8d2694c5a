"""dvsof_encode_event_image_static (include/dvsof.h): the encoder's direct path with its scratch in a
static __device__ array -- no workspace argument, no allocation inside the call.  No counterpart in
the reference (added for callers that cannot cudaMalloc).  Checked through the C ABI against the
existing entry point dvsof_encode_event_image, bit-exact (the encoder's bar), on two calls in a row
that share the scratch array.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest
import torch

from dvs_of_training_framework_b200 import _lib
from tests.helpers import synth_events


def test_static_entry_rejects_more_samples_than_the_scratch_holds(lib):
    # host-side check, before any CUDA call: runs without a GPU
    cap = lib.dvsof_encode_static_max_samples()
    assert cap == 4096
    dummy = C.c_void_p(256)
    rc = lib.dvsof_encode_event_image_static(None, None, None, None, None, 0, None, cap + 1, 16, 16,
                                             dummy, dummy, None)
    assert rc == -3 and b'static scratch' in lib.dvsof_last_error()
    rc = lib.dvsof_encode_event_image_static(None, None, None, None, None, -1, None, 1, 16, 16,
                                             dummy, dummy, None)
    assert rc == -1


def _encode_both(seed, B, H, W, n, check_no_alloc):
    ev, ts, _ = synth_events(seed, B, H, W, n, ragged=(seed % 2 == 1))
    ev = {k: v.cuda() for k, v in ev.items()}
    ts = ts.cuda()
    N = ev['x'].numel()
    lib = _lib.load()
    status = torch.zeros(2, dtype=torch.int32, device='cuda')
    want = torch.full((B, 4, H, W), -1.0, device='cuda')
    ws = torch.empty(lib.dvsof_encode_workspace_bytes(0, B, H, W), dtype=torch.uint8, device='cuda')
    args = [_lib.ptr(ev[k]) for k in ('x', 'y', 'timestamp', 'polarity', 'sample_index')] + \
        [N, _lib.ptr(ts), B, H, W]
    stream = _lib.stream_ptr(torch.device('cuda', torch.cuda.current_device()))
    _lib.call('dvsof_encode_event_image', *args, _lib.ptr(want), _lib.ptr(ws), ws.numel(),
              _lib.ptr(status[:1]), stream)
    got = torch.full((B, 4, H, W), -2.0, device='cuda')
    torch.cuda.synchronize()
    torch_bytes, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    _lib.call('dvsof_encode_event_image_static', *args, _lib.ptr(got), _lib.ptr(status[1:]), stream)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == torch_bytes
    if check_no_alloc:              # the first call on a device may load the module lazily
        # a cudaMalloc shows up in whole 2 MiB pages; anything smaller is not ours
        assert free_before - torch.cuda.mem_get_info()[0] < (1 << 20), 'device memory shrank across the call'
    assert status.tolist() == [0, 0]
    assert want[:, :2].sum().item() == N        # the comparison is not between two empty images
    assert torch.equal(got, want)
    return got


@pytest.mark.gpu
def test_static_entry_matches_workspace_entry_on_two_calls_in_a_row():
    torch.cuda.set_device(0)
    first = _encode_both(8, 8, 256, 256, 131072, check_no_alloc=False)       # cfg2 shape
    # fewer samples, another image size, ragged with an empty sample: the scratch still holds the
    # bounds of the 8 samples above beyond what this call rewrites
    _encode_both(3, 3, 48, 80, 1001, check_no_alloc=True)
    again = _encode_both(8, 8, 256, 256, 131072, check_no_alloc=True)
    assert torch.equal(first, again)


@pytest.mark.gpu
def test_static_entry_empty_stream_gives_zeros():
    torch.cuda.set_device(0)
    out = torch.full((2, 4, 16, 16), 7.0, device='cuda')
    status = torch.zeros(1, dtype=torch.int32, device='cuda')
    ts = torch.tensor([0., .04, 0., .04], device='cuda')
    null = C.c_void_p(0)
    _lib.call('dvsof_encode_event_image_static', null, null, null, null, null, 0, _lib.ptr(ts), 2, 16, 16,
              _lib.ptr(out), _lib.ptr(status), _lib.stream_ptr(torch.device('cuda', 0)))
    torch.cuda.synchronize()
    assert not out.any() and int(status) == 0


@pytest.mark.gpu
def test_static_entry_at_its_sample_limit():
    torch.cuda.set_device(0)
    _encode_both(12, 4096, 8, 16, 50, check_no_alloc=True)
