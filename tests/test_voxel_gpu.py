"""GPU: the voxel-grid EXTENSION (csrc/voxel.cuh, `dv.compute_voxel_grid`) through the C ABI against
oracle/voxel_oracle.py.  The reference has no voxel grid: parity is UNPINNED, the oracle restates the
published definition (Zhu et al., CVPR 2019).  Bar: the device sums fp32 weights in no fixed order,
the oracle sums the same fp32 weights exactly; |got - want| <= ATOL + RTOL * sum|w| per voxel."""
import numpy as np
import pytest
import torch

import dvs_of_training_framework_b200 as dv
from oracle import voxel_oracle as vo
from tests.helpers import np_events, synth_events

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 1e-6


@pytest.fixture(params=['direct', 'binned'])
def path(request, monkeypatch):
    monkeypatch.setenv('DVSOF_ENCODE_PATH', request.param)
    return request.param


def _check(ev, ts, sidx, H, W, C):
    want, mag = vo.voxel_grid(np_events(ev), ts.numpy(), sidx.numpy(), (H, W), C, with_abs_sum=True)
    got = dv.compute_voxel_grid({k: torch.as_tensor(v).cuda() for k, v in ev.items()}, ts.cuda(), sidx.cuda(),
                                (H, W), C).cpu().numpy()
    assert got.shape == want.shape
    err = np.abs(got.astype(np.float64) - want)
    assert (err <= ATOL + RTOL * mag).all(), float((err - RTOL * mag).max())
    return got, want


@pytest.mark.parametrize('B,H,W,n,C,contended', [
    (1, 272, 352, 32768, 5, False),       # cfg1: DAVIS, 5 bins
    (8, 256, 256, 131072, 9, False),      # cfg2: MVSEC, 9 bins
    (8, 256, 256, 131072, 9, True),       # contended pixels (long sums)
    (2, 480, 640, 200000, 15, False),     # cfg4 shape: DSEC, 15 bins (one 150 KB slab per SM)
    (3, 48, 80, 1001, 2, False),          # n % 4 != 0 tail, two bins
    (2, 32, 32, 5000, 1, False),          # one bin: every weight 1 - f with t* = 0 -> plain polarity sum
])
def test_matches_oracle(B, H, W, n, C, contended, path):
    got, want = _check(*synth_events(5 + B + C, B, H, W, n, contended=contended), H, W, C)
    assert np.abs(want).max() > 0


def test_ragged_empty_samples_and_unaligned_views(path):
    ev, ts, sidx = synth_events(11, 6, 32, 48, 4000, ragged=True)
    _check(ev, ts, sidx, 32, 48, 7)
    ev, ts, sidx = synth_events(3, 2, 64, 64, 5000)
    want = vo.voxel_grid(np_events({k: v[1:] for k, v in ev.items()}), ts.numpy(), sidx.numpy(), (64, 64), 4)
    got = dv.compute_voxel_grid({k: v.cuda()[1:] for k, v in ev.items()}, ts.cuda(), sidx.cuda(), (64, 64), 4)
    np.testing.assert_allclose(got.cpu().numpy(), want, rtol=0, atol=1e-5)


def test_polarity_sum_and_known_answer(path):
    # summed over the bins every in-range event contributes exactly its polarity (partition of unity)
    ev, ts, sidx = synth_events(21, 4, 64, 96, 20000)
    got, _ = _check(ev, ts, sidx, 64, 96, 9)
    e = np_events(ev)
    net = np.zeros((4, 64, 96))
    np.add.at(net, (e['sample_index'], e['y'], e['x']), e['polarity'])
    np.testing.assert_allclose(got.sum(1), net, rtol=0, atol=1e-4)
    # hand-made: start 1.0, last event at 2.0 -> dt 1; C = 3: t* = 2 (t - 1)
    ev = {'x': [1, 1, 2, 2, 3], 'y': [0] * 5, 'polarity': [1, -1, 1, 1, -1], 'sample_index': [0] * 5,
          'timestamp': np.array([1.25, 1.5, 0.75, 2.25, 2.0], np.float32)}
    ev = {k: torch.as_tensor(np.asarray(v)).cuda() for k, v in ev.items()}
    out = dv.compute_voxel_grid(ev, torch.tensor([1.0, 2.0]).cuda(), torch.tensor([0, 0]).cuda(), (4, 8), 3)
    want = np.zeros((1, 3, 4, 8), np.float32)
    want[0, 0, 0, 1], want[0, 1, 0, 1] = 0.5, 0.5 - 1.0          # t* = 0.5 (+), t* = 1 (-)
    want[0, 0, 0, 2] = 0.5                                        # t* = -0.5: half of its tent is inside
    # t* = 2.5 at (2, 0): beyond the last bin by 0.5 -> 0.5 into bin 2;  t* = 2 (-) at (3, 0)
    want[0, 2, 0, 2], want[0, 2, 0, 3] = 0.5, -1.0
    assert np.array_equal(out.cpu().numpy(), want)


def test_empty_stream_and_argument_errors():
    ev = {k: torch.zeros(0, dtype=torch.long).cuda() for k in ('x', 'y', 'polarity', 'sample_index')}
    ev['timestamp'] = torch.zeros(0).cuda()
    out = dv.compute_voxel_grid(ev, torch.tensor([0., .04]).cuda(), torch.tensor([0, 0]).cuda(), (16, 16), 5)
    assert out.shape == (1, 5, 16, 16) and not out.any()
    with pytest.raises(ValueError, match='n_bins'):
        dv.compute_voxel_grid(ev, torch.tensor([0., .04]).cuda(), torch.tensor([0, 0]).cuda(), (16, 16), 65)


def test_bad_events_raise_like_the_event_image(path):
    ev, ts, sidx = synth_events(2, 2, 32, 32, 3000)
    ev = {k: v.cuda() for k, v in ev.items()}
    ev['polarity'][5] = 0
    with pytest.raises(AssertionError, match='polarity'):
        dv.compute_voxel_grid(ev, ts.cuda(), sidx.cuda(), (32, 32), 5)
        torch.cuda.synchronize()


def test_full_size_binned_path_against_direct():
    """cfg3 density on fewer samples (4 x 5 M events, 9 bins): the binned path is the default there;
    its result agrees with the direct path to the accumulation tolerance and keeps the polarity sums."""
    B, H, W, n, C = 4, 256, 256, 5_000_000, 9
    ev, ts, sidx = synth_events(1, B, H, W, n, device='cuda')
    lib = dv._lib.load()
    N = ev['x'].numel()
    assert lib.dvsof_voxel_path(N, B, H, W, C, lib.dvsof_voxel_workspace_bytes(N, B, H, W, C)) == 1
    binned = dv.compute_voxel_grid(ev, ts.cuda(), sidx.cuda(), (H, W), C)
    import os
    os.environ['DVSOF_ENCODE_PATH'] = 'direct'
    try:
        direct = dv.compute_voxel_grid(ev, ts.cuda(), sidx.cuda(), (H, W), C)
    finally:
        del os.environ['DVSOF_ENCODE_PATH']
    assert torch.allclose(binned, direct, rtol=0, atol=2e-4)      # ~76 events per pixel, |w| <= 1
    net = torch.zeros(B, H, W, device='cuda')
    net.view(-1).index_add_(0, (ev['sample_index'] * H + ev['y']) * W + ev['x'], ev['polarity'].float())
    assert torch.allclose(binned.sum(1), net, rtol=0, atol=1e-3)


def test_bin_count_limits(path):
    """64 bins (the maximum) and a slab too large for shared memory (640 wide x 64 bins: 160 KB per row
    pair... the plan refuses and the call takes the direct path even when the binned one is asked for)."""
    ev, ts, sidx = synth_events(31, 2, 24, 40, 6000)
    _check(ev, ts, sidx, 24, 40, 64)
    lib = dv._lib.load()
    assert lib.dvsof_voxel_path(10 ** 7, 2, 480, 1280, 64, lib.dvsof_voxel_workspace_bytes(10 ** 7, 2, 480, 1280, 64)) == 0
    ev, ts, sidx = synth_events(32, 1, 8, 1280, 4000)
    _check(ev, ts, sidx, 8, 1280, 64)
