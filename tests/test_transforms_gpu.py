"""GPU parity (bit-exact) for count image, EventCrop, map_events / RandomRotation."""
import numpy as np
import pytest
import torch

import dvs_of_training_framework_b200 as dv
from oracle import encoder_oracle as eo

pytestmark = pytest.mark.gpu


def test_golden_crop_count_rotation(golden):
    g = np.load(golden / 'transforms.npz')
    box = g['box'].tolist()
    cropped = dv.EventCrop(box)(g['events'].copy())
    assert cropped.dtype == np.float64 and np.array_equal(cropped, g['cropped'])
    cnt = dv.get_count_image([cropped[:, 0], cropped[:, 1]], (box[2], box[3]))
    assert cnt.dtype == np.uint64 and np.array_equal(cnt, g['count'])
    e5 = np.hstack([g['events'], np.zeros((g['events'].shape[0], 1))])
    rot = dv.RandomRotation(30, (260, 346))
    for ang in ('0.0', '90.0', '17.5'):
        rim, rev, a = rot(g['images'], e5.copy(), angle=float(ang))
        assert a == float(ang)
        assert np.array_equal(rim, g[f'rot_images_{ang}'])
        assert rev.dtype == np.float32 and np.array_equal(rev, g[f'rot_events_{ang}'])


def test_map_identity_exact():
    # tests/dataset/test_dataset.py:42-73: angle 0 == identity, exact equality
    H, W = 260, 346
    rng = np.random.default_rng(1)
    n = 20000
    ev = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), np.sort(rng.random(n)),
                   rng.choice([-1, 1], n), np.zeros(n)], 1).astype(np.float32)
    idx = np.arange(H * W, dtype=np.uint64)
    out = dv.transformation.map(ev, (H, W), idx, idx)
    assert np.array_equal(out, ev)


def test_map_fanout_order_and_drops():
    H, W = 40, 50
    rng = np.random.default_rng(2)
    M = 6000
    src = rng.integers(0, H * W // 2, M).astype(np.uint64)   # many duplicates, half unmapped
    dst = rng.integers(0, H * W, M).astype(np.uint64)
    n = 30000
    ev = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), rng.random(n),
                   rng.choice([-1, 1], n), rng.integers(0, 3, n)], 1).astype(np.float32)
    want = eo.map_events(ev, (H, W), src, dst)
    got = dv.transformation.map(ev, (H, W), src, dst)
    assert got.shape == want.shape and np.array_equal(got, want)


def test_map_argument_errors():
    ev = np.zeros((4, 5), np.float32)
    idx = np.arange(10, dtype=np.uint64)
    with pytest.raises(ValueError):
        dv.transformation.map(ev, (4, 4), idx, idx[:5])
    with pytest.raises(TypeError):
        dv.transformation.map(ev.astype(np.float64), (4, 4), idx, idx)
    with pytest.raises(TypeError):
        dv.transformation.map(ev, (4, 4), idx.astype(np.int64), idx)
    assert dv.transformation.map(np.zeros((0, 5), np.float32), (4, 4), idx, idx).shape == (0, 5)


@pytest.mark.parametrize('dtype', [np.float32, np.float64])
def test_crop_random_boxes(dtype):
    rng = np.random.default_rng(3)
    n = 100003
    ev = np.stack([rng.integers(0, 346, n), rng.integers(0, 260, n), np.sort(rng.random(n)),
                   rng.choice([-1, 1], n)], 1).astype(dtype)
    for box in ([0, 0, 260, 346], [2, 45, 256, 256], [100, 100, 1, 1], [0, 0, 0, 0]):
        want = eo.event_crop(ev, box)
        got = dv.EventCrop(box)(ev.copy())
        assert got.dtype == dtype and np.array_equal(got, want)
    assert dv.EventCrop([0, 0, 5, 5])(np.zeros((0, 4), dtype)).shape == (0, 4)


def test_count_image_large_and_bad_pixel():
    rng = np.random.default_rng(4)
    n = 3_000_000
    x, y = rng.integers(0, 640, n), rng.integers(0, 480, n)
    want = eo.get_count_image([x, y], (480, 640))
    got = dv.get_count_image([x, y], (480, 640))
    assert np.array_equal(got, want) and int(got.sum()) == n
    with pytest.raises(ValueError):
        dv.get_count_image([np.array([640]), np.array([0])], (480, 640))
    # CUDA tensors in -> CUDA tensor out
    out = dv.get_count_image([torch.tensor(x).cuda(), torch.tensor(y).cuda()], (480, 640))
    assert out.is_cuda and np.array_equal(out.cpu().numpy().astype(np.uint64), want)


@pytest.mark.parametrize('dtype', [np.float32, np.float64])
def test_frame_generator_matches_reference_slicing(dtype):
    """utils/data.py:139-152 (used by utils/testing.py:64): searchsorted(side='right') windows,
    including duplicates of a boundary value, empty windows and windows past the last event."""
    from dvs_of_training_framework_b200 import data as dd
    rng = np.random.default_rng(4)
    n = 5000
    t = np.sort(rng.uniform(0, 1, n)).astype(dtype)
    t[100:110] = t[100]                                    # ties on a boundary
    ev = [rng.integers(0, 346, n), rng.integers(0, 260, n), t, rng.choice([-1, 1], n)]
    frames = [(0.0, 0.1), (float(t[100]), 0.5), (0.5, 0.5), (0.9, 2.0), (-1.0, 0.0)]
    want = np.searchsorted(t, np.array(frames).ravel(), side='right').reshape(-1, 2)
    got = dd.frame_offsets(t, frames).cpu().numpy()
    assert np.array_equal(got, want)
    for (e, start, stop), (a, b), (fs, fe) in zip(dd.frame_generator(ev, frames), want, frames):
        assert start == fs and stop == fe
        for comp, ref in zip(e, ev):
            assert np.array_equal(comp, ref[a:b])
    assert dd.frame_offsets(t[:0], frames).sum() == 0


# ------------------------------------------------------------------------------------------
# fused augmentation (flip + rotation LUT + crop in one pass), utils/dataset.py:753-769
# ------------------------------------------------------------------------------------------
def _reference_sequence(ev, shape, flip, angle, box):
    """The three reference steps one after the other, on the oracle (numpy)."""
    H, W = shape
    ev = ev.copy()
    if flip:
        ev[:, 0] = W - ev[:, 0] - 1                                   # dataset.py:757
    if angle is not None:
        src, dst = eo.rotation_lut(shape, angle)
        ev = eo.map_events(ev, shape, src, dst)
    if box is not None:
        ev = eo.event_crop(ev, box)
    return ev


@pytest.mark.parametrize('flip,angle,box', [
    (False, None, None), (True, None, None), (False, 17.5, None), (False, None, [2, 45, 256, 256]),
    (True, -23.0, [2, 45, 256, 256]), (True, 90.0, [0, 0, 260, 346]), (False, 5.0, [100, 100, 1, 1]),
])
def test_augment_events_equals_reference_sequence(flip, angle, box):
    H, W = 260, 346
    rng = np.random.default_rng(11)
    n = 50001
    ev = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), np.sort(rng.random(n)),
                   rng.choice([-1, 1], n), rng.integers(0, 2, n)], 1).astype(np.float32)
    want = _reference_sequence(ev, (H, W), flip, angle, box)
    lut = dv.rotation_lut((H, W), angle) if angle is not None else None
    got = dv.augment_events(ev, (H, W), is_flip=flip, lut=lut, box=box)
    assert got.dtype == np.float32 and got.shape == want.shape and np.array_equal(got, want)
    # and the same through the separate device ops (each pinned against the reference goldens)
    seq = torch.tensor(ev).cuda()
    if flip:
        seq[:, 0] = W - seq[:, 0] - 1
    if lut is not None:
        seq = lut.apply(seq.contiguous())
    if box is not None:
        seq = dv.EventCrop(box)(seq)
    assert np.array_equal(seq.cpu().numpy(), got)


def test_augment_events_golden_rotation_then_crop(golden):
    """Reference outputs (tests/golden/transforms.npz: RandomRotation of utils/data.py run
    unmodified) followed by the reference's EventCrop semantics."""
    g = np.load(golden / 'transforms.npz')
    e5 = np.hstack([g['events'], np.zeros((g['events'].shape[0], 1))]).astype(np.float32)
    box = g['box'].tolist()
    for ang in ('0.0', '90.0', '17.5'):
        want = eo.event_crop(g[f'rot_events_{ang}'], box)
        got = dv.augment_events(e5, (260, 346), lut=dv.rotation_lut((260, 346), float(ang)), box=box)
        assert np.array_equal(got, want)
    assert dv.augment_events(np.zeros((0, 5), np.float32), (260, 346), is_flip=True, box=box).shape == (0, 5)


@pytest.mark.parametrize('kind', ['ref', 'port'])
def test_map_equals_native_reference_source(kind):
    """A6 against the native component itself: the CUDA remap == the reference's own
    transformation.cpp:35-85 compiled from the reference tree (oracle/_ref/libmap_events_ref.so,
    built by oracle/Makefile where /root/reference exists and shipped prebuilt) and == the plain
    C++ restatement, on random LUTs with fan-out > 1, dropped pixels, fractional coordinates and
    the RandomRotation LUT at DAVIS size."""
    from oracle import map_events_native as mn
    if kind == 'port':
        mn.build()
    elif not mn.available('ref'):
        pytest.skip('oracle/_ref/libmap_events_ref.so not present')
    fn = mn.ref_map if kind == 'ref' else mn.port_map
    rng = np.random.default_rng(11)
    for H, W, n, M, E, frac in [(40, 50, 30000, 6000, 5, False), (260, 346, 100000, 90000, 5, False),
                                (33, 47, 4000, 3000, 6, True), (16, 16, 300, 100, 4, True)]:
        cols = [rng.integers(0, W, n), rng.integers(0, H, n), np.sort(rng.random(n)), rng.choice([-1, 1], n)]
        cols += [rng.integers(0, 4, n) for _ in range(E - 4)]
        ev = np.stack(cols[:E], 1).astype(np.float32)
        if frac:
            ev[:, :2] += rng.random((n, 2)).astype(np.float32) * 0.999
        src = rng.integers(0, H * W // 2, M).astype(np.uint64)
        dst = rng.integers(0, H * W, M).astype(np.uint64)
        want = fn(ev, (H, W), src, dst)
        got = dv.transformation.map(ev, (H, W), src, dst)
        assert got.shape == want.shape and np.array_equal(got, want)
    H, W = 260, 346
    src, dst = eo.rotation_lut((H, W), 23.0)
    ev = np.stack([rng.integers(0, W, 50000), rng.integers(0, H, 50000), np.sort(rng.random(50000)),
                   rng.choice([-1, 1], 50000), np.zeros(50000)], 1).astype(np.float32)
    assert np.array_equal(dv.transformation.map(ev, (H, W), src, dst), fn(ev, (H, W), src, dst))
