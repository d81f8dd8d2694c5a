"""CPU: the oracle (oracle/*.py) against the golden vectors produced by the reference's own code
(tests/golden/make_golden.py) and against the reference's known-answer tests."""
import numpy as np
import pytest
import torch

from oracle import encoder_oracle as eo
from oracle import loss_oracle as lo


@pytest.mark.parametrize('name', ['fixture', 'dense', 'ragged'])
def test_encoder_oracle_matches_reference_output(golden, name):
    g = np.load(golden / f'encoder_{name}.npz')
    ev = {k[3:]: g[k] for k in g.files if k.startswith('ev_')}
    out = eo.compute_event_image(ev, g['timestamps'], g['sample_idx'], (int(g['H']), int(g['W'])))
    assert out.dtype == np.float32
    assert np.array_equal(out, g['out'])          # bit-exact


def test_encoder_oracle_fixture_facts(golden):
    # SURVEY 8c probe: fixtures 0+1 -> 255 positive + 105 negative events, maxima [3,2,1,1]
    g = np.load(golden / 'encoder_fixture.npz')
    out = g['out']
    assert out.shape == (2, 4, 272, 352)
    assert out[:, 0].sum() == 255 and out[:, 1].sum() == 105
    assert out.max(axis=(0, 2, 3)).tolist() == [3, 2, 1, 1]


def test_encoder_oracle_empty():
    ev = {k: np.zeros(0, np.int64) for k in ('x', 'y', 'polarity', 'sample_index')}
    ev['timestamp'] = np.zeros(0, np.float32)
    out = eo.compute_event_image(ev, np.array([0, 0.04], np.float32), np.array([0, 0]), (16, 32))
    assert out.shape == (1, 4, 16, 32) and not out.any()


def test_encoder_oracle_rejects_bad_input():
    ev = {'x': np.array([1, 2]), 'y': np.array([1, 2]), 'polarity': np.array([1, 0]),
          'sample_index': np.array([0, 0]), 'timestamp': np.array([0.1, 0.2], np.float32)}
    with pytest.raises(AssertionError):
        eo.compute_event_image(ev, np.array([0, 1], np.float32), np.array([0, 0]), (4, 4))
    ev['polarity'] = np.array([1, -1])
    ev['sample_index'] = np.array([1, 0])
    with pytest.raises(AssertionError):
        eo.compute_event_image(ev, np.array([0, 1, 0, 1], np.float32), np.array([0, 0, 1, 1]), (4, 4))


def _fixture_images(golden, H, W):
    f = np.load(golden / 'fixture_000001.npz')
    images = torch.tensor(np.stack([f['image1'][:H, :W], f['image2'][:H, :W]])[:, None]).float()
    ts = torch.tensor([0, float(f['stop']) - float(f['start'])], dtype=torch.float32)
    return f, images, ts


def test_loss_oracle_reference_kat_no_changes():
    # tests/loss/test_loss.py:8-22
    H, W = 5, 6
    images = torch.zeros(2, 1, H, W)
    ts = torch.tensor([0, 0.4])
    out = lo.multiscale_loss([torch.zeros(1, 2, H, W)], ts.view(1, 2), torch.LongTensor([0]),
                             images, ts, torch.LongTensor([0, 0]))
    for term, gt in zip(out, [0.002, 0.002, 0]):
        assert len(term) == 1 and abs(float(term[0]) - gt) < 5e-6


def test_loss_oracle_reference_kat_fixture(golden):
    # tests/loss/test_loss.py:25-65 (goldens 0.622660; 0.002120 / 0.652659 / 0.007802)
    H, W = 246, 340
    f, images, ts = _fixture_images(golden, H, W)
    sidx, fs = torch.LongTensor([0, 0]), torch.LongTensor([0])
    out = lo.multiscale_loss([torch.zeros(1, 2, H, W)], ts.view(1, 2), fs, images, ts, sidx)
    for term, gt in zip(out, [0.002, 0.622660, 0]):
        assert abs(float(term[0]) - gt) < 5e-6
    flow = torch.tensor(f['flow'][:H, :W]).permute(2, 0, 1)[None].contiguous()
    out = lo.multiscale_loss([flow], ts.view(1, 2), fs, images, ts, sidx)
    for term, gt in zip(out, [0.002120, 0.652659, 0.007802]):
        assert abs(float(term[0]) - gt) < 5e-6
    kat = np.load(golden / 'loss_kat.npz')
    np.testing.assert_allclose([float(t[0]) for t in out], kat['pred'], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize('tag', ['a', 'b', 'c'])
def test_loss_oracle_matches_reference_values_and_grads(golden, tag):
    g = np.load(golden / f'loss_{tag}.npz')
    flows = [torch.tensor(g[f'flow{k}']).requires_grad_() for k in range(4)]
    terms = lo.multiscale_loss(flows, torch.tensor(g['flow_ts']), torch.tensor(g['flow_sample_idx']),
                               torch.tensor(g['images']).float(), torch.tensor(g['timestamps']),
                               torch.tensor(g['sample_idx']))
    got = np.array([[float(v) for v in t] for t in terms])
    np.testing.assert_allclose(got, g['terms'], rtol=1e-5, atol=1e-6)
    lo.combined(terms).backward()
    for k, f in enumerate(flows):
        np.testing.assert_allclose(f.grad.numpy(), g[f'grad{k}'], rtol=1e-5, atol=1e-6)


def test_transforms_oracle_match_reference(golden):
    g = np.load(golden / 'transforms.npz')
    box = g['box'].tolist()
    cropped = eo.event_crop(g['events'], box)
    assert np.array_equal(cropped, g['cropped'])
    cnt = eo.get_count_image([cropped[:, 0], cropped[:, 1]], (box[2], box[3]))
    assert cnt.dtype == np.uint64 and np.array_equal(cnt, g['count'])
    e5 = np.hstack([g['events'], np.zeros((g['events'].shape[0], 1))])
    for ang in ('0.0', '90.0', '17.5'):
        rim, rev = eo.rotate(g['images'], e5, (260, 346), float(ang))
        assert np.array_equal(rim, g[f'rot_images_{ang}'])
        assert np.array_equal(rev, g[f'rot_events_{ang}'])


def test_map_events_identity_and_fanout():
    # tests/dataset/test_dataset.py:42-73 use angle 0 == identity mapping, exact equality
    H, W = 6, 7
    rng = np.random.default_rng(0)
    ev = np.stack([rng.integers(0, W, 50), rng.integers(0, H, 50), rng.random(50),
                   rng.choice([-1, 1], 50), np.zeros(50)], 1).astype(np.float32)
    idx = np.arange(H * W, dtype=np.uint64)
    assert np.array_equal(eo.map_events(ev, (H, W), idx, idx), ev)
    # one-to-many + dropped pixels, LUT order preserved
    src = np.array([3, 3, 5, 3], np.uint64)
    dst = np.array([10, 2, 7, 40], np.uint64)
    e = np.array([[3, 0, .5, 1, 0], [4, 0, .6, -1, 0], [5, 0, .7, 1, 0]], np.float32)
    out = eo.map_events(e, (H, W), src, dst)
    assert out[:, 0].tolist() == [3, 2, 5, 0] and out[:, 1].tolist() == [1, 0, 5, 1]
    assert out[:, 2].tolist() == [.5, .5, .5, np.float32(.7)]
    with pytest.raises(ValueError):
        eo.map_events(e, (H, W), src, dst[:2])


def test_decode_events_matches_reference_layout():
    # utils/dataset.py:335-373 semantics on the golden dict of tests/dataset/test_encoding.py style
    out = eo.decode_events(np.array([1, 2, 3, 4, 5], np.int16), np.array([5, 4, 3, 2, 1], np.int16),
                           np.array([1, 0, 1, 1, 0], bool), np.arange(5, dtype=np.float32),
                           events_per_element=[2, 0, 3], elements_per_sample=[2, 1])
    assert out['polarity'].tolist() == [1, -1, 1, 1, -1]
    assert out['element_index'].tolist() == [0, 0, 0, 0, 0]
    assert out['sample_index'].tolist() == [0, 0, 1, 1, 1]
    assert out['x'].dtype == np.int64


def test_torch_baseline_port_equals_numpy_oracle():
    from oracle import encoder_oracle_torch as eot
    from tests.helpers import np_events, synth_events
    ev, ts, sidx = synth_events(17, 3, 40, 56, 5000, ragged=True)
    a = eot.compute_event_image(ev, ts, sidx, (40, 56)).numpy()
    b = eo.compute_event_image(np_events(ev), ts.numpy(), sidx.numpy(), (40, 56))
    assert np.array_equal(a, b)


def test_aten_bilinear_operation_order():
    """Pins the arithmetic the CUDA warp kernel copies (csrc/loss.cu warp_sample): ATen's CPU
    grid_sample (the reference's F.grid_sample, utils/loss.py:70) evaluates
    ix = (g + 1) * ((W-1)/2), weights (1-wy)(1-wx).., and the tap sum as mul + 3 chained FMAs.
    Emulated in numpy (fp32 fma through float64) this must reproduce torch bit for bit."""
    import torch.nn.functional as F
    torch.manual_seed(0)
    H, W = 64, 80
    img = torch.randint(0, 256, (1, 1, H, W)).float()
    flow = torch.randn(1, 2, H, W) * 3
    grid = lo.normalised_grid(flow)
    out = F.grid_sample(img, grid.permute(0, 2, 3, 1), mode='bilinear', padding_mode='zeros',
                        align_corners=True)[0, 0].numpy()
    f32 = np.float32
    gx, gy = grid[0, 0].numpy(), grid[0, 1].numpy()
    ix, iy = (gx + f32(1)) * f32((W - 1) / 2.), (gy + f32(1)) * f32((H - 1) / 2.)
    xw, yn = np.floor(ix), np.floor(iy)
    wx, wy = ix - xw, iy - yn
    ex, ey = f32(1) - wx, f32(1) - wy
    im = img[0, 0].numpy()

    def tap(yy, xx):
        ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
        return np.where(ok, im[np.clip(yy, 0, H - 1).astype(int), np.clip(xx, 0, W - 1).astype(int)],
                        f32(0)).astype(f32)

    def fma(a, b, c):
        return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)

    nw, ne, sw, se = tap(yn, xw), tap(yn, xw + 1), tap(yn + 1, xw), tap(yn + 1, xw + 1)
    val = fma(se, wy * wx, fma(sw, wy * ex, fma(ne, ey * wx, nw * (ey * ex))))
    assert np.array_equal(val, out)


def test_aten_upsample_operation_order():
    """Pins the arithmetic csrc/loss.cu resize_bilinear_kernel copies for the UPSAMPLING levels of
    the chained pyramid (utils/loss.py:210): src index = i * (in-1)/(out-1) in fp32, lambda =
    index - floor, and out = fma(hy, fma(hx, a, lx*b), ly * fma(hx, c, lx*e)).  (The first,
    downsampling level of the chain is not reproducible bit for bit from outside ATen; it agrees to
    1 ulp, which the GPU gradient tests account for through the reference's own sensitivity.)"""
    import torch.nn.functional as F
    f32 = np.float32

    def fma(a, b, c):
        return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)

    torch.manual_seed(1)
    img = F.interpolate(torch.randint(0, 256, (1, 1, 260, 346)).float(), size=(32, 43), mode='bilinear',
                        align_corners=True)
    for ho, wo in ((65, 86), (130, 173), (260, 346)):
        out = F.interpolate(img, size=(ho, wo), mode='bilinear', align_corners=True)
        im = img[0, 0].numpy()
        hi, wi = im.shape
        ys = np.arange(ho, dtype=f32) * (f32(hi - 1) / f32(ho - 1))
        xs = np.arange(wo, dtype=f32) * (f32(wi - 1) / f32(wo - 1))
        y0, x0 = np.minimum(ys.astype(np.int64), hi - 1), np.minimum(xs.astype(np.int64), wi - 1)
        y1, x1 = np.minimum(y0 + 1, hi - 1), np.minimum(x0 + 1, wi - 1)
        ly, lx = (ys - y0.astype(f32))[:, None], (xs - x0.astype(f32))[None, :]
        a, b, c, e = im[y0][:, x0], im[y0][:, x1], im[y1][:, x0], im[y1][:, x1]
        hx, hy = (f32(1) - lx) + 0 * a, (f32(1) - ly) + 0 * a
        lx, ly = lx + 0 * a, ly + 0 * a
        val = fma(hy, fma(hx, a, lx * b), ly * fma(hx, c, lx * e))
        assert np.array_equal(val, out[0, 0].numpy())
        img = out


def test_loss_oracle_matches_reference_dummynet_cfg1(golden):
    """BASELINE configs[0]: the reference's DummyNet (zero flow, 4 scales) through its own
    init_losses + process_minibatch at 260x346, batch 1 (tests/golden/make_golden_dummynet.py)."""
    g = np.load(golden / 'dummynet_cfg1.npz')
    flows = [torch.zeros(1, 2, int(h), int(w)) for h, w in g['shapes']]
    terms = lo.multiscale_loss(flows, torch.tensor(g['flow_ts']), torch.tensor(g['flow_sample_idx']),
                               torch.tensor(g['images']).float(), torch.tensor(g['timestamps']),
                               torch.tensor(g['sample_idx']))
    got = np.array([[float(v) for v in t] for t in terms])
    np.testing.assert_allclose(got, g['terms'], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(float(lo.combined(terms)), float(g['loss']), rtol=1e-6, atol=1e-7)


def _random_map_case(seed, H, W, n, M, E=5, frac=False):
    rng = np.random.default_rng(seed)
    cols = [rng.integers(0, W, n), rng.integers(0, H, n), np.sort(rng.random(n)), rng.choice([-1, 1], n)]
    cols += [rng.integers(0, 4, n) for _ in range(E - 4)]
    ev = np.stack(cols[:E], 1).astype(np.float32)
    if frac:                                       # float -> int truncation (transformation.cpp:62,74)
        ev[:, :2] += rng.random((n, 2)).astype(np.float32) * 0.999
    src = rng.integers(0, max(1, H * W // 2), M).astype(np.uint64)   # fan-out > 1, half the pixels dropped
    dst = rng.integers(0, H * W, M).astype(np.uint64)
    return ev, src, dst


@pytest.mark.parametrize('kind', ['port', 'ref'])
def test_map_events_numpy_oracle_equals_native(kind):
    """A6: numpy oracle == plain-C++ restatement (oracle/map_events_ref.cpp) == the reference's own
    transformation.cpp:35-85 compiled from /root/reference (oracle/_ref, prebuilt where the tree
    exists), on LUTs with fan-out > 1, dropped pixels, fractional coordinates, E = 4/5/6."""
    from oracle import map_events_native as mn
    if kind == 'port':
        mn.build()
    elif not mn.available('ref'):
        pytest.skip('oracle/_ref/libmap_events_ref.so not built (needs /root/reference once)')
    fn = mn.port_map if kind == 'port' else mn.ref_map
    for seed, (H, W, n, M, E, frac) in enumerate([(20, 30, 500, 700, 5, False), (260, 346, 20000, 90000, 5, False),
                                                  (16, 16, 300, 100, 4, True), (33, 47, 4000, 3000, 6, True),
                                                  (8, 8, 0, 10, 5, False), (8, 8, 50, 0, 5, False)]):
        ev, src, dst = _random_map_case(seed, H, W, n, M, E, frac)
        want = eo.map_events(ev, (H, W), src, dst)
        got = fn(ev, (H, W), src, dst)
        assert got.dtype == np.float32 and got.shape == want.shape and np.array_equal(got, want), seed
    # identity (tests/dataset/test_dataset.py:42-73) and the 90-degree LUT of RandomRotation (:122-170)
    H, W = 260, 346
    ev, _, _ = _random_map_case(9, H, W, 5000, 1)
    idx = np.arange(H * W, dtype=np.uint64)
    assert np.array_equal(fn(ev, (H, W), idx, idx), ev)
    src, dst = eo.rotation_lut((H, W), 90.0)
    assert np.array_equal(fn(ev, (H, W), src.astype(np.uint64), dst.astype(np.uint64)),
                          eo.map_events(ev, (H, W), src, dst))
    with pytest.raises(ValueError):
        fn(ev, (H, W), idx, idx[:5])
