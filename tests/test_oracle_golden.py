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
