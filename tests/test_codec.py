"""Batch codec: utils/dataset.py encode_batch / decode_batch.

CPU: the numpy oracle against vectors produced by the reference's own functions
(tests/golden/codec.npz, made by tests/golden/make_golden_codec.py; `fixture` is the batch of
tests/dataset/test_encoding.py:21-80).  GPU: the device codec against the same vectors and against
the oracle on larger seeded batches, bit for bit."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import encoder_oracle as eo

G = np.load(Path(__file__).parent / 'golden' / 'codec.npz')
CASES = ['fixture', 'ragged', 'droptail', 'single', 'bulk']
DEC_KEYS = ['x', 'y', 'polarity', 'element_index', 'sample_index']


def _case(name):
    dec = {k: G[f'{name}/dec/{k}'] for k in DEC_KEYS + ['timestamp']}
    enc = {k: G[f'{name}/enc/{k}'] for k in ['x', 'y', 'timestamp', 'polarity', 'events_per_element']}
    back = {k: G[f'{name}/back/{k}'] for k in DEC_KEYS + ['timestamp']}
    return dec, enc, G[f'{name}/enc/elements_per_sample'], back, G[f'{name}/sample_idx'], int(G[f'{name}/size'])


def test_fixture_is_the_reference_test_vector():
    # tests/dataset/test_encoding.py:57-67
    _, enc, eps, _, _, _ = _case('fixture')
    assert enc['events_per_element'].tolist() == [1, 2, 1, 1, 0, 1, 1]
    assert eps.tolist() == [2, 1, 4] and eps.dtype == np.uint8
    assert enc['polarity'].tolist() == [False, True, False, True, True, True, False]
    assert enc['x'].dtype == np.int16


@pytest.mark.parametrize('name', CASES)
def test_oracle_matches_reference(name):
    dec, enc, eps, back, sample_idx, size = _case(name)
    assert np.array_equal(eo.elements_per_sample(sample_idx, size), eps)
    got = eo.encode_events(dec['x'], dec['y'], dec['polarity'], dec['element_index'], dec['sample_index'], eps)
    for k in ('x', 'y', 'polarity', 'events_per_element'):
        assert got[k].dtype == enc[k].dtype and np.array_equal(got[k], enc[k]), k
    out = eo.decode_events(enc['x'], enc['y'], enc['polarity'], enc['timestamp'], enc['events_per_element'], eps)
    for k in DEC_KEYS:
        assert out[k].dtype == np.int64 and np.array_equal(out[k], back[k]), k
    for k in DEC_KEYS:                        # the round trip is the identity on these batches
        assert np.array_equal(back[k], dec[k]), k


def _cuda(d):
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in d.items()}


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_device_codec_matches_reference(name):
    import dvs_of_training_framework_b200 as dv
    dec, enc, eps, back, sample_idx, size = _case(name)
    n_img = sample_idx.size
    e = dv.encode_batch(_cuda(dec), torch.zeros(n_img).cuda(), torch.from_numpy(sample_idx).cuda(),
                        torch.zeros(n_img, 1, 2, 2).cuda(), {}, size)
    assert e['elements_per_sample'].dtype == torch.uint8
    assert e['elements_per_sample'].cpu().numpy().tolist() == eps.tolist()
    assert e['images'].dtype == torch.uint8
    for k in ('x', 'y', 'polarity', 'events_per_element'):
        got = e['events'][k].cpu().numpy()
        assert got.dtype == enc[k].dtype and np.array_equal(got, enc[k]), k
    assert e['events']['timestamp'].data_ptr() == e['events']['timestamp'].data_ptr()
    d = dv.decode_batch(e)
    assert d['size'] == size and np.array_equal(d['sample_idx'].cpu().numpy(), G[f'{name}/back/sample_idx'])
    assert d['images'].dtype == torch.float32 and d['timestamps'].dtype == torch.float32
    for k in DEC_KEYS:
        got = d['events'][k].cpu().numpy()
        assert got.dtype == np.int64 and np.array_equal(got, back[k]), k
    assert np.array_equal(d['events']['timestamp'].cpu().numpy(), dec['timestamp'])


def _big(seed, n_samples, n_events):
    g = np.random.default_rng(seed)
    eps = g.integers(1, 6, n_samples)
    E = int(eps.sum())
    epe = g.multinomial(n_events, g.dirichlet(np.ones(E) * 0.3))
    epe[g.random(E) < 0.2] = 0
    epe[-1] = max(epe[-1], 1)
    n = int(epe.sum())
    enc = {'x': g.integers(-5, 700, n).astype(np.int16), 'y': g.integers(0, 500, n).astype(np.int16),
           'timestamp': g.random(n).astype(np.float32), 'polarity': g.random(n) < 0.7,
           'events_per_element': epe.astype(np.int64)}
    return enc, eps.astype(np.uint8)


@pytest.mark.gpu
@pytest.mark.parametrize('seed,n_samples,n_events', [(1, 8, 100_003), (2, 3, 4099), (3, 32, 1_000_001)])
def test_device_decode_encode_against_oracle(seed, n_samples, n_events):
    import dvs_of_training_framework_b200 as dv
    enc, eps = _big(seed, n_samples, n_events)
    want = eo.decode_events(enc['x'], enc['y'], enc['polarity'], enc['timestamp'], enc['events_per_element'], eps)
    got = dv.decode_events(_cuda(enc), torch.from_numpy(eps).cuda())
    for k in DEC_KEYS:
        assert np.array_equal(got[k].cpu().numpy(), want[k]), k
    # and back: the device encoder on the decoded form reproduces the storage form
    sample_idx = np.repeat(np.arange(n_samples), eps.astype(np.int64) + 1)
    e = dv.encode_batch(got, torch.zeros(sample_idx.size).cuda(), torch.from_numpy(sample_idx).cuda(),
                        torch.zeros(sample_idx.size, 1, 1, 1).cuda(), {}, n_samples)
    ref = eo.encode_events(want['x'], want['y'], want['polarity'], want['element_index'], want['sample_index'], eps)
    for k in ('x', 'y', 'polarity', 'events_per_element'):
        assert np.array_equal(e['events'][k].cpu().numpy(), ref[k]), k


@pytest.mark.gpu
def test_device_codec_unaligned_views_and_unsorted_histogram():
    import dvs_of_training_framework_b200 as dv
    enc, eps = _big(7, 4, 5003)
    n = enc['x'].size
    pad = {k: torch.from_numpy(np.concatenate([v[:1], v])).cuda()[1:] for k, v in enc.items() if k != 'events_per_element'}
    pad['events_per_element'] = torch.from_numpy(enc['events_per_element']).cuda()
    assert pad['x'].data_ptr() % 16 != 0
    want = eo.decode_events(enc['x'], enc['y'], enc['polarity'], enc['timestamp'], enc['events_per_element'], eps)
    got = dv.decode_events(pad, torch.from_numpy(eps).cuda())
    for k in DEC_KEYS:
        assert np.array_equal(got[k].cpu().numpy(), want[k]), k
    # np.add.at semantics: the histogram does not need grouped events
    perm = np.random.default_rng(0).permutation(n)
    perm[-1], perm[np.argmax(perm == n - 1)] = n - 1, perm[-1]          # keep the last event last (:298)
    shuf = {k: torch.from_numpy(want[k][perm]).cuda() for k in DEC_KEYS}
    shuf['timestamp'] = torch.from_numpy(enc['timestamp'][perm]).cuda()
    sample_idx = np.repeat(np.arange(4), eps.astype(np.int64) + 1)
    e = dv.encode_batch(shuf, torch.zeros(sample_idx.size).cuda(), torch.from_numpy(sample_idx).cuda(),
                        torch.zeros(sample_idx.size, 1, 1, 1).cuda(), {}, 4)
    ref = eo.encode_events(want['x'][perm], want['y'][perm], want['polarity'][perm], want['element_index'][perm],
                           want['sample_index'][perm], eps)
    for k in ('x', 'y', 'polarity', 'events_per_element'):
        assert np.array_equal(e['events'][k].cpu().numpy(), ref[k]), k


@pytest.mark.gpu
def test_device_codec_flags_inconsistent_counts():
    import dvs_of_training_framework_b200 as dv
    enc, eps = _big(9, 2, 1000)
    enc['events_per_element'] = enc['events_per_element'].copy()
    enc['events_per_element'][0] += 1
    with pytest.raises(AssertionError):
        dv.decode_events(_cuda(enc), torch.from_numpy(eps).cuda())
    with pytest.raises(RuntimeError):
        dv.decode_events({k: torch.from_numpy(v) for k, v in enc.items()}, torch.from_numpy(eps))


@pytest.mark.gpu
def test_compact_encoder_equals_decode_then_encode_image():
    """The fused path (storage form straight into the event image) equals decode_batch followed
    by the int64 encoder."""
    import dvs_of_training_framework_b200 as dv
    g = np.random.default_rng(5)
    B, H, W = 4, 64, 96
    eps = np.ones(B, dtype=np.uint8)
    epe = g.integers(1000, 3000, B).astype(np.int64)
    n = int(epe.sum())
    t = np.concatenate([np.sort(g.random(k)).astype(np.float32) for k in epe])
    enc = {'x': g.integers(0, W, n).astype(np.int16), 'y': g.integers(0, H, n).astype(np.int16),
           'timestamp': t, 'polarity': g.random(n) < 0.7, 'events_per_element': epe}
    ts = torch.tensor([0., 1.] * B).cuda()
    sidx = torch.arange(B).repeat_interleave(2).cuda()
    dec = dv.decode_events(_cuda(enc), torch.from_numpy(eps).cuda())
    a = dv.compute_event_image(dec, ts, sidx, (H, W))
    c = _cuda(enc)
    b = dv.compute_event_image_compact(c['x'], c['y'], c['timestamp'], c['polarity'], c['events_per_element'], ts, (H, W))
    assert torch.equal(a, b)


def _encoded_host_batch(seed, n_samples, n_events):
    enc, eps = _big(seed, n_samples, n_events)
    n_img = int(eps.astype(np.int64).sum()) + n_samples
    g = np.random.default_rng(seed)
    return {'events': {k: torch.from_numpy(v) for k, v in enc.items()},
            'elements_per_sample': torch.from_numpy(eps),
            'timestamps': torch.from_numpy(g.random(n_img).astype(np.float32)),
            'images': torch.from_numpy(g.integers(0, 256, (n_img, 1, 4, 6)).astype(np.uint8)),
            'augmentation_params': None}


@pytest.mark.parametrize('world', [2, 3])
def test_shard_encoded_batch_equals_shard_of_decoded(world):
    """Host logic of the storage-form feed: sharding the encoded batch and decoding the shard gives
    what sharding the decoded batch gives (training.shard_batch)."""
    from dvs_of_training_framework_b200 import training as tr
    batch = _encoded_host_batch(11, 5, 3000)
    eps = batch['elements_per_sample'].numpy()
    e = {k: v.numpy() for k, v in batch['events'].items()}
    full = eo.decode_events(e['x'], e['y'], e['polarity'], e['timestamp'], e['events_per_element'], eps)
    decoded = {'events': {k: torch.from_numpy(v) for k, v in full.items()},
               'timestamps': batch['timestamps'], 'images': batch['images'], 'size': 5,
               'sample_idx': torch.from_numpy(np.repeat(np.arange(5), eps.astype(np.int64) + 1)),
               'augmentation_params': None}
    total = 0
    for rank in range(world):
        lo, hi = tr.shard_bounds(5, rank, world)
        sh = tr.shard_encoded_batch(batch, rank, world)
        se = {k: v.numpy() for k, v in sh['events'].items()}
        got = eo.decode_events(se['x'], se['y'], se['polarity'], se['timestamp'], se['events_per_element'],
                               sh['elements_per_sample'].numpy())
        sel = (full['sample_index'] >= lo) & (full['sample_index'] < hi)
        for k in DEC_KEYS:
            want = full[k][sel] - (lo if k == 'sample_index' else 0)
            assert np.array_equal(got[k], want), (rank, k)
        img = (decoded['sample_idx'] >= lo) & (decoded['sample_idx'] < hi)
        assert torch.equal(sh['images'], batch['images'][img]) and torch.equal(sh['timestamps'], batch['timestamps'][img])
        total += got['x'].size
    assert total == full['x'].size


@pytest.mark.gpu
def test_prefetcher_accepts_encoded_batches():
    """DevicePrefetcher on storage-form host batches yields the decoded training form on device."""
    from dvs_of_training_framework_b200 import training as tr
    batch = _encoded_host_batch(12, 4, 5000)
    eps = batch['elements_per_sample'].numpy()
    e = {k: v.numpy() for k, v in batch['events'].items()}
    full = eo.decode_events(e['x'], e['y'], e['polarity'], e['timestamp'], e['events_per_element'], eps)
    got = list(tr.DevicePrefetcher([batch, batch], 'cuda', True))
    assert len(got) == 2
    for b in got:
        assert b['size'] == 4 and b['global_size'] == 4 and b['global_events'] == full['x'].size
        assert b['images'].dtype == torch.float32 and b['images'].is_cuda
        assert torch.equal(b['images'].cpu(), batch['images'].float())
        assert b['sample_idx'].cpu().tolist() == np.repeat(np.arange(4), eps.astype(np.int64) + 1).tolist()
        for k in DEC_KEYS:
            assert np.array_equal(b['events'][k].cpu().numpy(), full[k]), k
    sh = list(tr.DevicePrefetcher([batch], 'cuda', True, rank=1, world=2))[0]
    sel = full['sample_index'] >= 2
    assert sh['size'] == 2 and np.array_equal(sh['events']['x'].cpu().numpy(), full['x'][sel])
    assert np.array_equal(sh['events']['sample_index'].cpu().numpy(), full['sample_index'][sel] - 2)


def _quant_case():
    inp = {k.split('/')[-1]: torch.from_numpy(G[k]) for k in G.files if k.startswith('quant/in/')}
    enc = {k.split('/')[-1]: G[k] for k in G.files if k.startswith('quant/enc/')}
    dec = {k.split('/')[-1]: G[k] for k in G.files if k.startswith('quant/dec/')}
    return inp, enc, dec


def test_quantized_batch_codec_matches_reference_on_cpu():
    """utils/dataset.py:429-502 on host tensors (pure data plumbing: runs without the library)."""
    from dvs_of_training_framework_b200 import dataset as ds
    inp, enc, dec = _quant_case()
    got = ds.encode_quantized_batch(dict(inp, augmentation_params={}, size=3))
    for k, v in enc.items():
        assert got[k].dtype == torch.from_numpy(v).dtype and np.array_equal(got[k].numpy(), v), k
    back = ds.decode_quantized_batch(got)
    for k in ('data', 'timestamps', 'sample_idx', 'images'):
        assert back[k].dtype == torch.from_numpy(dec[k]).dtype and np.array_equal(back[k].numpy(), dec[k]), k
    assert back['size'] == int(dec['size'])


@pytest.mark.gpu
def test_quantized_batch_codec_from_device_and_quantize_batch():
    """encode_quantized_batch on a device batch (scripts/quantize_preprocessed.py:77-93) lands on the
    CPU with the reference's values; quantize_batch gives the same event image from the training form
    and from the storage form."""
    import dvs_of_training_framework_b200 as dv
    inp, enc, dec = _quant_case()
    got = dv.encode_quantized_batch({**{k: v.cuda() for k, v in inp.items()}, 'augmentation_params': {}, 'size': 3})
    for k, v in enc.items():
        assert not got[k].is_cuda and np.array_equal(got[k].numpy(), v), k
    g = np.random.default_rng(8)
    B, H, W, n = 3, 40, 56, 2000
    ev = {'x': torch.from_numpy(g.integers(0, W, B * n)), 'y': torch.from_numpy(g.integers(0, H, B * n)),
          'timestamp': torch.from_numpy(np.concatenate([np.sort(g.random(n)) for _ in range(B)]).astype(np.float32)),
          'polarity': torch.from_numpy(g.integers(0, 2, B * n) * 2 - 1),
          'element_index': torch.zeros(B * n, dtype=torch.int64),
          'sample_index': torch.arange(B).repeat_interleave(n)}
    batch = {'events': {k: v.cuda() for k, v in ev.items()}, 'timestamps': torch.tensor([0., 1.] * B).cuda(),
             'sample_idx': torch.arange(B).repeat_interleave(2).cuda(),
             'images': torch.zeros(2 * B, 1, H, W).cuda(), 'augmentation_params': {}, 'size': B}
    q = dv.quantize_batch(batch)
    want = eo.compute_event_image({k: v.numpy() for k, v in ev.items()}, batch['timestamps'].cpu().numpy(),
                                  batch['sample_idx'].cpu().numpy(), dv.extend_size((H, W)))
    assert 'events' not in q and np.array_equal(q['data'].cpu().numpy(), want)
    stored = dv.encode_batch(batch['events'], batch['timestamps'], batch['sample_idx'], batch['images'], {}, B)
    q2 = dv.quantize_batch(stored)
    assert torch.equal(q2['data'], q['data']) and q2['size'] == B
    host = dv.encode_quantized_batch(q2)
    assert host['data'].shape == (B * 4,) + tuple(dv.extend_size((H, W))) and host['data'].is_pinned()


# ---- batch assembly: collate_wrapper (utils/dataset.py:961-1020), host code, runs without a GPU ----
def _collate_samples(g, name):
    n = int(g[f'{name}.n'])
    out = []
    for i in range(n):
        ev = {k[len(f'{name}.in{i}.ev_'):]: g[k] for k in g.files if k.startswith(f'{name}.in{i}.ev_')}
        aug = (i, g[f'{name}.in{i}.ts'].size - 1, 1, np.array([0, 0, 260, 346]), 0, np.array([False]))
        out.append((ev, g[f'{name}.in{i}.ts'], g[f'{name}.in{i}.images'].astype(np.float32), aug))
    return out


@pytest.mark.parametrize('name', ['pair', 'mixed', 'with_empty'])
def test_collate_wrapper_matches_reference(golden, name):
    """The batch the reference's own collate_wrapper builds from the same per-sample tuples
    (tests/golden/make_golden_collate.py; 'pair' is the batch of tests/dataset/test_dataset.py:260-297)."""
    from dvs_of_training_framework_b200.dataset import collate_wrapper
    g = np.load(golden / 'collate.npz')
    batch = collate_wrapper(_collate_samples(g, name))
    assert batch['size'] == int(g[f'{name}.n'])
    assert set(batch['events']) == {'x', 'y', 'timestamp', 'polarity', 'element_index', 'sample_index'}
    for k, v in batch['events'].items():
        want = torch.tensor(g[f'{name}.ev_{k}'])
        assert v.dtype == want.dtype and torch.equal(v, want), k
    for k in ('timestamps', 'sample_idx'):
        want = torch.tensor(g[f'{name}.{k}'])
        assert batch[k].dtype == want.dtype and torch.equal(batch[k], want), k
    assert batch['images'].dtype == torch.float32
    assert torch.equal(batch['images'], torch.tensor(g[f'{name}.images_u8']).float())
    for k, v in batch['augmentation_params'].items():
        want = torch.tensor(g[f'{name}.aug_{k}'])
        assert v.dtype == want.dtype and torch.equal(v, want), k


@pytest.mark.parametrize('name', ['pair', 'mixed'])
def test_collate_wrapper_encoded_equals_reference_encode_batch(golden, name):
    """encoded=True == the reference's encode_batch of the reference's collated batch."""
    from dvs_of_training_framework_b200.dataset import collate_wrapper
    g = np.load(golden / 'collate.npz')
    enc = collate_wrapper(_collate_samples(g, name), encoded=True)
    for k, v in enc['events'].items():
        want = torch.tensor(g[f'{name}.enc_ev_{k}'])
        assert v.dtype == want.dtype and torch.equal(v, want), k
    for k in ('elements_per_sample', 'timestamps'):
        want = torch.tensor(g[f'{name}.enc_{k}'])
        assert enc[k].dtype == want.dtype and torch.equal(enc[k], want), k
    assert enc['images'].dtype == torch.uint8 and torch.equal(enc['images'], torch.tensor(g[f'{name}.images_u8']))


@pytest.mark.gpu
@pytest.mark.parametrize('name', ['pair', 'mixed'])
def test_encoded_collate_through_device_prefetcher_equals_reference_batch(golden, name):
    """Device-side batch assembly: the storage-form batch of
    collate_wrapper(encoded=True) uploaded and widened on the device by training.DevicePrefetcher is
    the batch the reference's collate_wrapper builds (int64 events with sample_index, float32 images)."""
    from dvs_of_training_framework_b200 import training as tr
    from dvs_of_training_framework_b200.dataset import collate_wrapper
    g = np.load(golden / 'collate.npz')
    enc = collate_wrapper(_collate_samples(g, name), encoded=True)
    assert enc['events']['x'].is_pinned()
    (dev_batch,) = list(tr.DevicePrefetcher([enc], 'cuda', True))
    for k in ('x', 'y', 'timestamp', 'polarity', 'element_index', 'sample_index'):
        want = torch.tensor(g[f'{name}.ev_{k}'])
        got = dev_batch['events'][k]
        assert got.is_cuda and got.dtype == want.dtype and torch.equal(got.cpu(), want), k
    assert torch.equal(dev_batch['timestamps'].cpu(), torch.tensor(g[f'{name}.timestamps']))
    assert torch.equal(dev_batch['sample_idx'].cpu(), torch.tensor(g[f'{name}.sample_idx']))
    assert dev_batch['images'].dtype == torch.float32
    assert torch.equal(dev_batch['images'].cpu(), torch.tensor(g[f'{name}.images_u8']).float())
    assert dev_batch['global_size'] == int(g[f'{name}.n'])
