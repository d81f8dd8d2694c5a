"""The preprocessed-dataset loader (utils/dataset.py:19-198, 400-425, 505-549, 799-954): index
arithmetic against goldens produced by the reference's own functions
(tests/golden/make_golden_ranges.py), the reference's known answers of
tests/dataset/test_encoding.py:142-268, and the loader over file boundaries / wrap-around the way
tests/dataset/test_encoding.py:291-313 exercises it -- with .npz-backed stand-ins for the HDF5 files
(h5py is not installed offline; the reading code only needs `descriptor[key][begin:end]`)."""
import json

import numpy as np
import pytest
import torch

from dvs_of_training_framework_b200 import preprocessed as pp


def _fixture_parts():
    """tests/dataset/test_encoding.py:81-128 `encoded_parts` (two files: samples 0-1 and sample 2)."""
    def imgs(vals):
        return torch.tensor(vals, dtype=torch.uint8).view(-1, 1, 1, 1).tile(1, 1, 10, 10)
    a = {'events': {'x': torch.tensor([1, 2, 2, 1], dtype=torch.short), 'y': torch.tensor([2, 1, 3, 4], dtype=torch.short),
                    'timestamp': torch.tensor([0.02, 0.06, 0.07, 0.015]), 'polarity': torch.tensor([False, True, False, True]),
                    'events_per_element': torch.tensor([1, 2, 1])},
         'timestamps': torch.tensor([0, 0.04, 0.08, 0, 0.03]), 'elements_per_sample': torch.tensor([2, 1], dtype=torch.uint8),
         'images': imgs([0, 1, 2, 3, 4]),
         'augmentation_params': {'idx': torch.tensor([0, 1]), 'sequence_length': torch.tensor([2, 1], dtype=torch.short),
                                 'collapse_length': torch.tensor([1, 2], dtype=torch.short),
                                 'box': torch.tensor([[0, 0, 10, 10], [0, 1, 10, 10]]),
                                 'angle': torch.tensor([0.1, 0.2]), 'is_flip': torch.tensor([True, False])}}
    b = {'events': {'x': torch.tensor([4, 6, 7], dtype=torch.short), 'y': torch.tensor([1, 6, 8], dtype=torch.short),
                    'timestamp': torch.tensor([0.01, 0.05, 0.07]), 'polarity': torch.tensor([True, True, False]),
                    'events_per_element': torch.tensor([1, 0, 1, 1])},
         'timestamps': torch.tensor([0, 0.02, 0.04, 0.06, 0.08]), 'elements_per_sample': torch.tensor([4], dtype=torch.uint8),
         'images': imgs([5, 6, 7, 8, 9]),
         'augmentation_params': {'idx': torch.tensor([2]), 'sequence_length': torch.tensor([4], dtype=torch.short),
                                 'collapse_length': torch.tensor([3], dtype=torch.short),
                                 'box': torch.tensor([[1, 0, 10, 10]]), 'angle': torch.tensor([0.3]),
                                 'is_flip': torch.tensor([True])}}
    return [a, b]


def _same(a, b, path=''):
    assert type(a) is type(b) or (isinstance(a, dict) and isinstance(b, dict)), path
    if isinstance(a, dict):
        assert set(a) == set(b), (path, set(a), set(b))
        for k in a:
            _same(a[k], b[k], f'{path}/{k}')
    elif isinstance(a, torch.Tensor):
        assert a.dtype == b.dtype and torch.equal(a.cpu(), b.cpu()), path
    else:
        assert a == b, path


def test_select_ranges_match_reference(golden):
    cases = json.loads((golden / 'ranges.json').read_text())
    n = 0
    for c in cases:
        epe = torch.tensor(c['events_per_element'], dtype=torch.long)
        eps = torch.tensor(c['elements_per_sample'], dtype=torch.uint8)
        cps = torch.tensor(c['channels_per_sample'], dtype=torch.long)
        for s in c['selections']:
            assert pp.select_encoded_ranges(epe, eps, s['begin'], s['end']) == s['encoded']
            assert pp.select_quantized_ranges(cps, eps, s['begin'], s['end']) == s['quantized']
            n += 1
    assert n == 100


def test_select_encoded_ranges_reference_known_answers():
    # tests/dataset/test_encoding.py:142-268 on events_per_element [1,2,1,1,0,1,1], elements_per_sample [2,1,4]
    epe, eps = torch.tensor([1, 2, 1, 1, 0, 1, 1]), torch.tensor([2, 1, 4], dtype=torch.uint8)
    for (b, e), (ev, el, ts) in {(0, 1): ((0, 3), (0, 2), (0, 3)), (1, 2): ((3, 4), (2, 3), (3, 5)),
                                 (2, 3): ((4, 7), (3, 7), (5, 10)), (0, 2): ((0, 4), (0, 3), (0, 5)),
                                 (1, 3): ((3, 7), (2, 7), (3, 10)), (0, 3): ((0, 7), (0, 7), (0, 10))}.items():
        r = pp.select_encoded_ranges(epe, eps, b, e)
        assert r['events']['x'] == {'begin': ev[0], 'end': ev[1]} == r['events']['polarity']
        assert r['events']['events_per_element'] == {'begin': el[0], 'end': el[1]}
        assert r['timestamps'] == {'begin': ts[0], 'end': ts[1]} == r['images']
        assert r['elements_per_sample'] == {'begin': b, 'end': e} == r['augmentation_params']['box']
    with pytest.raises(AssertionError):
        pp.select_encoded_ranges(epe, eps, 1, 1)


def test_join_batches_reference_vector():
    # tests/dataset/test_encoding.py:138-140
    parts = _fixture_parts()
    joined = pp.join_batches(parts)
    assert joined['events']['events_per_element'].tolist() == [1, 2, 1, 1, 0, 1, 1]
    assert joined['elements_per_sample'].tolist() == [2, 1, 4] and joined['elements_per_sample'].dtype == torch.uint8
    assert joined['events']['x'].tolist() == [1, 2, 2, 1, 4, 6, 7] and joined['events']['x'].dtype == torch.short
    assert joined['timestamps'].numel() == 10 and joined['images'].shape == (10, 1, 10, 10)
    assert joined['augmentation_params']['idx'].tolist() == [0, 1, 2]
    assert pp.join_batches(parts[:1]) is parts[0]
    assert pp.join_batches([])['events']['polarity'].dtype == torch.bool


def _write(dirname, parts):
    """One flat .npz per part: nested keys joined with '/', like the groups of the HDF5 files."""
    def flat(d, prefix=''):
        out = {}
        for k, v in d.items():
            out.update(flat(v, f'{prefix}{k}/') if isinstance(v, dict) else {f'{prefix}{k}': v.numpy()})
        return out
    for i, part in enumerate(parts):
        np.savez(dirname / f'{i}.npz', **flat(part))


def _opener(path):
    z = np.load(path)
    tree = {}
    for k in z.files:
        node = tree
        *groups, leaf = k.split('/')
        for g in groups:
            node = node.setdefault(g, {})
        node[leaf] = z[k]
    return tree


def test_read_encoded_batch_and_loader_over_files(tmp_path):
    """tests/dataset/test_encoding.py:270-313: sub-ranges of one file, a batch per file, a batch over
    both files, a batch that wraps around to the first file, set_index."""
    parts = _fixture_parts()
    _write(tmp_path, parts)
    whole = pp.join_batches(parts)
    d = tmp_path / 'whole'
    d.mkdir()
    _write(d, [whole])
    f = _opener(d / '0.npz')
    epe, eps = torch.as_tensor(f['events']['events_per_element']), torch.as_tensor(f['elements_per_sample'])
    _same(pp.read_encoded_batch(f, epe, eps, 0, 2), parts[0])
    _same(pp.read_encoded_batch(f, epe, eps, 2, 3), parts[1])

    def loader(bs):
        return pp.PreprocessedDataloader(tmp_path, bs, is_raw=True, opener=_opener, suffix='.npz')
    ld = loader(2)
    assert len(ld) == 3 and (tmp_path / '0.info').is_file()
    _same(next(ld), parts[0])
    ld = loader(1)
    ld.set_index(2)
    _same(next(ld), parts[1])
    _same(next(loader(3)), whole)
    _same(next(loader(5)), pp.join_batches(parts + [parts[0]]))
    ld = loader(2)                                   # consecutive batches keep their place across files
    next(ld)
    _same(next(ld), pp.join_batches([parts[1], pp.read_encoded_batch(_opener(tmp_path / '0.npz'),
                                                                   parts[0]['events']['events_per_element'],
                                                                   parts[0]['elements_per_sample'], 0, 1)]))


def test_missing_h5py_is_loud(tmp_path):
    (tmp_path / '0.hdf5').write_bytes(b'')
    try:
        import h5py  # noqa: F401
        pytest.skip('h5py is installed here')
    except ImportError:
        pass
    with pytest.raises(RuntimeError, match='h5py'):
        pp.PreprocessedDataloader(tmp_path, 1, is_raw=True)


@pytest.mark.gpu
def test_loader_on_device_returns_the_decoded_training_batch(tmp_path):
    """device='cuda': what the reference's loader returns (decode_batch of the joined batch,
    tests/dataset/test_encoding.py:291-313), decoded on the device; and the storage-form batch of the
    default mode through training.DevicePrefetcher gives the same."""
    from dvs_of_training_framework_b200 import dataset, training as tr
    parts = _fixture_parts()
    _write(tmp_path, parts)
    ld = pp.PreprocessedDataloader(tmp_path, 3, is_raw=True, opener=_opener, suffix='.npz', device='cuda')
    got = next(ld)
    ev = got['events']
    assert ev['x'].is_cuda and ev['x'].dtype == torch.int64
    assert ev['x'].tolist() == [1, 2, 2, 1, 4, 6, 7] and ev['polarity'].tolist() == [-1, 1, -1, 1, 1, 1, -1]
    assert ev['element_index'].tolist() == [0, 1, 1, 0, 0, 2, 3] and ev['sample_index'].tolist() == [0, 0, 0, 1, 2, 2, 2]
    assert got['sample_idx'].tolist() == [0, 0, 0, 1, 1, 2, 2, 2, 2, 2] and got['images'].dtype == torch.float32
    want = dataset.decode_batch({k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v.cuda())
                                 for k, v in pp.join_batches(parts).items()})
    for k in ev:
        assert torch.equal(ev[k], want['events'][k]), k
    host = pp.PreprocessedDataloader(tmp_path, 3, is_raw=True, opener=_opener, suffix='.npz')
    (staged,) = list(tr.DevicePrefetcher([next(host)], 'cuda', True))
    for k in ev:
        assert torch.equal(staged['events'][k], ev[k]), k
    assert torch.equal(staged['images'], got['images'])


def test_dataloader_factories_keep_the_reference_wiring(tmp_path, golden):
    """utils/dataloader.py:31-108: parameter objects, the preprocessed branch (positioned with
    set_index) and the DataLoader branch with this package's collate_wrapper as collate_fn."""
    from pathlib import Path
    from types import SimpleNamespace
    from dvs_of_training_framework_b200 import dataloader as dl
    from dvs_of_training_framework_b200.dataset import collate_wrapper
    args = SimpleNamespace(shape=[256, 256], mbs=2, num_workers=0, min_sequence_length=1, max_sequence_length=1,
                           dynamic_sample_length=False, data_path=Path('/data'), cl=6, is_raw=True, cache_size=0)
    tp, vp = dl.get_trainset_params(args), dl.get_valset_params(args)
    assert tp.path == Path('/data/outdoor_day2') and tp.augmentation and tp.infinite and tp.shuffle
    assert tp.collate_fn is collate_wrapper and tp.preprocessed_dataset_path is None and tp.collapse_length == 6
    assert vp.path == Path('/data/outdoor_day1') and not vp.augmentation and not vp.infinite and vp.collapse_length == 1
    assert dl.choose_collate_function(False) is None
    assert dl.get_trainset_params(args, encoded=True).collate_fn.keywords == {'encoded': True}
    # preprocessed branch
    _write(tmp_path, _fixture_parts())
    tp.preprocessed_dataset_path, tp.cache_dir = tmp_path, None
    ld = dl.get_dataloader(tp, sample_idx=2, opener=_opener, suffix='.npz')
    assert isinstance(ld, pp.PreprocessedDataloader) and ld.batch_size == 2
    assert next(ld)['events']['x'].tolist()[:3] == [4, 6, 7]           # starts at sample 2, wraps to sample 0
    # DataLoader branch over a map-style dataset of reference-shaped samples
    g = np.load(golden / 'collate.npz')

    class TwoSamples(torch.utils.data.Dataset):
        def __init__(self, **kw):
            self.kw = kw

        def __len__(self):
            return 2

        def __getitem__(self, i):
            ev = {k[len(f'pair.in{i}.ev_'):]: g[k] for k in g.files if k.startswith(f'pair.in{i}.ev_')}
            return ev, g[f'pair.in{i}.ts'], g[f'pair.in{i}.images'].astype(np.float32), \
                (i, 1, 1, np.array([0, 0, 260, 346]), 0, np.array([False]))
    vp.pin_memory = False
    batch = next(iter(dl.get_dataloader(vp, dataset_classes=(TwoSamples, None))))
    assert batch['size'] == 2 and torch.equal(batch['events']['sample_index'], torch.tensor(g['pair.ev_sample_index']))
    assert torch.equal(batch['events']['x'], torch.tensor(g['pair.ev_x']))
