"""Golden vectors for the batch codec, produced by the REFERENCE's own utils/dataset.py
(encode_batch :240-305, decode_batch :335-373) in the build container.

    python tests/golden/make_golden_codec.py        -> tests/golden/codec.npz (committed)

utils/dataset.py is imported unmodified; `h5py` (absent here, only used by the HDF5 readers) and
the sibling modules the codec does not touch are empty stand-ins.  Cases: the fixture of
tests/dataset/test_encoding.py:21-80 and seeded random batches with empty elements, empty
trailing elements (dropped by encode_batch, :298) and one-element samples."""
import importlib.util
import sys
import types
from pathlib import Path

import numpy as np
import torch

REF = Path('/root/reference')
OUT = Path(__file__).parent


def load_dataset_module():
    if 'h5py' not in sys.modules:                     # only named in annotations of the HDF5 readers
        h5 = types.ModuleType('h5py')
        h5.File = object
        sys.modules['h5py'] = h5
    pkg = types.ModuleType('utils')
    pkg.__path__ = [str(REF / 'utils')]
    sys.modules['utils'] = pkg
    # siblings imported at module level: utils/common.py is the real one (cumsum_with_prefix is used
    # by decode_batch); the others are not touched by the codec functions
    spec = importlib.util.spec_from_file_location('utils.common', REF / 'utils' / 'common.py')
    common = importlib.util.module_from_spec(spec)
    common.__package__ = 'utils'
    sys.modules['utils.common'] = common
    spec.loader.exec_module(common)
    stubs = {'EV_FlowNet': [], 'EV_FlowNet.net': ['compute_event_image'],
             'utils.data': ['EventCrop', 'ImageRandomCrop', 'RandomRotation', 'ImageCentralCrop'],
             'utils.file_iterators': ['create_file_iterator'], 'tqdm': [], 'yaml': []}
    for name, syms in stubs.items():
        if name in sys.modules and name in ('tqdm', 'yaml'):
            continue
        try:
            if name in ('tqdm', 'yaml'):
                __import__(name)
                continue
        except ImportError:
            pass
        m = types.ModuleType(name)
        for sym in syms:
            setattr(m, sym, object)
        sys.modules[name] = m
    spec = importlib.util.spec_from_file_location('utils.dataset', REF / 'utils' / 'dataset.py')
    mod = importlib.util.module_from_spec(spec)
    mod.__package__ = 'utils'
    sys.modules['utils.dataset'] = mod
    spec.loader.exec_module(mod)
    return mod


def random_decoded(seed, n_samples, max_elements, max_events, drop_tail):
    g = np.random.default_rng(seed)
    eps = g.integers(1, max_elements + 1, n_samples)
    if drop_tail:
        eps[-1] = max(eps[-1], 2)
    x, y, t, p, e, s = [], [], [], [], [], []
    for i, ne in enumerate(eps):
        for j in range(ne):
            k = int(g.integers(0, max_events + 1)) if g.random() > 0.25 else 0
            if i == n_samples - 1 and j == 0:
                k = max(k, 1)       # a last sample without events makes the reference's decode_batch raise (:360)
            if drop_tail and i == n_samples - 1 and j == ne - 1:
                k = 0                                           # trailing empty element: dropped by :298
            x += g.integers(0, 346, k).tolist(); y += g.integers(0, 260, k).tolist()
            t += np.sort(g.random(k)).astype(np.float32).tolist()
            p += (g.integers(0, 2, k) * 2 - 1).tolist()
            e += [j] * k; s += [i] * k
    if not x:                                                   # the reference indexes the last event
        x, y, t, p, e, s = [1], [1], [0.5], [1], [0], [0]
    timestamps = np.concatenate([np.linspace(0, 1, ne + 1, dtype=np.float32) for ne in eps])
    sample_idx = np.concatenate([np.full(ne + 1, i) for i, ne in enumerate(eps)])
    L = lambda a: torch.tensor(a, dtype=torch.long)
    return {'events': {'x': L(x), 'y': L(y), 'timestamp': torch.tensor(t, dtype=torch.float32), 'polarity': L(p),
                       'element_index': L(e), 'sample_index': L(s)},
            'timestamps': torch.tensor(timestamps), 'sample_idx': L(sample_idx),
            'images': torch.zeros(sample_idx.size, 1, 2, 2), 'augmentation_params': {}, 'size': n_samples}


def main():
    ds = load_dataset_module()
    L = lambda a: torch.tensor(a, dtype=torch.long)
    fixture = {'events': {'x': L([1, 2, 2, 1, 4, 6, 7]), 'y': L([2, 1, 3, 4, 1, 6, 8]),
                          'timestamp': torch.tensor([0.02, 0.06, 0.07, 0.015, 0.01, 0.05, 0.07]),
                          'polarity': L([-1, 1, -1, 1, 1, 1, -1]), 'element_index': L([0, 1, 1, 0, 0, 2, 3]),
                          'sample_index': L([0, 0, 0, 1, 2, 2, 2])},
               'timestamps': torch.tensor([0, 0.04, 0.08, 0, 0.03, 0, 0.02, 0.04, 0.06, 0.08]),
               'sample_idx': L([0, 0, 0, 1, 1, 2, 2, 2, 2, 2]), 'images': torch.zeros(10, 1, 2, 2),
               'augmentation_params': {}, 'size': 3}            # tests/dataset/test_encoding.py:21-50
    cases = {'fixture': fixture,
             'ragged': random_decoded(1, 5, 4, 40, False),
             'droptail': random_decoded(2, 3, 5, 25, True),
             'single': random_decoded(3, 1, 1, 9, False),
             'bulk': random_decoded(4, 8, 3, 700, False)}
    out = {}
    for name, dec in cases.items():
        enc = ds.encode_batch(dec['events'], dec['timestamps'], dec['sample_idx'], dec['images'],
                              dec['augmentation_params'], dec['size'])
        back = ds.decode_batch(enc)
        for k, v in dec['events'].items():
            out[f'{name}/dec/{k}'] = v.numpy()
        out[f'{name}/sample_idx'] = dec['sample_idx'].numpy()
        out[f'{name}/size'] = np.int64(dec['size'])
        for k, v in enc['events'].items():
            out[f'{name}/enc/{k}'] = v.numpy()
        out[f'{name}/enc/elements_per_sample'] = enc['elements_per_sample'].numpy()
        for k, v in back['events'].items():
            out[f'{name}/back/{k}'] = v.numpy()
        out[f'{name}/back/sample_idx'] = back['sample_idx'].numpy()
    # utils/dataset.py:429-502 encode_quantized_batch / decode_quantized_batch on a small quantized batch
    g = torch.Generator().manual_seed(5)
    qb = {'data': torch.rand(3, 4, 6, 8, generator=g), 'timestamps': cases['fixture']['timestamps'],
          'sample_idx': cases['fixture']['sample_idx'], 'images': torch.randint(0, 256, (10, 1, 6, 8), generator=g).float(),
          'augmentation_params': {}, 'size': 3}
    qe = ds.encode_quantized_batch({k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in qb.items()})
    qd = ds.decode_quantized_batch(qe)
    for k in ('data', 'timestamps', 'sample_idx', 'images'):
        out[f'quant/in/{k}'] = qb[k].numpy()
    for k in ('data', 'channels_per_sample', 'elements_per_sample', 'timestamps', 'images'):
        out[f'quant/enc/{k}'] = qe[k].numpy()
    for k in ('data', 'timestamps', 'sample_idx', 'images'):
        out[f'quant/dec/{k}'] = qd[k].numpy()
    out['quant/dec/size'] = np.int64(qd['size'])
    np.savez_compressed(OUT / 'codec.npz', **out)
    print('wrote', OUT / 'codec.npz', {k: out[f'{k}/dec/x'].size for k in cases})


if __name__ == '__main__':
    main()
