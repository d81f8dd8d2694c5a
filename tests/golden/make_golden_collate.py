"""Golden vectors for batch assembly, produced by the REFERENCE's own utils/dataset.py:
`collate_wrapper` (:961-1020) on per-sample tuples of the shape DatasetImpl.__getitem__ returns
(:776-797), and `encode_batch` (:240-305) of that batch.  The samples are fixtures 0..2 exactly as
tests/dataset/test_dataset.py:260-297 builds its expected batch (timestamps aligned to the start
frame), plus a two-element sample (fixtures 3+4 collapsed, element_index 0/1) and an empty one.

    python tests/golden/make_golden_collate.py        -> tests/golden/collate.npz (committed)"""
import sys
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).parent
sys.path.insert(0, str(OUT))
from make_golden_codec import REF, load_dataset_module  # noqa: E402
from _h5min import read_all  # noqa: E402


def sample_from_fixtures(ids, idx):
    """(events dict, image_ts, images, augmentation tuple) like DatasetImpl.__getitem__ (is_raw, no
    augmentation): elements `ids` concatenated, timestamps aligned to the first start."""
    ev, ts, ims = [], [], []
    t0 = None
    for e, i in enumerate(ids):
        d = read_all(REF / f'tests/data/seq/{i:06d}.hdf5')
        if t0 is None:
            t0 = float(d['start'])
            ts.append(0.0)
            ims.append(d['image1'][None])
        rows = d['events'].copy()
        rows[:, 2] -= t0
        ev.append(np.hstack([rows, np.full((rows.shape[0], 1), e)]))
        ts.append(float(d['stop']) - t0)
        ims.append(d['image2'][None])
    ev = np.vstack(ev).astype(np.float32)
    events = {'x': ev[:, 0].astype(np.int_), 'y': ev[:, 1].astype(np.int_), 'timestamp': ev[:, 2],
              'polarity': ev[:, 3].astype(np.int_), 'element_index': ev[:, 4].astype(np.int_)}
    images = np.concatenate(ims, 0).astype(np.float32)
    aug = (idx, len(ids), 1, np.array([0, 0, 260, 346]), 0, np.array([False]))
    return events, np.array(ts), images, aug


def empty_sample(idx):
    events = {'x': np.zeros(0, np.int_), 'y': np.zeros(0, np.int_), 'timestamp': np.zeros(0, np.float32),
              'polarity': np.zeros(0, np.int_), 'element_index': np.zeros(0, np.int_)}
    return events, np.array([0.0, 0.03]), np.zeros((2, 260, 346), np.float32), \
        (idx, 1, 1, np.array([0, 0, 260, 346]), 0, np.array([False]))


def main():
    ds = load_dataset_module()
    out = {}
    cases = {'pair': [sample_from_fixtures([0], 0), sample_from_fixtures([1], 1)],
             'mixed': [sample_from_fixtures([2], 0), sample_from_fixtures([3, 4], 1), sample_from_fixtures([5], 2)],
             'with_empty': [sample_from_fixtures([6], 0), empty_sample(1), sample_from_fixtures([7], 2)]}
    for name, samples in cases.items():
        for i, (ev, ts, im, aug) in enumerate(samples):
            for k, v in ev.items():
                out[f'{name}.in{i}.ev_{k}'] = v
            out[f'{name}.in{i}.ts'] = ts
            out[f'{name}.in{i}.images'] = im.astype(np.uint8)
        out[f'{name}.n'] = np.int64(len(samples))
        batch = ds.collate_wrapper([(dict(ev), ts.copy(), im.copy(), aug) for ev, ts, im, aug in samples])
        for k, v in batch['events'].items():
            out[f'{name}.ev_{k}'] = v.numpy()
        for k in ('timestamps', 'sample_idx'):
            out[f'{name}.{k}'] = batch[k].numpy()
        out[f'{name}.images_u8'] = batch['images'].numpy().astype(np.uint8)
        assert torch.equal(batch['images'], torch.tensor(out[f'{name}.images_u8']).float())
        for k, v in batch['augmentation_params'].items():
            out[f'{name}.aug_{k}'] = v.numpy()
        if name != 'with_empty':           # encode_batch indexes new_e[-1]: needs a last event
            enc = ds.encode_batch(**{k: batch[k] for k in ('events', 'timestamps', 'sample_idx', 'images',
                                                          'augmentation_params', 'size')})
            for k, v in enc['events'].items():
                out[f'{name}.enc_ev_{k}'] = v.numpy()
            out[f'{name}.enc_elements_per_sample'] = enc['elements_per_sample'].numpy()
            out[f'{name}.enc_timestamps'] = enc['timestamps'].numpy()
        print(name, {k: tuple(v.shape) for k, v in batch['events'].items()}, batch['size'])
    np.savez_compressed(OUT / 'collate.npz', **out)


if __name__ == '__main__':
    main()
