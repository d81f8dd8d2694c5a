"""Golden vector for BASELINE.json configs[0]: the reference's DummyNet (DummyNet/net.py:43-80, a
zero flow at 4 scales + timestamp bookkeeping) through the reference's own `init_losses`
(utils/loss.py:217-240) and `process_minibatch` (utils/training.py:37-86) on a 260x346 DAVIS
batch of 1 -- fixture 1's two frames, its 190-odd events -- on the CPU, unmodified.

Run where /root/reference exists:   python tests/golden/make_golden_dummynet.py
Output (committed): tests/golden/dummynet_cfg1.npz
"""
import sys
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).parent
sys.path.insert(0, str(OUT))
from make_golden import REF, _load, load_reference  # noqa: E402
from _h5min import read_all  # noqa: E402


def main():
    ref_loss, _, _ = load_reference()
    _load('utils.common', REF / 'utils/common.py', 'utils')
    training = _load('utils.training', REF / 'utils/training.py', 'utils')
    dummy = _load('ref_dummynet_net', REF / 'DummyNet/net.py')
    timer = sys.modules['utils.timer']

    d = read_all(REF / 'tests/data/seq/000001.hdf5')
    H, W = d['image1'].shape                                   # 260 x 346
    ev = d['events']                                           # (n, 4) float64: x, y, t, p
    t0 = float(d['start'])
    events = {'x': torch.tensor(ev[:, 0]).long(), 'y': torch.tensor(ev[:, 1]).long(),
              'timestamp': torch.tensor(ev[:, 2] - t0).float(), 'polarity': torch.tensor(ev[:, 3]).long(),
              'element_index': torch.zeros(len(ev), dtype=torch.long),
              'sample_index': torch.zeros(len(ev), dtype=torch.long)}
    images = torch.tensor(np.stack([d['image1'], d['image2']])[:, None]).float()
    timestamps = torch.tensor([0.0, float(d['stop']) - t0], dtype=torch.float32)
    sample_idx = torch.zeros(2, dtype=torch.long)
    batch = {'events': events, 'timestamps': timestamps, 'sample_idx': sample_idx, 'images': images,
             'augmentation_params': None, 'size': 1}

    model = dummy.Model('cpu')
    evaluator = ref_loss.init_losses((H, W), 1, model, 'cpu', 1)
    shapes = [(l.H, l.W) for l in evaluator.losses]
    loss, terms, tags, info = training.process_minibatch(model, batch, timer.FakeTimer(), 'cpu', True,
                                                         evaluator, [0.5, 1, 1], return_prediction=True)
    terms = np.array([[v for v in group] for group in terms], dtype=np.float64)
    np.savez_compressed(OUT / 'dummynet_cfg1.npz', images=images.numpy().astype(np.uint8),
                        timestamps=timestamps.numpy(), sample_idx=sample_idx.numpy(),
                        **{f'ev_{k}': v.numpy() for k, v in events.items()},
                        shapes=np.array(shapes), loss=np.float64(loss.item()), terms=terms,
                        tags=np.array(list(tags)), flow_ts=info['flow_ts'].numpy(),
                        flow_sample_idx=info['flow_sample_idx'].numpy())
    print('shapes', shapes, 'loss', float(loss), 'terms', terms)


if __name__ == '__main__':
    main()
