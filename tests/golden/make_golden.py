"""Generate golden vectors by running the REFERENCE's own code in the build container.

Run once where /root/reference exists:   python tests/golden/make_golden.py
Outputs (committed): tests/golden/*.npz.  The GPU box has no /root/reference; tests only
read the .npz files.

What is run unmodified from /root/reference:
  utils/loss.py           (Losses, Loss, charbonier_loss)  -- imported without the
                          `utils/__init__` chain (needs only utils/timer.py, utils/logging.py)
  EV_FlowNet/net.py       compute_event_image -- with a `torch_scatter` stand-in providing
                          scatter_sum / scatter_max through Tensor.scatter_add_ /
                          Tensor.scatter_reduce_('amax', include_self=True)
  utils/data.py           EventCrop, get_count_image, RandomRotation -- the native
                          `transformation.map` it imports is the reference's own
                          src/transformation.cpp compiled by oracle/Makefile (Eigen container
                          stand-in: oracle/ref_shim/transformation.h)
"""
import importlib.util
import sys
import types
from pathlib import Path

import numpy as np
import torch

REF = Path('/root/reference')
OUT = Path(__file__).parent
sys.path.insert(0, str(OUT))
from _h5min import read_all  # noqa: E402


def _load(name, path, package=None):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    if package:
        mod.__package__ = package
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference():
    # --- torch_scatter stand-in (semantics: oracle/encoder_oracle.py header) ---
    ts = types.ModuleType('torch_scatter')

    def scatter_sum(src, index, out):
        return out.scatter_add_(0, index, src)

    def scatter_max(src, index, out):
        out.scatter_reduce_(0, index, src, 'amax', include_self=True)
        return out, None
    ts.scatter_sum, ts.scatter_max = scatter_sum, scatter_max
    sys.modules['torch_scatter'] = ts

    # --- `utils` package shell so relative imports resolve without utils/__init__ ---
    pkg = types.ModuleType('utils')
    pkg.__path__ = [str(REF / 'utils')]
    sys.modules['utils'] = pkg

    # the native `transformation.map`: the reference's own src/transformation.cpp (:35-85),
    # compiled where it lies by oracle/Makefile against a container stand-in for the absent Eigen
    # headers (oracle/ref_shim/transformation.h) -- not a restatement
    tr = types.ModuleType('utils.transformation')
    sys.path.insert(0, str(OUT.parent.parent))
    from oracle import map_events_native as native
    native.build()
    assert native.available('ref'), 'oracle/_ref/libmap_events_ref.so was not built'

    def map_events(events, shape, src, dst):
        assert events.dtype == np.float32 and events.flags.c_contiguous      # module.cpp: Eigen::Ref, no cast
        assert src.dtype == np.uint64 and dst.dtype == np.uint64
        return native.ref_map(events, shape, src, dst)
    tr.map = map_events
    sys.modules['utils.transformation'] = tr

    _load('utils.logging', REF / 'utils/logging.py', 'utils')
    _load('utils.timer', REF / 'utils/timer.py', 'utils')
    loss = _load('utils.loss', REF / 'utils/loss.py', 'utils')
    data = _load('utils.data', REF / 'utils/data.py', 'utils')
    net = _load('ref_evflownet_net', REF / 'EV_FlowNet/net.py')
    return loss, data, net


def synth_events(g, B, H, W, n_per_sample, T=0.032, ragged=False):
    xs, ys, ts, ps, bs = [], [], [], [], []
    for b in range(B):
        n = n_per_sample if not ragged else int(torch.randint(0, n_per_sample + 1, (1,), generator=g))
        if ragged and b == 1:
            n = 0
        xs.append(torch.randint(0, W, (n,), generator=g))
        ys.append(torch.randint(0, H, (n,), generator=g))
        ts.append(torch.sort(torch.rand(n, generator=g) * T)[0])
        ps.append((torch.rand(n, generator=g) < 0.7).long() * 2 - 1)
        bs.append(torch.full((n,), b, dtype=torch.long))
    ev = {'x': torch.cat(xs), 'y': torch.cat(ys), 'timestamp': torch.cat(ts).float(),
          'polarity': torch.cat(ps), 'sample_index': torch.cat(bs)}
    ev['element_index'] = torch.zeros_like(ev['x'])
    timestamps = torch.tensor([0, T] * B, dtype=torch.float32)
    sample_idx = torch.arange(B).repeat_interleave(2)
    return ev, timestamps, sample_idx


def main():
    ref_loss, ref_data, ref_net = load_reference()
    torch.manual_seed(0)

    # ---------------- fixtures (tests/data/seq, tests/data/pred) ----------------
    for i in (0, 1):
        d = read_all(REF / f'tests/data/seq/{i:06d}.hdf5')
        p = read_all(REF / f'tests/data/pred/{i:06d}.hdf5')
        np.savez_compressed(OUT / f'fixture_{i:06d}.npz', events=d['events'],
                            start=d['start'], stop=d['stop'], image1=d['image1'],
                            image2=d['image2'], flow=p['flow'])

    # ---------------- loss: reference KATs on fixture 1 (tests/loss/test_loss.py) ------------
    d = read_all(REF / 'tests/data/seq/000001.hdf5')
    p = read_all(REF / 'tests/data/pred/000001.hdf5')
    H, W = 246, 340
    images = torch.tensor(np.stack([d['image1'][:H, :W], d['image2'][:H, :W]])[:, None]).float()
    ts = torch.tensor([0, float(d['stop']) - float(d['start'])], dtype=torch.float32)
    sidx = torch.LongTensor([0, 0])
    fsidx = torch.LongTensor([0])
    ev = ref_loss.Losses([(H, W)], 1, 'cpu')
    kat = {}
    for name, flow in (('zero', torch.zeros(1, 2, H, W)),
                       ('pred', torch.tensor(p['flow'][:H, :W]).permute(2, 0, 1)[None].contiguous())):
        out = ev([flow], ts.view(1, 2), fsidx, images, ts, sidx)
        kat[name] = np.array([float(o[0]) for o in out], dtype=np.float64)
        print('KAT', name, kat[name])
    np.savez(OUT / 'loss_kat.npz', **kat)

    # ---------------- loss: seeded random multiscale, values + gradients ----------------
    for tag, (B, H, W, sigma) in {'a': (2, 64, 80, 3.0), 'b': (3, 40, 56, 20.0),
                                  'c': (1, 260, 346, 3.0)}.items():
        g = torch.Generator().manual_seed(4321 + ord(tag))
        images = torch.randint(0, 256, (2 * B, 1, H, W), generator=g).float()
        ts = torch.tensor([0.0, 0.032] * B)
        sidx = torch.arange(B).repeat_interleave(2)
        flow_ts = ts.view(-1, 2).clone()
        fsidx = torch.arange(B)
        shapes = [(H // 2 ** k, W // 2 ** k) for k in (3, 2, 1, 0)]
        flows = [(torch.randn(B, 2, h, w, generator=g) * sigma / 2 ** k).requires_grad_()
                 for k, (h, w) in zip((3, 2, 1, 0), shapes)]
        ev = ref_loss.Losses(shapes, B, 'cpu')
        terms = ev(flows, flow_ts, fsidx, images, ts, sidx)
        total = sum(w * sum(t) / len(t) for w, t in zip((0.5, 1, 1), terms))
        total.backward()
        save = {'images': images.numpy().astype(np.uint8), 'timestamps': ts.numpy(),
                'sample_idx': sidx.numpy(), 'flow_ts': flow_ts.numpy(),
                'flow_sample_idx': fsidx.numpy(),
                'terms': np.array([[float(v) for v in t] for t in terms], dtype=np.float64),
                'total': float(total)}
        for k, f in enumerate(flows):
            save[f'flow{k}'] = f.detach().numpy()
            save[f'grad{k}'] = f.grad.numpy()
        np.savez_compressed(OUT / f'loss_{tag}.npz', **save)
        print('loss', tag, save['terms'].tolist())

    # ---------------- encoder: reference compute_event_image + scatter stand-in ----------------
    # (a) fixtures 0+1 as a batch of 2, the way collate_wrapper lays them out
    evs = []
    tss = []
    for i in (0, 1):
        d = read_all(REF / f'tests/data/seq/{i:06d}.hdf5')
        e = d['events']
        t0 = float(d['start'])
        evs.append({'x': e[:, 0].astype(np.int64), 'y': e[:, 1].astype(np.int64),
                    'timestamp': (e[:, 2] - t0).astype(np.float32),
                    'polarity': e[:, 3].astype(np.int64),
                    'sample_index': np.full(e.shape[0], i, np.int64)})
        tss += [0.0, float(d['stop']) - t0]
    evd = {k: torch.tensor(np.concatenate([e[k] for e in evs])) for k in evs[0]}
    ts = torch.tensor(tss, dtype=torch.float32)
    sidx = torch.LongTensor([0, 0, 1, 1])
    res = ref_net.compute_event_image(evd, ts, sidx, (272, 352))
    print('encoder fixture sums', res.sum((0, 2, 3)).tolist())
    np.savez_compressed(OUT / 'encoder_fixture.npz', timestamps=ts.numpy(), sample_idx=sidx.numpy(),
                        H=272, W=352, out=res.numpy(), **{f'ev_{k}': v.numpy() for k, v in evd.items()})
    # (b) synthetic: dense collisions, ragged (incl. an empty sample)
    for tag, (B, H, W, n, ragged) in {'dense': (3, 32, 48, 20000, False),
                                      'ragged': (5, 48, 64, 3000, True)}.items():
        g = torch.Generator().manual_seed(99 + len(tag))
        evd, ts, sidx = synth_events(g, B, H, W, n, ragged=ragged)
        res = ref_net.compute_event_image(evd, ts, sidx, (H, W))
        np.savez_compressed(OUT / f'encoder_{tag}.npz', timestamps=ts.numpy(), sample_idx=sidx.numpy(),
                            H=H, W=W, out=res.numpy(),
                            **{f'ev_{k}': v.numpy() for k, v in evd.items()})
        print('encoder', tag, res.shape, float(res.sum()))

    # ---------------- transforms: EventCrop, get_count_image, RandomRotation ----------------
    d = read_all(REF / 'tests/data/seq/000002.hdf5')
    e = d['events']
    box = [20, 30, 200, 256]
    cropped = ref_data.EventCrop(box)(e.copy())
    cnt = ref_data.get_count_image([cropped[:, 0], cropped[:, 1]], (200, 256))
    imgs = np.stack([d['image1'], d['image2']])
    e5 = np.hstack([e, np.zeros((e.shape[0], 1))])
    rot = ref_data.RandomRotation(30, (260, 346))
    out = {'events': e, 'box': np.array(box), 'cropped': cropped, 'count': cnt, 'images': imgs}
    for ang in (0.0, 90.0, 17.5):
        rim, rev, _ = rot(imgs, e5.copy(), angle=ang)
        out[f'rot_images_{ang}'] = rim
        out[f'rot_events_{ang}'] = rev
        print('rot', ang, rev.shape)
    np.savez_compressed(OUT / 'transforms.npz', **out)


if __name__ == '__main__':
    main()
