"""Experiment harness (GPU box): time the encoder scatter kernel on cfg2/cfg3/cfg4 for the
variant selected by DVSOF_ENCODE_VARIANT, and check the image against a torch scatter on GPU."""
import json
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402


def torch_ref(ev, ts, B, H, W):
    start = ts[::2]
    b, t = ev['sample_index'], ev['timestamp']
    n = b.numel() // B
    last = torch.arange(1, B + 1, device=b.device) * n - 1
    dt = (t[last] - start).clamp_(min=1e-9)
    tn = (t - start[b]) / dt[b]
    c = (1 - ev['polarity']) // 2
    idx = ((b * 4 + c) * H + ev['y']) * W + ev['x']
    res = torch.zeros(B * 4 * H * W, device=b.device)
    res.scatter_add_(0, idx, torch.ones_like(tn))
    res.scatter_reduce_(0, idx + 2 * H * W, tn, 'amax', include_self=True)
    return res.view(B, 4, H, W)


def main():
    dev = torch.device('cuda', 0)
    dv._lib.Status.mode = 'deferred'
    variant = os.environ.get('DVSOF_ENCODE_VARIANT', '0')
    res = {'variant': variant}
    if sys.argv[1:2] == ['sweep']:
        for n in (16384, 32768, 65536, 131072, 262144, 524288, 1048576, 4194304):
            bench.CFG['cfg9'] = dict(B=8, H=256, W=256, n=n, T=0.032)
            r = bench.bench_encoder('cfg9', 10, 3, dev)
            res[f'n{8 * n}'] = [round(r['kernel_ms'] * 1e3, 2), round(r['step_ms'] * 1e3, 2)]   # scatter kernel, whole call (us)
        print(json.dumps(res))
        return
    for name in sys.argv[1:] or ['cfg2', 'cfg3', 'cfg4']:
        c = bench.CFG[name]
        r = bench.bench_encoder(name, 10, 3, dev)
        ev, ts, sidx = bench.fast_events(1234 + int(name[3:]), c['B'], c['H'], c['W'], c['n'], c['T'], dev)
        size = dv.extend_size((c['H'], c['W']))
        got = dv.compute_event_image(ev, ts, sidx, size)
        ok = bool(torch.equal(got, torch_ref(ev, ts, c['B'], *size)))
        res[name] = {'kernel_ms': round(r['kernel_ms'], 4), 'step_ms': round(r['step_ms'], 4),
                     'GBs': round(r['bytes'] / r['kernel_ms'] / 1e6, 1),
                     'Gev_s': round(r['events'] / r['kernel_ms'] / 1e6, 2), 'exact': ok}
        del ev, got
        torch.cuda.empty_cache()
    print(json.dumps(res))


if __name__ == '__main__':
    main()
