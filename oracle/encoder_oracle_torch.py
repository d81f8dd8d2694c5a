"""TEST INFRASTRUCTURE -- torch-CPU restatement of the reference encoder, used ONLY for the
timed CPU baseline (`bench.py` cpu_baseline / `--impl reference`) and checked against the numpy
oracle in tests/test_oracle_golden.py.

It issues the same ATen operator sequence the reference's CPU path runs
(EV_FlowNet/net.py:113-150: unique/cumsum bookkeeping, the 4 assert reductions, elementwise
normalisation, then the two scatters -- `torch_scatter.scatter_sum/scatter_max` with `out=`
being `scatter_add_` / `scatter_reduce_(amax, include_self=True)` on CPU), so its wall-clock on
the box's host cores is a fair stand-in for the reference, which cannot travel to the GPU box.
"""
import torch


def compute_event_image(events, timestamps, sample_idx, imsize):
    start = timestamps[::2]
    bs = start.numel()
    H, W = imsize
    res = torch.zeros(bs * 4 * H * W, dtype=torch.float32)
    x, y, t = events['x'], events['y'], events['timestamp']
    p, b = events['polarity'], events['sample_index']
    if x.numel() == 0:
        return res.view(bs, 4, H, W)
    assert bool((p.abs() == 1).all())
    uniq_b, counts = torch.unique(b, return_counts=True)
    last = torch.cumsum(counts, 0) - 1
    assert bool((b[last] == uniq_b).all())
    assert bool((b[last[:-1] + 1] > uniq_b[:-1]).all())
    assert int(last[-1]) + 1 == b.numel()
    dt = torch.zeros(bs, dtype=torch.float32)
    dt[uniq_b] = t[last] - start[uniq_b]
    dt.clamp_(min=1e-9)
    tn = (t - start[b]) / dt[b]
    c = torch.div(1 - p, 2, rounding_mode='floor')
    idx = ((b * 4 + c) * H + y) * W + x
    res.scatter_add_(0, idx, torch.ones(idx.numel(), dtype=torch.float32))
    res.scatter_reduce_(0, idx + 2 * H * W, tn, 'amax', include_self=True)
    return res.view(bs, 4, H, W)
