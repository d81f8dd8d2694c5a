"""Shared synthetic-input generators for tests and bench."""
import numpy as np
import torch


def synth_events(seed, B, H, W, n_per_sample, T=0.032, contended=False, device='cpu',
                 ragged=False):
    """Per sample: x~U{0..W-1}, y~U{0..H-1}, p=+-1 with P(+1)=0.7, t sorted U[0,T);
    grouped by ascending sample_index; timestamps=[0,T]*B; sample_idx=[0,0,1,1,...]."""
    g = torch.Generator().manual_seed(seed)
    xs, ys, ts, ps, bs = [], [], [], [], []
    for b in range(B):
        n = n_per_sample
        if ragged:
            n = int(torch.randint(0, n_per_sample + 1, (1,), generator=g))
            if b == 1:
                n = 0
        x = torch.randint(0, W, (n,), generator=g)
        y = torch.randint(0, H, (n,), generator=g)
        if contended:  # 80 % of events on a random 10 % of pixels
            hot = torch.randperm(H * W, generator=g)[:max(1, H * W // 10)]
            pick = hot[torch.randint(0, hot.numel(), (n,), generator=g)]
            use = torch.rand(n, generator=g) < 0.8
            x = torch.where(use, pick % W, x)
            y = torch.where(use, pick // W, y)
        xs.append(x)
        ys.append(y)
        ts.append(torch.sort(torch.rand(n, generator=g) * T)[0].float())
        ps.append((torch.rand(n, generator=g) < 0.7).long() * 2 - 1)
        bs.append(torch.full((n,), b, dtype=torch.long))
    ev = {'x': torch.cat(xs), 'y': torch.cat(ys), 'timestamp': torch.cat(ts),
          'polarity': torch.cat(ps), 'sample_index': torch.cat(bs)}
    ev['element_index'] = torch.zeros_like(ev['x'])
    ev = {k: v.to(device) for k, v in ev.items()}
    timestamps = torch.tensor([0, T] * B, dtype=torch.float32, device=device)
    sample_idx = torch.arange(B, device=device).repeat_interleave(2)
    return ev, timestamps, sample_idx


def synth_loss_inputs(seed, B, H, W, sigma=3.0, device='cpu', n_scales=4):
    g = torch.Generator().manual_seed(seed)
    images = torch.randint(0, 256, (2 * B, 1, H, W), generator=g).float()
    ts = torch.tensor([0.0, 0.032] * B)
    sidx = torch.arange(B).repeat_interleave(2)
    ks = list(range(n_scales - 1, -1, -1))
    shapes = [(H // 2 ** k, W // 2 ** k) for k in ks]
    flows = [(torch.randn(B, 2, h, w, generator=g) * sigma / 2 ** k)
             for k, (h, w) in zip(ks, shapes)]
    out = dict(images=images, timestamps=ts, sample_idx=sidx, flow_ts=ts.view(-1, 2).clone(),
               flow_sample_idx=torch.arange(B), flows=flows, shapes=shapes)
    return {k: ([t.to(device) for t in v] if k == 'flows' else
                (v.to(device) if isinstance(v, torch.Tensor) else v)) for k, v in out.items()}


def np_events(ev):
    return {k: v.cpu().numpy() for k, v in ev.items()}
