"""TEST INFRASTRUCTURE -- CPU oracle (torch fp32, device='cpu') for the self-supervised
multiscale warp loss of the reference (`utils/loss.py`).

Used ONLY as the checker by `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
cpu_baseline / `--impl reference` legs; the product package never imports it.

Parity pin: (1) the reference's own known-answer tests `tests/loss/test_loss.py:20,41,63`
(0.002/0.002/0; 0.002/0.622660/0; 0.002120/0.652659/0.007802, +-5e-6), restated in
tests/test_loss_oracle.py on the committed fixture tests/golden/fixture_000001.npz;
(2) values AND gradients of the reference's `Losses` run in the build container on
seeded random inputs (tests/golden/make_golden.py -> tests/golden/loss_*.npz).

A floating-point path: this is the "torch fp32 reference" the task allows for a
floating-point kernel.  Functional restatement; autograd of it is the gradient oracle.
"""
import torch
import torch.nn.functional as F

ALPHA = 0.45
EPS = 1e-3


def charbonnier(delta):
    """(delta^2 + eps^2)^alpha elementwise -- utils/loss.py:31."""
    return (delta.pow(2) + EPS * EPS).pow(ALPHA)


def resize(images, shape):
    """utils/loss.py:20-21: bilinear, align_corners=True."""
    return F.interpolate(images, size=tuple(shape), mode='bilinear', align_corners=True)


def normalised_grid(flow):
    """utils/loss.py:150-156: g = (pix + flow) / ((S-1)/2) - 1, IEEE division, x then y."""
    n, _, h, w = flow.shape
    ys = torch.arange(h, dtype=torch.float32).view(1, h, 1).expand(n, h, w)
    xs = torch.arange(w, dtype=torch.float32).view(1, 1, w).expand(n, h, w)
    gx = (xs + flow[:, 0]) / ((w - 1) / 2.) - 1
    gy = (ys + flow[:, 1]) / ((h - 1) / 2.) - 1
    return torch.stack((gx, gy), dim=1)


def photometric(prev_images, next_images, grid):
    """utils/loss.py:58-74: warp `next` by the grid, Charbonnier mean of (warped - prev)."""
    warped = F.grid_sample(next_images, grid.permute(0, 2, 3, 1), mode='bilinear',
                           padding_mode='zeros', align_corners=True)
    return charbonnier(warped - prev_images).mean()


def smoothness(flow):
    """utils/loss.py:76-90: 4 neighbour directions, each its own mean, averaged."""
    horiz = flow[..., :, 1:] - flow[..., :, :-1]
    vert = flow[..., 1:, :] - flow[..., :-1, :]
    diag_a = flow[..., 1:, 1:] - flow[..., :-1, :-1]
    diag_b = flow[..., :-1, 1:] - flow[..., 1:, :-1]
    return sum(charbonnier(d).mean() if d.numel() else d.new_zeros(())
               for d in (horiz, vert, diag_a, diag_b)) / 4


def outborder(flow, grid):
    """utils/loss.py:92-119: sum over out-of-border pixels of c(flow)/(2 n_i N)."""
    n = flow.shape[0]
    with torch.no_grad():
        bad = ((grid < -1) | (grid > 1)).any(dim=1)                 # (N,H,W)
        n_bad = bad.view(n, -1).sum(dim=1)                          # n_i
    if int(n_bad.sum()) == 0:
        return torch.tensor(0, dtype=torch.float32)
    denom = (2 * n_bad * n).to(torch.float32).view(n, 1, 1, 1).clamp(min=1)
    contrib = charbonnier(flow) / denom
    return (contrib * bad.unsqueeze(1)).sum()


def scale_loss(prev_images, next_images, flow):
    """One scale: returns (smoothness, photometric, outborder) -- utils/loss.py:121-171."""
    grid = normalised_grid(flow)
    return (smoothness(flow), photometric(prev_images, next_images, grid),
            outborder(flow, grid))


def match_images(flow_ts, flow_sample_idx, timestamps, sample_idx):
    """utils/loss.py:182-206: exact fp32 timestamp equality within the same sample."""
    P = flow_sample_idx.shape[0]
    same = sample_idx.view(-1, 1) == flow_sample_idx.view(1, -1)          # (D,P)
    out = []
    for k in range(2):
        hit = (timestamps.view(-1, 1) == flow_ts[:, k].view(1, -1)) & same
        assert (hit.sum(0) == 1).all()
        out.append(hit.to(torch.int64).argmax(dim=0))
    assert out[0].numel() == P
    return out[0], out[1]


def multiscale_loss(flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx):
    """utils/loss.py:179-214.  The image pyramid is CHAINED (line 210): scale k is resized
    from scale k-1.  Returns ((smooth...), (photo...), (outborder...)) per scale."""
    start, stop = match_images(flow_ts, flow_sample_idx, timestamps, sample_idx)
    per_scale = []
    for flow in flows:
        with torch.no_grad():
            images = resize(images, flow.shape[-2:])
        per_scale.append(scale_loss(images[start], images[stop], flow))
    return tuple(zip(*per_scale))


def combined(terms, weights=(0.5, 1.0, 1.0)):
    """utils/training.py:12-24: sum_k w_k * mean_s(term_k,s); order (smooth, photo, outborder)."""
    total = 0
    for w, t in zip(weights, terms):
        total = total + w * sum(t) / len(t)
    return total
