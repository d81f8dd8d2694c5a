"""TEST INFRASTRUCTURE -- CPU oracle for the voxel-grid EXTENSION (`dv.compute_voxel_grid`,
csrc/voxel.cuh).  Used only by `tests/` and `bench.py`'s cpu_baseline leg.

PARITY UNPINNED: the reference has no voxel grid (its encoder is the 4-channel event image of
EV_FlowNet/net.py:88-150, restated in encoder_oracle.py), so there is no reference output, test or
golden vector to pin this against.  It restates the published definition the "bins" wording of
BASELINE.json points at -- Zhu, Yuan, Chaney, Daniilidis, "Unsupervised Event-based Learning of
Optical Flow, Depth and Egomotion", CVPR 2019, eq. (1)-(3):

    t*_i = (C - 1) (t_i - t_0) / (t_N - t_0)
    V(x, y, k) = sum_i p_i  max(0, 1 - |x - x_i|)  max(0, 1 - |y - y_i|)  max(0, 1 - |k - t*_i|)

(integer pixel coordinates: the spatial kernels are 1 at the event's own pixel) -- with t_0 and
t_N - t_0 taken per sample exactly as the reference normalises its timestamp planes
(net.py:136-141: start = timestamps[2b], dt = max(t[last event of b] - start, 1e-9)).

Per-event arithmetic is fp32, one rounding per operation, as the kernels do it; the sums are taken
in float64, so `voxel_grid` is the exact sum of the fp32 weights (rounded once at the end) and
`abs_sum` its conditioning -- the fp32 accumulation on the device may differ by ~eps * abs_sum."""
import numpy as np

F32 = np.float32


def voxel_grid(events, timestamps, sample_idx, imsize, n_bins, with_abs_sum=False):
    timestamps = np.asarray(timestamps, dtype=F32)
    start = timestamps[::2]
    bs = start.size
    H, W = imsize
    C = int(n_bins)
    acc = np.zeros(bs * C * H * W, dtype=np.float64)
    mag = np.zeros_like(acc)
    x = np.asarray(events['x'], dtype=np.int64)
    if x.size:
        y = np.asarray(events['y'], dtype=np.int64)
        t = np.asarray(events['timestamp'], dtype=F32)
        p = np.asarray(events['polarity'], dtype=np.int64)
        b = np.asarray(events['sample_index'], dtype=np.int64)
        assert (np.abs(p) == 1).all() and (np.diff(b) >= 0).all()
        uniq_b, num = np.unique(b, return_counts=True)
        last = np.cumsum(num) - 1
        dt = np.zeros(bs, dtype=F32)
        dt[uniq_b] = t[last] - start[uniq_b]                      # net.py:136-138
        dt = np.maximum(dt, F32(1e-9))
        tn = ((t - start[b]).astype(F32) / dt[b]).astype(F32)      # net.py:141
        ts = (tn * F32(C - 1)).astype(F32)
        fl = np.floor(ts)
        f = (ts - fl).astype(F32)
        sign = p.astype(F32)
        w0 = (sign * (F32(1) - f).astype(F32)).astype(F32)
        w1 = (sign * f).astype(F32)
        k0 = np.where(np.isfinite(fl), np.clip(fl, -2, C + 1), -2).astype(np.int64)
        plane = H * W
        base = b * (C * plane) + y * W + x
        for k, w in ((k0, w0), (k0 + 1, w1)):
            ok = (k >= 0) & (k < C)
            np.add.at(acc, base[ok] + k[ok] * plane, w[ok].astype(np.float64))
            np.add.at(mag, base[ok] + k[ok] * plane, np.abs(w[ok]).astype(np.float64))
    out = acc.astype(F32).reshape(bs, C, H, W)
    return (out, mag.reshape(bs, C, H, W)) if with_abs_sum else out
