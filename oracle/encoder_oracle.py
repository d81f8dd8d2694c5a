"""TEST INFRASTRUCTURE -- CPU oracle for the event-image encoder and the event transforms.

This is a numpy restatement of the reference's algorithm, used ONLY as the checker by
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs.  Nothing under `dvs_of_training_framework_b200/` may import it.

Parity pin: the reference's tests hold NO value assertion for `compute_event_image`
(SURVEY.md F2); this restatement is pinned against outputs of the reference's own
`EV_FlowNet/net.py::compute_event_image` run in the build container with a
`torch_scatter` stand-in (tests/golden/make_golden.py -> tests/golden/encoder_*.npz).
`map_events` / `EventCrop` / `get_count_image` are pinned by the reference's exact
equality + invariance tests (tests/dataset/test_dataset.py:42-73,122-170,173-218),
restated in tests/test_transforms_oracle.py.

Third-party dependency on the path: PyPI `torch-scatter` (unpinned in
requirements.txt:11; the reference test image resolves the torch-1.10.0+cpu wheel,
Dockerfile_test:24).  Published semantics relied on (call sites EV_FlowNet/net.py:147-149):
  scatter_sum(src, index, out=res): res[index[i]] += src[i]
  scatter_max(src, index, out=res): res[index[i]] = max(res[index[i]], src[i]) (the
  existing zeros in `out` take part in the max), returns (out, argmax) -- argmax unused.
"""
import numpy as np

F32 = np.float32


def compute_event_image(events, timestamps, sample_idx, imsize):
    """events -> (B, 4, H, W) fp32 [count(p=+1), count(p=-1), max t^(p=+1), max t^(p=-1)].

    Follows EV_FlowNet/net.py:88-150 step by step.
    events: dict of 1-D arrays x, y, polarity, sample_index (int64), timestamp (fp32).
    """
    timestamps = np.asarray(timestamps, dtype=F32)
    sample_idx = np.asarray(sample_idx)
    assert timestamps.ndim == 1 and sample_idx.ndim == 1          # net.py:99-100
    assert (int(sample_idx[-1]) + 1) * 2 == timestamps.size        # net.py:101
    start = timestamps[::2]                                        # net.py:102 (stop unused)
    bs = start.size
    H, W = imsize
    res = np.zeros(bs * 4 * H * W, dtype=F32)                      # net.py:113-114
    x = np.asarray(events['x'], dtype=np.int64)
    if x.size == 0:                                                # net.py:115-116
        return res.reshape(bs, 4, H, W)
    y = np.asarray(events['y'], dtype=np.int64)
    t = np.asarray(events['timestamp'], dtype=F32)
    p = np.asarray(events['polarity'], dtype=np.int64)
    b = np.asarray(events['sample_index'], dtype=np.int64)
    assert (np.abs(p) == 1).all()                                  # net.py:124

    # index of the last event of every sample present (net.py:127-134)
    uniq_b, num_events = np.unique(b, return_counts=True)
    last = np.cumsum(num_events) - 1
    assert (b[last] == uniq_b).all()
    assert (b[last[:-1] + 1] > uniq_b[:-1]).all()
    assert last[-1] + 1 == b.size

    dt = np.zeros(bs, dtype=F32)                                   # net.py:136-138
    dt[uniq_b] = t[last] - start[uniq_b]
    dt = np.maximum(dt, F32(1e-9))

    tn = ((t - start[b]) / dt[b]).astype(F32)                      # net.py:141
    c = (1 - p) // 2                                               # net.py:142  (+1 -> 0, -1 -> 1)
    idx = ((b * 4 + c) * H + y) * W + x                            # net.py:144
    np.add.at(res, idx, F32(1))                                    # scatter_sum, net.py:147-148
    np.maximum.at(res, idx + 2 * H * W, tn)                        # scatter_max, net.py:149
    return res.reshape(bs, 4, H, W)


def extend_size(imsize):
    """EV_FlowNet/net.py:194-195: each side rounded up to a multiple of 16."""
    return tuple(((s - 1) // 16 + 1) * 16 for s in imsize)


def get_count_image(events, imsize):
    """utils/data.py:120-136: per-pixel event count, uint64 (H, W)."""
    x = np.asarray(events[0]).astype(int)
    y = np.asarray(events[1]).astype(int)
    H, W = imsize
    assert ((x >= 0) & (x < W) & (y >= 0) & (y < H)).all()  # ravel_multi_index raises
    return np.bincount(y * W + x, minlength=H * W).astype(np.uint64).reshape(H, W)


def event_crop(events, box):
    """utils/data.py:24-42: keep events inside box=[y0,x0,h,w], shift to box origin.

    events: (N, >=2) array, columns x, y, ...; order preserved.
    """
    y0, x0, h, w = box
    x, y = events[:, 0], events[:, 1]
    keep = (x >= x0) & (x < x0 + w) & (y >= y0) & (y < y0 + h)
    out = events[keep].copy()
    out[:, 0] -= x0
    out[:, 1] -= y0
    return out


def map_events(events, shape, src, dst):
    """transformation.cpp:35-85: one-to-many pixel remap of an event array.

    For each event (input order) and each LUT entry k with src[k] == pixel(event)
    (LUT order), emit a row with (x, y) = unravel(dst[k]) and columns 2.. copied.
    pixel(event) = int(y) * W + int(x) (float -> int truncation, :62,:74).
    """
    events = np.asarray(events)
    src = np.asarray(src)
    dst = np.asarray(dst)
    if src.shape[0] != dst.shape[0]:
        raise ValueError('src and dst should have the same size!')   # :41-42
    assert events.dtype == np.float32 and events.ndim == 2
    H, W = shape
    n, E = events.shape
    # stable grouping of LUT entries by src pixel == insertion order per key (:50-54)
    order = np.argsort(src, kind='stable')
    ssrc = src[order].astype(np.int64)
    sdst = dst[order].astype(np.int64)
    key = events[:, 1].astype(np.int32).astype(np.int64) * W + \
        events[:, 0].astype(np.int32).astype(np.int64)
    lo = np.searchsorted(ssrc, key, side='left')
    hi = np.searchsorted(ssrc, key, side='right')
    cnt = hi - lo
    total = int(cnt.sum())
    out = np.zeros((total, E), dtype=np.float32)
    if total == 0:
        return out
    ev_of_row = np.repeat(np.arange(n), cnt)
    first = np.cumsum(cnt) - cnt
    within = np.arange(total) - np.repeat(first, cnt)
    d = sdst[np.repeat(lo, cnt) + within]
    out[:, 0] = (d % W).astype(np.float32)
    out[:, 1] = (d // W).astype(np.float32)
    out[:, 2:] = events[ev_of_row, 2:]
    return out


def rotation_lut(shape, angle):
    """utils/data.py:155-205: nearest-neighbour rotation LUT (src=ridx, dst=idx)."""
    H, W = shape
    xs, ys = np.meshgrid(range(W), range(H))
    xs, ys = xs.ravel(), ys.ravel()
    idx = ys * W + xs
    xc = xs.astype(float) - W / 2
    yc = ys.astype(float) - H / 2
    rad = angle * np.pi / 180
    mat = np.array([[np.cos(rad), -np.sin(rad)], [np.sin(rad), np.cos(rad)]])
    r = mat.dot(np.vstack((xc[None], yc[None])))
    x1 = np.rint(r[0] + W / 2).astype(int)
    y1 = np.rint(r[1] + H / 2).astype(int)
    m = (x1 >= 0) & (x1 < W) & (y1 >= 0) & (y1 < H)
    return (y1[m] * W + x1[m]).astype(np.uint64), idx[m].astype(np.uint64)


def rotate(images, events, shape, angle):
    """utils/data.py:180-218: rotate images (C,H,W) and events (N,E) by `angle` degrees."""
    ridx, idx = rotation_lut(shape, angle)
    C = images.shape[0]
    flat = images.reshape(C, -1)
    out = np.zeros_like(flat)
    out[:, idx.astype(np.int64)] = flat[:, ridx.astype(np.int64)]
    rev = map_events(events.astype(np.float32).copy(), shape, ridx, idx)
    return out.reshape(images.shape), rev


def decode_events(x16, y16, pbool, t, events_per_element, elements_per_sample):
    """utils/dataset.py:335-373 (decode_batch, event part): storage form -> training form."""
    epe = np.asarray(events_per_element, dtype=np.int64)
    eps = np.asarray(elements_per_sample, dtype=np.int64)
    # :357-359 slices events_per_element by the prefix of elements_per_sample: elements past the
    # end of a shorter events_per_element (trailing empty ones, dropped by encode_batch :298) do
    # not exist
    sample_of_element = np.repeat(np.arange(eps.size), eps)[:epe.size]
    elem_in_sample = np.arange(epe.size) - np.repeat(np.cumsum(eps) - eps, eps)[:epe.size]
    return {'x': np.asarray(x16).astype(np.int64),
            'y': np.asarray(y16).astype(np.int64),
            'timestamp': np.asarray(t, dtype=F32),
            'polarity': np.asarray(pbool).astype(np.int64) * 2 - 1,
            'element_index': np.repeat(elem_in_sample, epe),
            'sample_index': np.repeat(sample_of_element, epe)}


def encode_events(x, y, polarity, element_index, sample_index, elements_per_sample):
    """utils/dataset.py:286-302 (encode_batch, event part): training form -> storage form.  The
    number of elements is taken from the LAST event (:298), so trailing empty elements vanish."""
    s = np.asarray(sample_index).astype(np.int16)                                 # :291
    shift = np.cumsum(np.concatenate([[0], np.asarray(elements_per_sample, dtype=np.int64)]))   # :294-296
    new_e = np.asarray(element_index, dtype=np.int64) + shift[s]                  # :297
    total = int(new_e[-1]) + 1                                                    # :298
    epe = np.zeros(total, dtype=np.int64)
    np.add.at(epe, new_e, 1)                                                      # :300-301
    return {'x': np.asarray(x).astype(np.int16), 'y': np.asarray(y).astype(np.int16),
            'polarity': ((np.asarray(polarity) + 1) / 2).astype(bool),            # :289
            'events_per_element': epe}


def elements_per_sample(sample_idx, size):
    """utils/dataset.py:231-233 (encode_batch_info): images per sample minus one, as uint8."""
    eps = np.zeros(size, dtype=np.int16) - 1
    np.add.at(eps, np.asarray(sample_idx, dtype=np.int64), 1)
    return eps.astype(np.uint8)
