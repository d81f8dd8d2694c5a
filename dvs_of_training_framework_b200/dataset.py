"""Batch codec of the reference's data pipeline on device: utils/dataset.py
`encode_batch_info` (:201-237), `encode_batch` (:240-305), `decode_batch_info` (:308-332) and
`decode_batch` (:335-373) with the same names, arguments and result dictionaries -- the tensors are
CUDA tensors and the per-event work (narrowing / widening, polarity mapping, the histogram and the
run-length expansion the reference does with numpy and Python loops) runs in libdvsof.

The 9 B/event storage form can also be fed to the encoder directly
(`encoder.compute_event_image_compact`), which skips the int64 form altogether."""
import torch

from . import _lib


def _cumsum_with_prefix(v):
    out = torch.zeros(v.numel() + 1, dtype=torch.int64, device=v.device)
    torch.cumsum(v.to(torch.int64), 0, out=out[1:])
    return out


def encode_batch_info(timestamps, sample_idx, images, augmentation_params, size):
    """utils/dataset.py:201-237: elements_per_sample = (images of the sample) - 1 as uint8, images
    as uint8."""
    sample_idx = _lib.require_cuda(sample_idx, 'sample_idx')
    counts = torch.bincount(sample_idx.to(torch.int64), minlength=int(size))
    elements_per_sample = (counts.to(torch.int16) - 1).to(torch.uint8)          # :231-233
    return {'timestamps': timestamps,
            'elements_per_sample': elements_per_sample,
            'images': images.to(torch.uint8),
            'augmentation_params': augmentation_params}


def encode_batch(events, timestamps, sample_idx, images, augmentation_params, size):
    """utils/dataset.py:240-305.  events: dict of CUDA tensors {x, y, timestamp, polarity,
    element_index, sample_index} (the training form).  Like the reference, the number of elements
    is read from the LAST event (`int(new_e[-1]) + 1`, :298): one two-value D2H read."""
    result = encode_batch_info(timestamps, sample_idx, images, augmentation_params, size)
    x = _lib.require_cuda(events['x'], 'x', torch.int64)
    dev = x.device
    y = _lib.require_cuda(events['y'], 'y', torch.int64)
    p = _lib.require_cuda(events['polarity'], 'polarity', torch.int64)
    e = _lib.require_cuda(events['element_index'], 'element_index', torch.int64)
    s = _lib.require_cuda(events['sample_index'], 'sample_index', torch.int64)
    n = x.numel()
    assert y.numel() == n and p.numel() == n and e.numel() == n and s.numel() == n
    eps = result['elements_per_sample'].to(torch.int64)
    bs = eps.numel()
    assert n > 0, 'encode_batch: the reference indexes the last event (IndexError on an empty batch)'
    last = torch.stack([e[-1], s[-1]]).tolist()                                   # :298
    shift = _cumsum_with_prefix(eps)
    last_s = ((last[1] + 2 ** 15) % 2 ** 16) - 2 ** 15                            # .to(torch.short), :291
    total_elements = int(last[0] + shift[last_s].item()) + 1
    ox = torch.empty(n, dtype=torch.int16, device=dev)
    oy = torch.empty(n, dtype=torch.int16, device=dev)
    op = torch.empty(n, dtype=torch.bool, device=dev)
    epe = torch.empty(max(total_elements, 0), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        st = _lib.status_for(dev)
        st.flush(IndexError)
        lib = _lib.load()
        ws = torch.empty(lib.dvsof_codec_workspace_bytes(epe.numel(), bs), dtype=torch.uint8, device=dev)
        _lib.call('dvsof_encode_events', _lib.ptr(x), _lib.ptr(y), _lib.ptr(p), _lib.ptr(e), _lib.ptr(s), n,
                  _lib.ptr(eps), bs, epe.numel(), _lib.ptr(ox), _lib.ptr(oy), _lib.ptr(op.view(torch.uint8)),
                  _lib.ptr(epe), _lib.ptr(ws), _lib.ptr(st.word), _lib.stream_ptr(dev))
        st.after_call(IndexError)
    result['events'] = {'x': ox, 'y': oy, 'timestamp': events['timestamp'], 'polarity': op,
                        'events_per_element': epe}
    return result


def decode_batch_info(encoded_batch_info):
    """utils/dataset.py:308-332."""
    timestamps = encoded_batch_info['timestamps']
    eps = encoded_batch_info['elements_per_sample']
    batch_size = eps.numel()
    # :322-324; a sample with e elements has e + 1 frames, so the length is known from `timestamps`
    # (without output_size= the op reads it back from the device: a host sync per batch)
    sample_idx = torch.repeat_interleave(torch.arange(batch_size, dtype=torch.int64, device=eps.device),
                                         eps.to(torch.int64) + 1, output_size=timestamps.numel())
    return {'timestamps': timestamps.to(torch.float32),
            'sample_idx': sample_idx,
            'images': encoded_batch_info['images'].to(torch.float32),
            'augmentation_params': encoded_batch_info['augmentation_params'],
            'size': batch_size}


def decode_events(events, elements_per_sample):
    """The event part of decode_batch (utils/dataset.py:349-372) in one device pass."""
    x = _lib.require_cuda(events['x'], 'x', torch.int16)
    dev = x.device
    y = _lib.require_cuda(events['y'], 'y', torch.int16)
    p = _lib.require_cuda(events['polarity'], 'polarity')
    if p.dtype == torch.bool:
        p = p.view(torch.uint8)
    assert p.dtype == torch.uint8
    epe = _lib.require_cuda(events['events_per_element'], 'events_per_element').to(torch.int64)
    eps = _lib.require_cuda(elements_per_sample, 'elements_per_sample').to(torch.int64)
    n = x.numel()
    assert y.numel() == n and p.numel() == n
    out = {k: torch.empty(n, dtype=torch.int64, device=dev)
           for k in ('x', 'y', 'polarity', 'element_index', 'sample_index')}
    with torch.cuda.device(dev):
        st = _lib.status_for(dev)
        st.flush()
        lib = _lib.load()
        ws = torch.empty(lib.dvsof_codec_workspace_bytes(epe.numel(), eps.numel()), dtype=torch.uint8, device=dev)
        _lib.call('dvsof_decode_events', _lib.ptr(x), _lib.ptr(y), _lib.ptr(p), n, _lib.ptr(epe), epe.numel(),
                  _lib.ptr(eps), eps.numel(), _lib.ptr(out['x']), _lib.ptr(out['y']), _lib.ptr(out['polarity']),
                  _lib.ptr(out['element_index']), _lib.ptr(out['sample_index']), _lib.ptr(ws),
                  _lib.ptr(st.word), _lib.stream_ptr(dev))
        st.after_call()
    return {'x': out['x'], 'y': out['y'], 'timestamp': events['timestamp'], 'polarity': out['polarity'],
            'element_index': out['element_index'], 'sample_index': out['sample_index']}


def decode_batch(encoded_batch):
    """utils/dataset.py:335-373: the batch in the training form (the input format of encode_batch)."""
    result = decode_batch_info(encoded_batch)
    result['events'] = decode_events(encoded_batch['events'], encoded_batch['elements_per_sample'])
    return result


def quantize_batch(batch, imsize=None):
    """The loop body of scripts/quantize_preprocessed.py:77-92: the batch with
    its events replaced by the event image `data` = quantize(events, ...) (EV_FlowNet/net.py:197-205).
    Accepts the training form (events with int64 x, y, ...) and the storage form (`encode_batch`
    output: the 9 B/event arrays go straight into the encoder, no int64 detour)."""
    from . import encoder
    if 'elements_per_sample' in batch:                      # storage form
        info = decode_batch_info(batch)
        ev = batch['events']
        eps = batch['elements_per_sample'].to(torch.int64)
        epe = ev['events_per_element'].to(torch.int64)
        # events per sample from the run lengths (elements may be truncated at the end, :298)
        elem_sample = torch.repeat_interleave(torch.arange(eps.numel(), device=eps.device), eps)[:epe.numel()]
        per_sample = torch.zeros(eps.numel(), dtype=torch.int64, device=eps.device).index_add_(0, elem_sample, epe)
        imsize = imsize if imsize is not None else batch['images'].shape[-2:]
        # EV_FlowNet/net.py:101-103: two image time stamps per sample (one element per sample)
        ts = info['timestamps']
        assert ts.numel() == 2 * eps.numel()
        data = encoder.compute_event_image_compact(ev['x'], ev['y'], ev['timestamp'], ev['polarity'], per_sample,
                                                   ts.contiguous(), encoder.extend_size(imsize))
        out = dict(info)
    else:
        imsize = imsize if imsize is not None else batch['images'].shape[-2:]
        with torch.no_grad():
            data = encoder.quantize(batch['events'], batch['timestamps'], batch['sample_idx'], imsize)
        out = {k: v for k, v in batch.items() if k != 'events'}
    out['data'] = data
    return out


def encode_quantized_batch(batch):
    """utils/dataset.py:429-479: everything to the CPU (one synchronisation for all tensors, through
    pinned buffers), `data` (B, C, H, W) -> (B*C, H, W), channels_per_sample = C as uint8, plus
    encode_batch_info."""
    B, C, H, W = batch['data'].size()
    info = encode_batch_info(batch['timestamps'], batch['sample_idx'], batch['images'],
                             batch['augmentation_params'], batch['size']) if batch['sample_idx'].is_cuda else None
    if info is None:                                        # already on the CPU: the reference's own arithmetic
        sidx = batch['sample_idx'].to(torch.int64)
        eps = (torch.bincount(sidx, minlength=int(batch['size'])).to(torch.int16) - 1).to(torch.uint8)
        info = {'timestamps': batch['timestamps'], 'elements_per_sample': eps,
                'images': batch['images'].to(torch.uint8), 'augmentation_params': batch['augmentation_params']}
    result = {'data': batch['data'].reshape(B * C, H, W),
              'channels_per_sample': torch.full((B,), C, dtype=torch.uint8)}
    result.update(info)

    pending = []

    def to_cpu(v):
        if isinstance(v, dict):
            return {k: to_cpu(x) for k, x in v.items()}
        if isinstance(v, torch.Tensor) and v.is_cuda:
            host = torch.empty(v.shape, dtype=v.dtype, pin_memory=True)
            host.copy_(v, non_blocking=True)
            pending.append(v.device)
            return host
        return v
    result = to_cpu(result)
    for dev in set(pending):
        torch.cuda.current_stream(dev).synchronize()
    return result


def decode_quantized_batch(batch):
    """utils/dataset.py:482-502."""
    result = decode_batch_info(batch)
    cps = batch['channels_per_sample']
    assert cps.numel() > 0
    assert bool((cps == cps[0]).all())
    B, C = result['size'], int(cps[0])
    _, H, W = batch['data'].size()
    result['data'] = batch['data'].view(B, C, H, W)
    return result


# ---------------------------------------------------------------------------------------------
# Batch assembly (utils/dataset.py:961-1020 `collate_wrapper`)
# ---------------------------------------------------------------------------------------------
def _pinned(shape, dtype):
    """A host tensor the copy engines can read asynchronously (plain when there is no GPU)."""
    t = torch.empty(shape, dtype=dtype)
    return t.pin_memory() if torch.cuda.is_available() else t


def _augmentation_info(batch):
    """utils/dataset.py:998-1013: the per-sample augmentation tuples as a dict of tensors."""
    import numpy as np
    if not (len(batch) > 0 and len(batch[0]) > 3):
        return None
    aug = [x[3] for x in batch]
    info = {'idx': np.array([a[0] for a in aug]), 'sequence_length': np.array([a[1] for a in aug]),
            'collapse_length': np.array([a[2] for a in aug]),
            'box': np.vstack([np.asarray(a[3]).reshape(1, -1) for a in aug]),
            'angle': np.array([a[4] for a in aug]), 'is_flip': np.array([a[5] for a in aug])}
    # utils/common.py:240-259 `to_tensor`: platform ints -> int64, everything else -> float32
    return {k: torch.tensor(v, dtype=torch.long) if v.dtype == np.int_ else torch.tensor(v, dtype=torch.float32)
            for k, v in info.items()}


def collate_wrapper(batch, encoded=False):
    """Drop-in for the reference's `collate_wrapper` (utils/dataset.py:961-1020): `batch` is a list of
    `(events dict, timestamps, images[, augmentation tuple])` as `DatasetImpl.__getitem__` returns
    them (:776-797) and the result is the same dictionary -- `events` {x, y, polarity, element_index,
    sample_index: int64; timestamp: float32} with `sample_index` synthesised per sample (:974-985),
    `timestamps`, `sample_idx`, `images (D, 1, H, W)` float32, `augmentation_params`, `size` -- as
    pinned host tensors, so the upload of `process_minibatch` is asynchronous.

    encoded=True (`collate_fn=functools.partial(collate_wrapper, encoded=True)`) assembles the SAME
    batch directly in the storage form of `encode_batch` (utils/dataset.py:240-305: x, y int16,
    polarity bool, timestamp float32, `events_per_element`, `elements_per_sample`, uint8 images)
    without the int64 detour: 9 B/event instead of 36 cross PCIe, and `training.DevicePrefetcher`
    (or `decode_batch` on the device) widens it there.  Equal to
    `encode_batch(**collate_wrapper(batch))` of the reference whenever that is defined."""
    import numpy as np
    n_samples = len(batch)
    counts = [int(sample[0]['x'].size) for sample in batch]
    total = sum(counts)
    pin = (lambda t: t.pin_memory()) if torch.cuda.is_available() else (lambda t: t)
    timestamps = pin(torch.tensor(np.hstack([np.asarray(s[1]) for s in batch]), dtype=torch.float32))
    frames = [np.asarray(s[1]).size for s in batch]
    sample_idx = pin(torch.tensor(np.repeat(np.arange(n_samples), frames), dtype=torch.long))
    images_np = np.vstack([s[2] for s in batch])
    aug = _augmentation_info(batch)
    if not encoded:
        events = {}
        for k in ('x', 'y', 'polarity', 'element_index'):
            out = _pinned((total,), torch.long)
            np.concatenate([s[0][k] for s in batch], out=out.numpy(), casting='unsafe')
            events[k] = out
        t = _pinned((total,), torch.float32)
        np.concatenate([s[0]['timestamp'] for s in batch], out=t.numpy(), casting='unsafe')
        events['timestamp'] = t
        si = _pinned((total,), torch.long)
        si.numpy()[:] = np.repeat(np.arange(n_samples), counts)                    # :974-985
        events['sample_index'] = si
        images = _pinned((images_np.shape[0], 1) + images_np.shape[1:], torch.float32)
        images.numpy()[:, 0] = images_np
        return {'events': events, 'timestamps': timestamps, 'sample_idx': sample_idx, 'images': images,
                'augmentation_params': aug, 'size': n_samples}
    # ---- storage form (encode_batch, :240-305), straight from the per-sample arrays ----
    x, y = _pinned((total,), torch.int16), _pinned((total,), torch.int16)
    p, t = _pinned((total,), torch.bool), _pinned((total,), torch.float32)
    np.concatenate([s[0]['x'] for s in batch], out=x.numpy(), casting='unsafe')    # .to(torch.short): wraps
    np.concatenate([s[0]['y'] for s in batch], out=y.numpy(), casting='unsafe')
    np.concatenate([s[0]['timestamp'] for s in batch], out=t.numpy(), casting='unsafe')
    off = 0
    pn = p.numpy()
    for s, n in zip(batch, counts):                                               # ((p + 1) / 2).to(bool), :289
        np.not_equal(s[0]['polarity'], -1, out=pn[off:off + n])
        off += n
    elements = [f - 1 for f in frames]                                            # :231-233
    per_element = [np.bincount(np.asarray(s[0]['element_index'], dtype=np.int64), minlength=e)[:max(e, 0)]
                   if s[0]['x'].size else np.zeros(max(e, 0), np.int64) for s, e in zip(batch, elements)]
    epe = np.concatenate(per_element) if per_element else np.zeros(0, np.int64)
    # the reference sizes events_per_element by the LAST event's element (:298): elements after it
    # (empty trailing elements) are not stored
    if total:
        last_sample = max(i for i, n in enumerate(counts) if n)
        keep = sum(elements[:last_sample]) + int(np.asarray(batch[last_sample][0]['element_index'])[-1]) + 1
        epe = epe[:keep]
    images = _pinned((images_np.shape[0], 1) + images_np.shape[1:], torch.uint8)
    images.numpy()[:, 0] = images_np                                              # images.to(torch.uint8), :235
    return {'events': {'x': x, 'y': y, 'timestamp': t, 'polarity': p,
                       'events_per_element': pin(torch.tensor(epe, dtype=torch.int64))},
            'timestamps': timestamps, 'elements_per_sample': pin(torch.tensor(elements, dtype=torch.uint8)),
            'images': images, 'augmentation_params': aug}
