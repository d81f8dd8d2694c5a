"""Training step -- drop-in for the reference's `utils/training.py` (`combined_loss` :12-24,
`process_minibatch` :37-86, `train` :89-235, `validate` :244-271) with the data-parallel
partition the reference lacks:

  * the batch is sharded BY SAMPLE across ranks (`shard_batch`): events are grouped by
    ascending `sample_index` (utils/dataset.py:974-985), so a shard is one contiguous event
    range plus the matching `2b` ranges of images / timestamps / sample_idx -- no data-path
    collective;
  * the only exchange is the gradient allreduce of the optimizer step (`GradientAllreducer`):
    flat fp32 buckets, NCCL over NVLink (gloo in the CPU tests), launched as each bucket's
    gradients become ready so the wire time hides behind cuDNN's backward.

One process per GPU (torchrun); world size 1 degenerates to the reference's loop.
"""
import torch
import torch.distributed as dist

from .timer import FakeTimer


def mean(values):
    """utils/common.py:22-23."""
    return sum(values) / len(values)


def _world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def combined_loss(evaluator, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx,
                  features, weights=[0.5, 1, 1]):
    """utils/training.py:12-24: loss = sum_k w_k * mean_s(term[k][s]), k over (smoothness,
    photometric, outborder).  The fused evaluator (`Losses.combined`) returns the weighted loss,
    the 3*S terms and -- when a flow requires grad -- d(loss)/d(flow) from ONE kernel sweep (one
    node in the autograd graph instead of ~30)."""
    if hasattr(evaluator, 'combined'):
        loss, flat = evaluator.combined(flows, flow_ts, flow_sample_idx, images, timestamps,
                                        sample_idx, weights)
        S = flat.numel() // 3
        if flat.requires_grad:                      # generic path: autograd-connected scalars
            parts = flat.unbind(0)
            return loss, (tuple(parts[:S]), tuple(parts[S:2 * S]), tuple(parts[2 * S:]))
        host = _HostTerms(flat)                     # logging values: fetched once, when first looked at
        return loss, tuple(_TermGroup(host, k * S, (k + 1) * S) for k in range(3))
    terms = evaluator(flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx)
    loss = sum(map(lambda v, w: w * mean(v), terms, weights))
    return loss, terms


class _HostTerms:
    """The 3*S detached terms of the fused pass, copied to the host in ONE transfer the first time
    any of them is looked at (the reference's logging does 12 `.item()` calls = 12 syncs)."""

    def __init__(self, flat):
        self.flat, self._host = flat, None

    def __getitem__(self, i):
        if self._host is None:
            self._host = self.flat.detach().cpu()
        return self._host[i]


class _TermGroup:
    """One of (smoothness, photometric, outborder): a sequence of S 0-dim tensors, as the
    reference's tuples, materialised on demand."""

    def __init__(self, host, lo, hi):
        self._h, self._lo, self._hi = host, lo, hi

    def __len__(self):
        return self._hi - self._lo

    def __getitem__(self, i):
        if isinstance(i, slice):
            return tuple(self[j] for j in range(*i.indices(len(self))))
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        return self._h[self._lo + i]

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def predictions2tag(predictions):
    """utils/training.py:33-34."""
    return (f'{x.shape[-2]}x{x.shape[-1]}' for x in predictions)


def shard_bounds(n_samples, rank, world):
    """Samples [lo, hi) of rank `rank`: equal split, remainder to the first ranks."""
    base, rem = divmod(n_samples, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(batch, rank, world):
    """Partition a collated batch (utils/dataset.py:961-1020 layout) by sample.  Returns a new
    dict with this rank's contiguous sample range, `sample_index` / `sample_idx` rebased to
    start at 0.  Frames per sample = len(timestamps) / size (2 for sequence length 1)."""
    if world == 1:
        return batch
    B = int(batch['size'])
    lo, hi = shard_bounds(B, rank, world)
    out = dict(batch)
    out['size'] = hi - lo
    per = batch['timestamps'].numel() // B
    for k in ('timestamps', 'sample_idx', 'images'):
        out[k] = batch[k][lo * per:hi * per]
    out['sample_idx'] = out['sample_idx'] - lo
    if 'events' in batch:
        ev = batch['events']
        b = ev['sample_index']
        # grouped + ascending -> the shard is one contiguous range (two binary searches)
        edges = torch.searchsorted(b, torch.tensor([lo, hi], dtype=b.dtype, device=b.device))
        e0, e1 = int(edges[0]), int(edges[1])
        sh = {k: v[e0:e1] for k, v in ev.items() if isinstance(v, torch.Tensor) and v.dim() == 1
              and v.numel() == b.numel()}
        sh['sample_index'] = sh['sample_index'] - lo
        for k, v in ev.items():
            sh.setdefault(k, v)
        out['events'] = sh
    if 'data' in batch:   # pre-quantised event images (B,4,H,W)
        out['data'] = batch['data'][lo:hi]
    if isinstance(batch.get('augmentation_params'), dict):
        out['augmentation_params'] = {k: v[lo:hi] for k, v in batch['augmentation_params'].items()}
    return out


def shard_encoded_batch(batch, rank, world):
    """The same partition for a batch in the storage form of utils/dataset.py:240-305 (what
    `PreprocessedDataloader` reads from disk): this rank's samples, found from the run lengths
    (cf. select_encoded_ranges, utils/dataset.py:67-115).  Host tensors; a few tiny prefix sums."""
    if world == 1:
        return batch
    eps = batch['elements_per_sample'].to(torch.int64)
    B = eps.numel()
    lo, hi = shard_bounds(B, rank, world)
    elem = torch.zeros(B + 1, dtype=torch.int64)
    torch.cumsum(eps, 0, out=elem[1:])
    ev = batch['events']
    epe = ev['events_per_element'].to(torch.int64)
    E = epe.numel()                                   # trailing empty elements may be missing (:298)
    el0, el1 = min(int(elem[lo]), E), min(int(elem[hi]), E)
    evb = torch.zeros(E + 1, dtype=torch.int64)
    torch.cumsum(epe, 0, out=evb[1:])
    e0, e1 = int(evb[el0]), int(evb[el1])
    img0, img1 = int(elem[lo]) + lo, int(elem[hi]) + hi      # a sample has elements + 1 images
    out = dict(batch)
    out['elements_per_sample'] = batch['elements_per_sample'][lo:hi]
    out['timestamps'] = batch['timestamps'][img0:img1]
    out['images'] = batch['images'][img0:img1]
    sh = {k: ev[k][e0:e1] for k in ('x', 'y', 'timestamp', 'polarity')}
    sh['events_per_element'] = ev['events_per_element'][el0:el1]
    out['events'] = sh
    if isinstance(batch.get('augmentation_params'), dict):
        out['augmentation_params'] = {k: v[lo:hi] for k, v in batch['augmentation_params'].items()}
    return out


class GradientAllreducer:
    """Flat-bucket gradient averaging for the data-parallel step.

    Every parameter's `.grad` is a view into one of a few flat fp32 buckets (filled in reverse
    registration order = the order backward produces them), so a bucket is reduced in place with
    one `all_reduce(SUM)` pre-scaled by 1/world, issued from a post-accumulate hook the moment
    its last gradient lands -- asynchronously, on the process group's own stream, while cuDNN
    continues with the earlier layers.  `wait()` before `optimizer.step()`."""

    def __init__(self, params, bucket_mb=16.0, group=None):
        self.group = group
        self.rank, self.world = _world()
        self.params = [p for p in params if p.requires_grad]
        self.buckets, self.where, self.handles = [], {}, []
        cap = int(bucket_mb * (1 << 20)) // 4
        cur, cur_n = [], 0
        for p in reversed(self.params):
            if cur and cur_n + p.numel() > cap:
                self._close(cur, cur_n)
                cur, cur_n = [], 0
            cur.append(p)
            cur_n += p.numel()
        if cur:
            self._close(cur, cur_n)
        self.hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.enabled = True          # False on non-boundary micro-batches (accumulation)

    def _close(self, plist, n):
        flat = torch.zeros(n, dtype=plist[0].dtype, device=plist[0].device)
        idx, off = len(self.buckets), 0
        for p in plist:
            piece = flat[off:off + p.numel()]
            # the view has the parameter's own strides (channels_last weights stay channels_last):
            # autograd's gradient layout contract holds, nothing is re-laid out when the gradient
            # is accumulated, and fused optimizers accept the pair
            dense = p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))
            p.grad = piece.as_strided(p.size(), p.stride()) if dense else piece.view_as(p)
            self.where[p] = idx
            off += p.numel()
        self.buckets.append({'flat': flat, 'params': plist, 'pending': len(plist)})

    def _on_grad(self, p):
        if not self.enabled or self.world == 1:
            return
        b = self.buckets[self.where[p]]
        b['pending'] -= 1
        if b['pending'] == 0:
            b['flat'].div_(self.world)
            self.handles.append(dist.all_reduce(b['flat'], op=dist.ReduceOp.SUM, group=self.group,
                                                async_op=True))

    def wait(self):
        """Block the current stream on every outstanding bucket; reduce buckets whose hooks did
        not all fire (parameters unused in this step keep their zero gradient)."""
        if self.world > 1 and self.enabled:
            for b in self.buckets:
                if b['pending'] != 0:    # same on every rank: same graph, same unused parameters
                    b['flat'].div_(self.world)
                    self.handles.append(dist.all_reduce(b['flat'], op=dist.ReduceOp.SUM,
                                                        group=self.group, async_op=True))
            for h in self.handles:
                h.wait()
        self.handles = []
        for b in self.buckets:
            b['pending'] = len(b['params'])

    def zero_grad(self):
        """Keep the flat views (set_to_none would detach them): zero in place."""
        for b in self.buckets:
            b['flat'].zero_()

    def broadcast_parameters(self, module):
        """Rank 0's weights and buffers everywhere (BatchNorm statistics stay per replica
        afterwards)."""
        if self.world == 1:
            return
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src=0, group=self.group)


def _upload(batch, device, is_raw):
    """H2D of what process_minibatch needs (utils/training.py:46-56), non-blocking."""
    out = dict(batch)
    for k in ('timestamps', 'sample_idx', 'images'):
        out[k] = batch[k].to(device, non_blocking=True)
    if is_raw:
        ev = dict(batch['events'])
        for k in set.difference(set(ev.keys()), {'size'}):
            ev[k] = ev[k].to(device, non_blocking=True)
        out['events'] = ev
    else:
        out['data'] = batch['data'].to(device, non_blocking=True)
    return out


def _tensors(batch):
    for v in batch.values():
        if isinstance(v, torch.Tensor):
            yield v
        elif isinstance(v, dict):
            yield from _tensors(v)


class DevicePrefetcher:
    """Wraps a loader of collated host batches: keeps this rank's shard (by sample) and uploads
    batch k+1 on a copy stream while the caller computes on batch k, so the per-step H2D
    (36 B/event in the int64 training form) hides behind cuDNN instead of preceding it -- with 8
    ranks sharing one host the serial copy was 10 % of the step.  Yields device-resident batches
    that `process_minibatch` takes as is (`.to(device)` of a device tensor is a no-op).  Pinned
    host tensors make the copies asynchronous; pageable ones still work (synchronous copy)."""

    def __init__(self, loader, device, is_raw=True, rank=0, world=1):
        self.loader, self.device, self.is_raw = loader, torch.device(device), is_raw
        self.rank, self.world = rank, world
        self.stream = torch.cuda.Stream(self.device) if self.device.type == 'cuda' else None

    def _stage_encoded(self, batch):
        """A batch in the storage form (utils/dataset.py:240-305; 9 B/event, uint8 images): upload
        that and expand it on the device (`dataset.decode_batch`) -- a quarter of the bytes of the
        int64 training form cross PCIe.  Device only (libdvsof has no CPU path)."""
        from . import dataset
        local = shard_encoded_batch(batch, self.rank, self.world)

        def up(v):
            return v.to(self.device, non_blocking=True) if isinstance(v, torch.Tensor) else v
        with torch.cuda.stream(self.stream):
            dev = {k: ({kk: up(vv) for kk, vv in v.items()} if isinstance(v, dict) else up(v))
                   for k, v in local.items()}
            out = dataset.decode_batch(dev)
        out['global_size'] = batch['elements_per_sample'].numel()
        out['global_events'] = batch['events']['x'].numel()
        return out

    def _stage(self, batch):
        if batch is None:
            return None
        if 'elements_per_sample' in batch:
            return self._stage_encoded(batch)
        local = dict(shard_batch(batch, self.rank, self.world))
        local['global_size'] = batch['size']                     # what the reference's counters see
        local['global_events'] = batch['events']['x'].numel() if self.is_raw else 0
        if self.stream is None:
            return _upload(local, self.device, self.is_raw)
        with torch.cuda.stream(self.stream):
            return _upload(local, self.device, self.is_raw)

    def __iter__(self):
        it = iter(self.loader)
        nxt = self._stage(next(it, None))
        while nxt is not None:
            cur = nxt
            if self.stream is not None:
                compute = torch.cuda.current_stream(self.device)
                compute.wait_stream(self.stream)                 # batch k has landed
                for t in _tensors(cur):
                    if t.is_cuda:
                        t.record_stream(compute)                 # allocated on the copy stream
            nxt = self._stage(next(it, None))                    # upload k+1 starts now
            yield cur

    def __len__(self):
        return len(self.loader)


def process_minibatch(model, batch, timers, device, is_raw, evaluator, weights,
                      return_prediction=False):
    """utils/training.py:37-86, unchanged in meaning: H2D of the batch, model forward (whose
    first act is `quantize` -> the CUDA encoder), fused multiscale loss."""
    timers('batch2gpu').start()
    timestamps, sample_idx, images = map(lambda x: x.to(device, non_blocking=True),
                                         (batch['timestamps'], batch['sample_idx'], batch['images']))
    if is_raw:
        events = dict(batch['events'])
        for k in set.difference(set(events.keys()), {'size'}):
            events[k] = events[k].to(device, non_blocking=True)
    else:
        events = batch['data'].to(device, non_blocking=True)
    timers('batch2gpu').stop()
    shape = images.size()[-2:]
    timers('forward').start()
    prediction, flow_ts, flow_sample_idx, features = model(events, timestamps, sample_idx, shape,
                                                           raw=is_raw, intermediate=True)
    tags = predictions2tag(prediction)
    timers('forward').stop()
    timers('loss').start()
    loss, terms = combined_loss(evaluator, prediction, flow_ts, flow_sample_idx, images,
                                timestamps, sample_idx, features, weights=weights)
    terms = ((y.item() for y in x) for x in terms)      # lazy, as the reference: no sync here
    timers('loss').stop()
    add_info = tuple()
    if return_prediction:
        add_info = ({'prediction': prediction, 'flow_ts': flow_ts,
                     'flow_sample_idx': flow_sample_idx, 'features': features}, )
    return (loss, terms, tags) + add_info


def add_loss(loss_sum, loss_values):
    """utils/training.py:238-241."""
    if len(loss_sum) == 0:
        return list(loss_values)
    return [x + y for x, y in zip(loss_sum, loss_values)]


def train(model, device, loader, optimizer, num_steps: int, scheduler, logger, evaluator,
          weights=[0.5, 1, 1], is_raw=True, accumulation_steps=1, timers=FakeTimer(), hooks={},
          init_step=0, init_samples_passed=0, max_events_per_batch: int = 350000,
          reducer: GradientAllreducer = None, prefetch: bool = False):
    """utils/training.py:89-235 with the data-parallel partition: under an initialised process
    group every rank receives the same global batch from `loader`, keeps its shard (by sample)
    and averages gradients through `reducer` before the optimizer step.  Logging, hooks,
    accumulation and the over-large-batch skip follow the reference; only rank 0 logs."""
    rank, world = _world()
    if world > 1 and reducer is None:
        reducer = GradientAllreducer(model.parameters())
        reducer.broadcast_parameters(model)
    model.train()
    samples_passed = init_samples_passed
    loss_sum, smooth_sum, photo_sum, out_reg_sum = 0, [], [], []
    if reducer is not None:
        reducer.zero_grad()
    else:
        optimizer.zero_grad(set_to_none=True)
    init_batch = init_step * accumulation_steps
    global_step = init_batch
    num_skipped = 0
    if prefetch:       # shard + upload one batch ahead on a copy stream (DevicePrefetcher)
        loader = DevicePrefetcher(loader, device, is_raw, rank, world)
    timers('batch_construction').start()
    for batch in loader:
        if global_step == num_steps * accumulation_steps:
            break
        num_events = batch['global_events'] if 'global_events' in batch else \
            (batch['events']['x'].numel() if is_raw else 0)
        if num_events > max_events_per_batch:                       # training.py:141-150
            num_skipped += 1
            num_processed = global_step - init_batch
            if rank == 0:
                print(f'Skipping batch with {num_events} events')
                print(f'Augmentation parameters {batch.get("augmentation_params")}')
                print(f'Processing rate is {num_processed / (num_processed + num_skipped):.2f}')
            continue
        global_step += 1
        timers('batch_construction').stop()
        samples_passed += batch.get('global_size', batch['size'])
        is_step_boundary = global_step % accumulation_steps == 0
        if reducer is not None:
            reducer.enabled = is_step_boundary     # reduce once per optimizer step
        local = batch if prefetch else shard_batch(batch, rank, world)
        loss, (smoothness, photometric, out_reg), tags = process_minibatch(
            model, local, timers, device, is_raw, evaluator, weights)
        # not `loss /= ...` as training.py:155 has it: the fused evaluator's loss is a view of the
        # buffer its autograd node owns, which torch refuses to modify in place
        loss = loss / accumulation_steps
        timers('backprop').start()
        loss.backward()
        timers('backprop').stop()
        photo_sum = add_loss(photo_sum, photometric)
        smooth_sum = add_loss(smooth_sum, smoothness)
        out_reg_sum = add_loss(out_reg_sum, out_reg)
        if is_step_boundary:
            timers('optimizer_step').start()
            if reducer is not None:
                reducer.wait()
            optimizer.step()
            if reducer is not None:
                reducer.zero_grad()
            else:
                optimizer.zero_grad(set_to_none=True)
            timers('optimizer_step').stop()
            scheduler.step()
            timers('logging').start()
            loss_sum += loss.item()
            if rank == 0 and logger is not None:
                for tag, s, p, o in zip(tags, smooth_sum, photo_sum, out_reg_sum):
                    logger.add_scalar(f'Train/photometric loss/{tag}', p / accumulation_steps, samples_passed)
                    logger.add_scalar(f'Train/smoothness loss/{tag}', s / accumulation_steps, samples_passed)
                    logger.add_scalar(f'Train/out regularization/{tag}', o / accumulation_steps, samples_passed)
                logger.add_scalar('General/Train loss', loss_sum, samples_passed)
                for i, lr in enumerate([p['lr'] for p in optimizer.param_groups]):
                    logger.add_scalar(f'General/learning rate/{i}', lr, samples_passed)
            loss_sum, smooth_sum, photo_sum, out_reg_sum = 0, [], [], []
            timers('logging').stop()
            step = global_step // accumulation_steps
            for k, hook in hooks.items():
                timers(k).start()
                hook(step, samples_passed)
                timers(k).stop()
            model.train()
        else:
            timers('optimizer_step').start()                        # training.py:206-211: every
            timers('optimizer_step').stop()                         # phase exists in every row
            for k in hooks:
                timers(k).start()
                timers(k).stop()
            timers('logging').start()
            loss_sum += loss.item()
            timers('logging').stop()
        # training.py:222-233 (the timer group prints on rank 0 only and resets everywhere)
        timers.log(names=['batch_construction', 'batch2gpu', 'forward', 'loss', 'grid_construction',
                          'photometric_loss', 'smoothness_loss', 'outborder_loss', 'backprop',
                          'optimizer_step', 'free', 'logging'] + list(hooks))
        timers('batch_construction').start()
    timers('batch_construction').stop()
    return global_step // accumulation_steps


def validate(model, device, loader, samples_passed, logger, evaluator, weights=[0.5, 1, 1],
             is_raw=True):
    """utils/training.py:244-271."""
    model.eval()
    n = len(loader)
    photo_sum, smooth_sum, out_reg_sum, loss_sum = [], [], [], 0
    tags = ()
    with torch.no_grad():
        for batch in loader:
            loss, (smoothness, photometric, out_reg), tags = process_minibatch(
                model, batch, FakeTimer(), device, is_raw, evaluator, weights)
            photo_sum = add_loss(photo_sum, photometric)
            smooth_sum = add_loss(smooth_sum, smoothness)
            out_reg_sum = add_loss(out_reg_sum, out_reg)
            loss_sum += loss.item()
    logger.add_scalar('General/Validation loss', loss_sum / n, samples_passed)
    for tag, s, p, o in zip(tags, smooth_sum, photo_sum, out_reg_sum):
        logger.add_scalar(f'Validation/smoothness loss/{tag}', s / n, samples_passed)
        logger.add_scalar(f'Validation/photometric loss/{tag}', p / n, samples_passed)
        logger.add_scalar(f'Validation/out regularization loss/{tag}', o / n, samples_passed)
    return loss_sum / n
