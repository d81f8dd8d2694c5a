"""Self-supervised multiscale warp loss -- drop-in for the reference's `utils/loss.py`
(`Loss` :38-171, `Losses` :174-214, `init_losses` :217-240, `charbonier_loss` :24-35,
`interpolate` :20-21), forward AND backward on the fused sm_100a kernels of csrc/loss.cu.

Same names, positional order, defaults and return structure as the reference; results are
autograd-connected 0-dim tensors.  There is no CPU path: inputs must be CUDA tensors."""
from typing import Optional

import torch

from . import _lib
from .timer import FakeTimer


def _f32(t, name):
    return _lib.require_cuda(t, name, torch.float32)


def interpolate(img, shape):
    """utils/loss.py:20-21: bilinear, align_corners=True.  img (D,1,H,W) or (D,H,W) fp32."""
    img = _f32(img, 'img')
    lead = img.shape[:-2]
    hin, win = img.shape[-2:]
    hout, wout = int(shape[0]), int(shape[1])
    out = torch.empty(lead + (hout, wout), dtype=torch.float32, device=img.device)
    d = img.numel() // (hin * win)
    if d:
        with torch.cuda.device(img.device):
            _lib.call('dvsof_resize_bilinear', _lib.ptr(img), d, hin, win, _lib.ptr(out), hout,
                      wout, _lib.stream_ptr(img.device))
    return out


def charbonier_loss(delta, alpha: float = 0.45, epsilon: float = 1e-3,
                    denominator: Optional[torch.Tensor] = None,
                    device: Optional[torch.device] = torch.device('cpu')):
    """utils/loss.py:24-35.  Kept for API compatibility (plain tensor ops, any device); the fused
    kernels evaluate the same penalty inline."""
    if delta.numel() == 0:
        return torch.tensor(0, dtype=torch.float32, device=device)
    delta = (delta.pow(2) + epsilon * epsilon).pow(alpha)
    if denominator is None:
        return delta.mean()
    assert denominator.numel() == delta.numel()
    return (delta / denominator).sum()


def _scale_struct(flows, levels, grads=None):
    sc = _lib.LossScales()
    sc.n_scales = len(flows)
    sc.n_pred = flows[0].shape[0]
    sc.n_images = levels[0].shape[0]
    for s, (f, im) in enumerate(zip(flows, levels)):
        sc.H[s], sc.W[s] = f.shape[-2], f.shape[-1]
        sc.flow[s] = f.data_ptr()
        sc.images[s] = im.data_ptr()
        sc.grad_flow[s] = grads[s].data_ptr() if grads is not None else None
    return sc


def _carve(buf, sizes):
    """Split a flat fp32 buffer into 16-byte aligned pieces of `sizes` elements."""
    out, off = [], 0
    for n in sizes:
        out.append(buf[off:off + n])
        off += (n + 3) & ~3
    return out


def _flat_len(sizes):
    return sum((n + 3) & ~3 for n in sizes)


def _backward(match_idx, n_bad, levels, flows, grad_terms):
    dev = flows[0].device
    sizes = [f.numel() for f in flows]
    gbuf = torch.empty(_flat_len(sizes), dtype=torch.float32, device=dev)
    grads = [g.view(f.shape) for g, f in zip(_carve(gbuf, sizes), flows)]
    with torch.cuda.device(dev):
        sc = _scale_struct(flows, levels, grads)
        _lib.call('dvsof_warp_loss_backward', sc, _lib.ptr(match_idx), _lib.ptr(grad_terms.contiguous()),
                  _lib.ptr(n_bad), _lib.stream_ptr(dev))
    return grads


class _WarpLoss(torch.autograd.Function):
    """terms[3*S] = [smoothness(s)..., photometric(s)..., outborder(s)...] of all scales in one
    fused launch, from ready-made pyramid levels; backward is the analytic gradient w.r.t. every
    flow (images get none)."""

    @staticmethod
    def forward(ctx, match_idx, n_scales, *tensors):
        levels, flows = tensors[:n_scales], tensors[n_scales:]
        dev = flows[0].device
        P = flows[0].shape[0]
        with _lib.on_device(dev):
            sc = _scale_struct(flows, levels)
            lib = _lib.load()
            ws = torch.empty(lib.dvsof_loss_workspace_bytes(sc), dtype=torch.uint8, device=dev)
            terms = torch.empty(3 * n_scales, dtype=torch.float32, device=dev)
            n_bad = torch.empty(n_scales * P, dtype=torch.int32, device=dev)
            _lib.call('dvsof_warp_loss_forward', sc, _lib.ptr(match_idx), _lib.ptr(terms),
                      _lib.ptr(n_bad), _lib.ptr(ws), _lib.stream_ptr(dev))
        ctx.n_scales = n_scales
        ctx.save_for_backward(match_idx, n_bad, *levels, *flows)
        return terms

    @staticmethod
    def backward(ctx, grad_terms):
        S = ctx.n_scales
        match_idx, n_bad = ctx.saved_tensors[:2]
        levels, flows = ctx.saved_tensors[2:2 + S], ctx.saved_tensors[2 + S:]
        return (None, None) + (None,) * S + tuple(_backward(match_idx, n_bad, levels, flows, grad_terms))


class _Plan:
    """Per (device, P, D, image size, scale shapes) constants of the fused call: the layout of the
    one fp32 buffer a forward allocates (pyramid levels | match_idx | n_bad | terms), the gradient
    buffer layout, and reusable ctypes argument blocks whose pointer fields are the only thing
    that changes from call to call (no tensor views, no struct rebuilds on the hot path)."""

    def __init__(self, dev, P, D, H, W, shapes):
        self.S = S = len(shapes)
        self.shapes = shapes
        sizes = [D * h * w for h, w in shapes] + [2 * P, S * P, 3 * S, 1]   # .. | fused: the loss
        self.offs, off = [], 0
        for n in sizes:
            self.offs.append(off)
            off += (n + 3) & ~3                              # 16-byte aligned pieces
        self.total = off
        self.gsizes = [P * 2 * h * w for h, w in shapes]
        self.goffs, off = [], 0
        for n in self.gsizes:
            self.goffs.append(off)
            off += (n + 3) & ~3
        self.gtotal = off
        self.fwd, self.bwd, self.frames = _lib.LossScales(), _lib.LossScales(), _lib.LossFrames()
        for sc in (self.fwd, self.bwd):
            sc.n_scales, sc.n_pred, sc.n_images = S, P, D
            for s, (h, w) in enumerate(shapes):
                sc.H[s], sc.W[s] = h, w
        self.frames.n_images, self.frames.H0, self.frames.W0 = D, H, W
        self.ws = torch.empty(_lib.load().dvsof_loss_workspace_bytes(self._sized(self.fwd)),
                              dtype=torch.uint8, device=dev)  # tile records + ticket, forward only
        self.ws_ptr = _lib.ptr(self.ws)

    @staticmethod
    def _sized(sc):
        for s in range(sc.n_scales):                         # workspace_bytes validates non-null
            sc.flow[s] = sc.images[s] = 1
        return sc

    def fill(self, sc, base, flows, grad_base=None):
        for s in range(self.S):
            sc.flow[s] = flows[s].data_ptr()
            sc.images[s] = base + 4 * self.offs[s]
            if grad_base is not None:
                sc.grad_flow[s] = grad_base + 4 * self.goffs[s]


def _fill_frames(plan, images, flow_ts, flow_sample_idx, timestamps, sample_idx):
    fr = plan.frames
    fr.images = images.data_ptr()
    fr.flow_ts, fr.flow_sample_idx = flow_ts.data_ptr(), flow_sample_idx.data_ptr()
    fr.timestamps, fr.sample_idx = timestamps.data_ptr(), sample_idx.data_ptr()
    return fr


class _FusedTrainingLoss(torch.autograd.Function):
    """utils/training.py:12-24 on top of utils/loss.py:179-214 as ONE sweep: the 3*S terms, the
    weighted loss and d(loss)/d(flow) of every scale come out of the same kernel
    (`dvsof_multiscale_loss_fused`); backward only rescales the stored gradient by the incoming
    d/d(loss) -- on device, and not at all when that is exactly 1."""

    @staticmethod
    def forward(ctx, owner, coef, images, flow_ts, flow_sample_idx, timestamps, sample_idx, *flows):
        dev = images.device
        D, _, H, W = images.shape
        plan = owner._plan(dev, flows[0].shape[0], D, H, W, flows)
        S, o = plan.S, plan.offs
        with _lib.on_device(dev):
            buf = torch.empty(plan.total, dtype=torch.float32, device=dev)
            # one gradient tensor per scale, straight from the caching allocator (an empty() costs a
            # third of a view of a shared buffer)
            grads = [torch.empty_like(f) for f in flows]
            base = buf.data_ptr()
            sc = plan.bwd
            for s in range(S):
                sc.flow[s] = flows[s].data_ptr()
                sc.images[s] = base + 4 * o[s]
                sc.grad_flow[s] = grads[s].data_ptr()
            fr = _fill_frames(plan, images, flow_ts, flow_sample_idx, timestamps, sample_idx)
            st = _lib.status_for(dev)
            st.flush()
            _lib.call('dvsof_multiscale_loss_fused', sc, fr, coef, base + 4 * o[S],
                      base + 4 * o[S + 2], base + 4 * o[S + 3], base + 4 * o[S + 1], plan.ws_ptr,
                      _lib.ptr(st.word), _lib.stream_ptr(dev))
            st.after_call()                                              # loss.py:193 asserts
        ctx.plan, ctx.grads, ctx.buf = plan, grads, buf
        loss = buf[o[S + 3]:o[S + 3] + 1].view(())
        terms = buf[o[S + 2]:o[S + 2] + 3 * S]
        ctx.mark_non_differentiable(terms)
        return loss, terms

    @staticmethod
    def backward(ctx, grad_loss, _grad_terms):
        plan, grads = ctx.plan, ctx.grads
        if grads is None:
            raise RuntimeError('the fused training loss stores its gradient once: a second backward '
                               '(retain_graph) needs Losses.__call__ / Losses.terms')
        dev = grads[0].device
        grad_loss = _f32(grad_loss, 'grad_loss')
        with _lib.on_device(dev):
            sc = plan.bwd       # only the gradient pointers are read by the rescale
            for s in range(plan.S):
                sc.grad_flow[s] = grads[s].data_ptr()
            _lib.call('dvsof_scale_gradients', sc, _lib.ptr(grad_loss), _lib.stream_ptr(dev))
        ctx.grads = ctx.buf = None
        return (None,) * 7 + tuple(grads)


class _MultiscaleLoss(torch.autograd.Function):
    """utils/loss.py:179-214 forward in ONE library call (matching + chained pyramid + fused
    terms, kernels chained on the stream with programmatic dependent launch)."""

    @staticmethod
    def forward(ctx, owner, images, flow_ts, flow_sample_idx, timestamps, sample_idx, *flows):
        dev = images.device
        D, _, H, W = images.shape
        plan = owner._plan(dev, flows[0].shape[0], D, H, W, flows)
        S = plan.S
        with _lib.on_device(dev):
            # one allocation per call: what backward needs (levels, match_idx, n_bad) cannot live
            # in a buffer shared between calls
            buf = torch.empty(plan.total, dtype=torch.float32, device=dev)
            base = buf.data_ptr()
            plan.fill(plan.fwd, base, flows)
            fr = _fill_frames(plan, images, flow_ts, flow_sample_idx, timestamps, sample_idx)
            st = _lib.status_for(dev)
            st.flush()
            o = plan.offs
            _lib.call('dvsof_multiscale_loss_forward', plan.fwd, fr, base + 4 * o[S], base + 4 * o[S + 2],
                      base + 4 * o[S + 1], plan.ws_ptr, _lib.ptr(st.word), _lib.stream_ptr(dev))
            st.after_call()                                              # loss.py:193 asserts
        ctx.plan = plan
        ctx.save_for_backward(buf, *flows)
        return buf[o[S + 2]:o[S + 2] + 3 * S]

    @staticmethod
    def backward(ctx, grad_terms):
        plan = ctx.plan
        buf, flows = ctx.saved_tensors[0], ctx.saved_tensors[1:]
        dev, S, o = buf.device, plan.S, plan.offs
        grad_terms = grad_terms.contiguous()
        with _lib.on_device(dev):
            gbuf = torch.empty(plan.gtotal, dtype=torch.float32, device=dev)
            base = buf.data_ptr()
            plan.fill(plan.bwd, base, flows, gbuf.data_ptr())
            _lib.call('dvsof_warp_loss_backward', plan.bwd, base + 4 * o[S], _lib.ptr(grad_terms),
                      base + 4 * o[S + 1], _lib.stream_ptr(dev))
        grads = tuple(gbuf[g:g + n].view(f.shape) for g, n, f in zip(plan.goffs, plan.gsizes, flows))
        return (None,) * 6 + grads


def _check_shapes(owner, N, H, W):
    # utils/loss.py:122-131 (same messages' meaning, AssertionError)
    assert owner.N >= N, f'This object should be used for batch of at most {owner.N} samples, ' \
                         f'but {N} samples are given'
    assert owner.H == H, f'This object should be used for images of height {owner.H}, ' \
                         f'but image of height {H} are given'
    assert owner.W == W, f'This object should be used for images of width {owner.W}, ' \
                         f'but image of width {W} are given'


class Loss:
    """One scale (utils/loss.py:38-171)."""

    def __init__(self, pred_shape, batch_size, device, timers=FakeTimer()):
        self.N = batch_size
        self.H, self.W = pred_shape
        self.device = device
        self.timers = timers

    def __call__(self, prev_images, next_images, flow):
        N, C, H, W = prev_images.size()
        _check_shapes(self, N, H, W)
        assert tuple(next_images.size()) == (N, C, H, W)                   # loss.py:133-137
        FN, FC, FH, FW = flow.size()
        assert FN == N, f'Number of images and flows should be the same {N} vs {FN}'
        assert FC == 2, 'Flow should contain 2 channels (dx and dy)'
        assert FH == H, f'Height of images and flows should be the same {H} vs {FH}'
        assert FW == W, f'Width of images and flows should be the same {W} vs {FW}'
        assert C == 1, 'the fused kernel warps single-channel (grayscale) frames'
        for name in ('grid_construction', 'smoothness_loss', 'outborder_loss'):
            self.timers(name).start()
            self.timers(name).stop()
        self.timers('photometric_loss').start()
        flow = _f32(flow, 'flow')
        level = torch.cat((_f32(prev_images, 'prev_images').view(N, H, W),
                           _f32(next_images, 'next_images').view(N, H, W)))
        match_idx = torch.arange(2 * N, dtype=torch.int32, device=flow.device)
        terms = _WarpLoss.apply(match_idx, 1, level, flow)
        self.timers('photometric_loss').stop()
        smoothness, photometric, outborder = terms.unbind(0)
        return smoothness, photometric, outborder


class Losses:
    """All scales (utils/loss.py:174-214): image<->flow timestamp matching, the CHAINED bilinear
    pyramid (line 210) and the per-scale terms, with no host synchronisation (Status.mode
    'deferred') or exactly one (mode 'sync', to raise the reference's AssertionError in place)."""

    def __init__(self, shapes, batch_size, device, timers=FakeTimer()):
        self.losses = [Loss(shape, batch_size, device, timers) for shape in shapes]
        self.timers = timers
        self._plans = {}
        self._coefs = {}

    def _plan(self, dev, P, D, H, W, flows):
        shapes = tuple((f.shape[-2], f.shape[-1]) for f in flows)
        # the plan owns the reduction workspace (tile records, tickets, per-tile counts): one per
        # stream, or two calls with equal shapes on different streams would share the tickets
        key = (dev, _lib.raw_stream(dev), P, D, H, W, shapes)
        plan = self._plans.get(key)
        if plan is None:
            plan = self._plans[key] = _Plan(dev, P, D, H, W, shapes)
        return plan

    def _checked(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx):
        S = len(flows)
        assert S == len(self.losses) and 1 <= S <= _lib.MAX_SCALES
        images = _f32(images, 'images')
        assert images.dim() == 4 and images.shape[1] == 1, \
            'the fused kernel warps single-channel (grayscale) frames'
        D = images.shape[0]
        P = flow_sample_idx.shape[0]
        flows = [_f32(f, 'flow') for f in flows]
        for loss, f in zip(self.losses, flows):
            assert f.dim() == 4 and f.shape[0] == P and f.shape[1] == 2
            _check_shapes(loss, P, f.shape[-2], f.shape[-1])
        flow_ts = _f32(flow_ts, 'flow_ts')
        timestamps = _f32(timestamps, 'timestamps')
        assert flow_ts.numel() == 2 * P and timestamps.numel() == D and sample_idx.numel() == D
        return (images, flow_ts, _lib.require_cuda(flow_sample_idx, 'flow_sample_idx', torch.int64),
                timestamps, _lib.require_cuda(sample_idx, 'sample_idx', torch.int64), *flows)

    def terms(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx):
        """The 3*S terms as ONE autograd-connected tensor
        [smoothness(s)..., photometric(s)..., outborder(s)...]; `__call__` returns views of it."""
        return _MultiscaleLoss.apply(self, *self._checked(flows, flow_ts, flow_sample_idx, images,
                                                          timestamps, sample_idx))

    def combined(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx,
                 weights=(0.5, 1, 1)):
        """(loss, terms) with loss = sum_k weights[k] * mean_s(term[k][s]) (utils/training.py:12-24).
        When a flow requires grad, the gradient is produced by the same kernel sweep as the terms
        (`_FusedTrainingLoss`); `terms` is then a detached tensor (logging only)."""
        args = self._checked(flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx)
        S = len(flows)
        key = (images.device, S) + tuple(float(w) for w in weights)
        coef = self._coefs.get(key)
        if coef is None:
            vals = [w / S for w in key[2:] for _ in range(S)]
            coef = self._coefs[key] = ((_lib.C.c_float * (3 * S))(*vals),
                                       torch.tensor(vals, dtype=torch.float32, device=images.device))
        if torch.is_grad_enabled() and any(f.requires_grad for f in args[5:]):
            return _FusedTrainingLoss.apply(self, coef[0], *args)
        flat = _MultiscaleLoss.apply(self, *args)
        return torch.dot(flat, coef[1]), flat

    def __call__(self, flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx):
        S = len(flows)
        flat = self.terms(flows, flow_ts, flow_sample_idx, images, timestamps, sample_idx).unbind(0)
        return tuple(flat[:S]), tuple(flat[S:2 * S]), tuple(flat[2 * S:])


def init_losses(shape, batch_size, model, device, sequence_length, timers=FakeTimer()):
    """utils/loss.py:217-240: one no-grad model call on EMPTY events to learn the output shapes."""
    def empty(dtype):
        return torch.tensor([], dtype=dtype, device=device)
    events = {'x': empty(torch.long), 'y': empty(torch.long), 'timestamp': empty(torch.float32),
              'polarity': empty(torch.long), 'element_index': empty(torch.long),
              'sample_index': empty(torch.long)}
    with torch.no_grad():
        num_timestamps = sequence_length + 1
        out = model(events,
                    torch.tensor([0.04 * i for i in range(num_timestamps)], dtype=torch.float32,
                                 device=device),
                    torch.tensor([0] * num_timestamps, dtype=torch.long, device=device),
                    shape, raw=True)
    out_shapes = tuple(tuple(flow.shape[2:]) for flow in out[0])
    return Losses(out_shapes, batch_size, device, timers=timers)
