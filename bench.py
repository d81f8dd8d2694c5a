#!/usr/bin/env python
"""bench.py -- hot-path benchmark of dvs_of_training_framework_b200 on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    (N > 1: launched by torchrun, one rank per GPU)

Headline metric (BASELINE.json): encoded ("voxelized") Mevents/s on configs[1] -- MVSEC-shaped
256x256 events, batch 8, 131072 events/sample.  A step = one pass of the encoder
over one synthetic batch, inputs resident in HBM, L2 flushed between steps.  The same JSON line
also carries the warp-loss fwd+bwd samples/s (`warp_loss`), larger workloads (`workloads`), the
end-to-end number through the public Python API from pinned host buffers (`e2e`), the roofline
of the dominant kernel and the CPU baseline (oracle port, timed on this box's host cores).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

CFG = {
    'cfg1': dict(B=1, H=260, W=346, n=32768, T=0.032),
    'cfg2': dict(B=8, H=256, W=256, n=131072, T=0.032),
    'cfg3': dict(B=32, H=256, W=256, n=5_000_000, T=0.032),
    'cfg4': dict(B=16, H=480, W=640, n=1_048_576, T=0.1),
}
HEADLINE = 'cfg2'
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the scatter kernel, from the committed
# `ncu --set full` capture (profiles/r01_ncu_full_v1_summary.csv)
TRAFFIC = {'cfg2': 46.15e6}


def peaks():
    f = ROOT / 'MEASURED_PEAKS.json'
    if f.exists():
        return float(json.loads(f.read_text())['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


def fast_events(seed, B, H, W, n, T, device):
    """Synthetic events of SURVEY 8d generated on `device` (same distribution as
    tests/helpers.synth_events; torch device RNG so cfg3's 160 M events take milliseconds)."""
    g = torch.Generator(device=device).manual_seed(seed)
    N = B * n
    ev = {'x': torch.randint(0, W, (N,), generator=g, device=device),
          'y': torch.randint(0, H, (N,), generator=g, device=device),
          'polarity': (torch.rand(N, generator=g, device=device) < 0.7).long() * 2 - 1,
          'sample_index': torch.arange(B, device=device).repeat_interleave(n)}
    t = torch.rand(B, n, generator=g, device=device) * T
    ev['timestamp'] = torch.sort(t, dim=1)[0].reshape(-1).float()
    timestamps = torch.tensor([0, T] * B, dtype=torch.float32, device=device)
    sample_idx = torch.arange(B, device=device).repeat_interleave(2)
    return ev, timestamps, sample_idx


def encoder_bytes(B, H, W, N):
    return 36 * N + 8 * B + 16 * B * H * W


def loss_bytes(B, H, W):
    tot, prev = 0, H * W
    for k in (3, 2, 1, 0):
        p = (H // 2 ** k) * (W // 2 ** k)
        tot += 40 * B * p + 8 * B * (prev + p)
        prev = p
    return tot


class Clocks:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md clocks line)."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def __enter__(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix='.csv')
            os.close(fd)
            self.proc = subprocess.Popen(
                ['nvidia-smi', f'--id={self.index}', f'--query-gpu={self.Q}',
                 '--format=csv,noheader,nounits', '-lms', '100'],
                stdout=open(self.path, 'w'), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.proc.wait()

    def summary(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': []}
        try:
            rows = [r.split(', ') for r in open(self.path).read().strip().splitlines() if r]
            sm = sorted(int(r[0]) for r in rows)
            out['sm_mhz'] = sm[len(sm) // 2]
            out['sm_max_mhz'] = int(rows[0][1])
            names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
            out['reasons'] = [n for i, n in enumerate(names) if any(r[2 + i].strip() == 'Active' for r in rows)]
            out['samples'] = len(rows)
        except Exception as e:  # nvidia-smi missing / no samples
            out['error'] = str(e)[:80]
        return out


class L2Flusher:
    """Evicts the 126 MB L2 between timed steps: a 256 MiB buffer is written (memset).  With
    DVSOF_FLUSH=write+read a second 256 MiB buffer is then READ, so that the lines left in L2 are
    clean -- otherwise the timed kernel also pays the write-back of the flusher's dirty lines."""
    MODE = os.environ.get('DVSOF_FLUSH', 'write')

    def __init__(self, device):
        self.buf = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2
        self.rd = torch.zeros(64 << 20, dtype=torch.int32, device=device) if self.MODE == 'write+read' else None

    def __call__(self):
        self.buf.zero_()
        if self.rd is not None:
            self.rd.sum()


def time_steps(fn, K, W, flush, inner=None):
    """W warm-ups, then K timed steps.  Each step has its own CUDA event pair on the launching
    (current) stream, the L2 flush sits between pairs.  Returns per-step ms list (+ inner ms)."""
    for _ in range(W):
        flush()
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for a, b in evs:
        flush()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in evs]


def bench_codec(name, K, W, dev):
    """Batch codec through the C ABI on the events of `name`: dvsof_decode_events (storage form ->
    int64 training form, 5 B read + 40 B written per event) and dvsof_encode_events (40 B read +
    5 B written), L2 flushed between steps."""
    import dvs_of_training_framework_b200 as dv
    L = dv._lib
    c = CFG[name]
    B, n = c['B'], c['n']
    N = B * n
    g = torch.Generator(device=dev).manual_seed(77)
    x = torch.randint(0, c['W'], (N,), generator=g, device=dev, dtype=torch.int16)
    y = torch.randint(0, c['H'], (N,), generator=g, device=dev, dtype=torch.int16)
    p = (torch.rand(N, generator=g, device=dev) < 0.7).view(torch.uint8)
    eps = torch.full((B,), 2, dtype=torch.int64, device=dev)              # two elements per sample
    epe = torch.full((2 * B,), n // 2, dtype=torch.int64, device=dev)
    epe[1::2] = n - n // 2
    out = [torch.empty(N, dtype=torch.int64, device=dev) for _ in range(5)]
    ws = torch.empty(L.load().dvsof_codec_workspace_bytes(2 * B, B), dtype=torch.uint8, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    sp = L.stream_ptr(dev)
    x2, y2, p2 = torch.empty_like(x), torch.empty_like(y), torch.empty_like(p)
    epe2 = torch.empty_like(epe)

    def dec():
        L.call('dvsof_decode_events', L.ptr(x), L.ptr(y), L.ptr(p), N, L.ptr(epe), 2 * B, L.ptr(eps), B,
               *[L.ptr(o) for o in out], L.ptr(ws), L.ptr(st), sp)

    def enc():
        L.call('dvsof_encode_events', *[L.ptr(o) for o in out], N, L.ptr(eps), B, 2 * B, L.ptr(x2), L.ptr(y2),
               L.ptr(p2), L.ptr(epe2), L.ptr(ws), L.ptr(st), sp)
    flush = L2Flusher(dev)
    dec_ms = float(np.mean(time_steps(dec, K, W, flush)))
    enc_ms = float(np.mean(time_steps(enc, K, W, flush)))
    assert int(st.item()) == 0 and torch.equal(x, x2) and torch.equal(p, p2) and torch.equal(epe, epe2)
    return {'events': N, 'decode_ms': dec_ms, 'encode_ms': enc_ms, 'bytes': 45 * N}


def bench_encoder(name, K, W, dev, seed_off=0):
    import dvs_of_training_framework_b200 as dv
    L = dv._lib
    c = CFG[name]
    B, H, Wd, n = c['B'], c['H'], c['W'], c['n']
    He, We = dv.extend_size((H, Wd))
    ev, ts, _ = fast_events(1234 + int(name[3:]) + seed_off, B, H, Wd, n, c['T'], dev)
    N = B * n
    out = torch.empty((B, 4, He, We), dtype=torch.float32, device=dev)
    ws = torch.empty(L.load().dvsof_encode_workspace_bytes(N, B, He, We), dtype=torch.uint8, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    sp = L.stream_ptr(dev)
    flush = L2Flusher(dev)

    def prepare():
        L.call('dvsof_encode_prepare', L.ptr(ev['timestamp']), L.ptr(ev['sample_index']), N,
               L.ptr(ts), B, He, We, L.ptr(out), L.ptr(ws), sp)

    def scatter():
        L.call('dvsof_encode_scatter', L.ptr(ev['x']), L.ptr(ev['y']), L.ptr(ev['timestamp']),
               L.ptr(ev['polarity']), L.ptr(ev['sample_index']), N, B, He, We, L.ptr(out),
               L.ptr(ws), L.ptr(status), sp)

    def whole():
        L.call('dvsof_encode_event_image', L.ptr(ev['x']), L.ptr(ev['y']), L.ptr(ev['timestamp']),
               L.ptr(ev['polarity']), L.ptr(ev['sample_index']), N, L.ptr(ts), B, He, We, L.ptr(out),
               L.ptr(ws), ws.numel(), L.ptr(status), sp)

    E = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    # (1) the step: one library call = prepare (zero + per-sample dt) -> scatter, PDL-chained
    for _ in range(W):
        flush(); whole()
    torch.cuda.synchronize()
    marks = [(E(), E()) for _ in range(K)]
    for e0, e1 in marks:
        flush()
        e0.record(); whole(); e1.record()
    torch.cuda.synchronize()
    step_ms = float(np.mean([e0.elapsed_time(e1) for e0, e1 in marks]))
    # (2) the dominant kernel alone (events on the launching stream around the scatter launch)
    for _ in range(W):
        flush(); prepare(); scatter()
    torch.cuda.synchronize()
    marks = [(E(), E()) for _ in range(K)]
    for k0, k1 in marks:
        flush(); prepare()
        k0.record(); scatter(); k1.record()
    torch.cuda.synchronize()
    assert int(status.item()) == 0, 'encoder flagged invalid input'
    assert int(out[:, :2].sum(dtype=torch.float64).item()) == N, 'event count not conserved'
    kern_ms = float(np.mean([k0.elapsed_time(k1) for k0, k1 in marks]))
    return dict(workload=name, events=N, step_ms=step_ms, kernel_ms=kern_ms,
                bytes=encoder_bytes(B, He, We, N), launches_per_step=2)


def bench_loss(name, K, W, dev):
    """fwd+bwd of the multiscale loss the way the train loop calls it
    (training.combined_loss -> Losses.terms -> one library call; loss.backward() -> one call)."""
    import dvs_of_training_framework_b200 as dv
    from dvs_of_training_framework_b200 import training as tr
    from tests.helpers import synth_loss_inputs
    c = CFG[name]
    B, H, Wd = c['B'], c['H'], c['W']
    inp = synth_loss_inputs(1234 + int(name[3:]), B, H, Wd, 3.0, device=dev)
    flows = [f.requires_grad_() for f in inp['flows']]
    ev = dv.Losses(inp['shapes'], B, dev)
    flush = L2Flusher(dev)

    def step():
        for f in flows:
            f.grad = None
        loss, _ = tr.combined_loss(ev, flows, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                                   inp['timestamps'], inp['sample_idx'], None, weights=[0.5, 1, 1])
        loss.backward()
        return loss

    ms = time_steps(step, K, W, flush)
    step_ms = float(np.mean(ms))
    # kernel level: the fused forward (all scales, ready-made pyramid) and backward launches alone
    L = dv._lib
    from dvs_of_training_framework_b200.loss import _scale_struct, interpolate
    fl = [f.detach() for f in flows]
    levels, cur = [], inp['images'].view(2 * B, H, Wd)
    for f in fl:
        cur = interpolate(cur, f.shape[-2:])
        levels.append(cur)
    grads = [torch.empty_like(f) for f in fl]
    sc = _scale_struct(fl, levels, grads)
    S = len(fl)
    ws = torch.empty(L.load().dvsof_loss_workspace_bytes(sc), dtype=torch.uint8, device=dev)
    terms = torch.empty(3 * S, device=dev)
    n_bad = torch.empty(S * B, dtype=torch.int32, device=dev)
    match = torch.cat((torch.arange(0, 2 * B, 2), torch.arange(1, 2 * B, 2))).to(torch.int32).to(dev)
    gt = torch.tensor([0.5 / S] * S + [1.0 / S] * S + [1.0 / S] * S, device=dev)
    sp = L.stream_ptr(dev)
    fwd = lambda: L.call('dvsof_warp_loss_forward', sc, L.ptr(match), L.ptr(terms), L.ptr(n_bad), L.ptr(ws), sp)  # noqa: E731
    bwd = lambda: L.call('dvsof_warp_loss_backward', sc, L.ptr(match), L.ptr(gt), L.ptr(n_bad), sp)  # noqa: E731
    fwd_ms = float(np.mean(time_steps(fwd, K, W, flush)))
    bwd_ms = float(np.mean(time_steps(bwd, K, W, flush)))
    px = sum(B * f.shape[-2] * f.shape[-1] for f in fl)
    # the fused training sweep alone (terms + d(loss)/d(flow) in one pass: 16 B read + 8 B written
    # per pixel), n_bad filled by one untimed call with the counting pre-pass
    import ctypes
    coef = (ctypes.c_float * (3 * S))(*([0.5 / S] * S + [1.0 / S] * S + [1.0 / S] * S))
    loss1 = torch.empty(1, device=dev)
    fus = lambda cp: L.call('dvsof_warp_loss_fused', sc, L.ptr(match), coef, L.ptr(terms), L.ptr(loss1),  # noqa: E731
                            L.ptr(n_bad), L.ptr(ws), cp, sp)
    fus(1)
    fused_ms = float(np.mean(time_steps(lambda: fus(0), K, W, flush)))
    fusedc_ms = float(np.mean(time_steps(lambda: fus(1), K, W, flush)))
    # the whole training pass through the C ABI alone (no Python autograd around it): one
    # dvsof_multiscale_loss_fused call (memset, match, 4 resize, border count, fused sweep) + the
    # backward's rescale launch -- what a compiled host (or a CUDA-graph replay) pays per step
    plan = ev._plan(dev, B, 2 * B, H, Wd, fl)
    o = plan.offs
    buf = torch.empty(plan.total, dtype=torch.float32, device=dev)
    gbuf = torch.empty(plan.gtotal, dtype=torch.float32, device=dev)
    from dvs_of_training_framework_b200.loss import _fill_frames
    fr = _fill_frames(plan, inp['images'], inp['flow_ts'], inp['flow_sample_idx'], inp['timestamps'],
                      inp['sample_idx'])
    plan.fill(plan.bwd, buf.data_ptr(), fl, gbuf.data_ptr())
    st = L.status_for(dev)
    one = torch.ones(1, device=dev)
    base = buf.data_ptr()

    def abi_step():
        L.call('dvsof_multiscale_loss_fused', plan.bwd, fr, coef, base + 4 * o[S], base + 4 * o[S + 2],
               base + 4 * o[S + 3], base + 4 * o[S + 1], plan.ws_ptr, L.ptr(st.word), sp)
        L.call('dvsof_scale_gradients', plan.bwd, L.ptr(one), sp)

    abi_ms = float(np.mean(time_steps(abi_step, K, W, flush)))
    st.check()
    return dict(workload=name, samples=B, step_ms=step_ms, abi_ms=abi_ms, bytes=loss_bytes(B, H, Wd),
                fwd_ms=fwd_ms, bwd_ms=bwd_ms, fwd_bytes=16 * px, bwd_bytes=24 * px,
                fused_ms=fused_ms, fused_with_count_ms=fusedc_ms, fused_bytes=24 * px,
                launches_per_step=9)


def fused_kernels(r, peak):
    """Kernel-level loss figures: the fused training sweep (24 B/pixel moved; SURVEY 8d counts the
    same work as 16 + 24 = 40 B/pixel for separate forward and backward passes) and the generic
    forward / backward kernels behind Losses.__call__."""
    frac = lambda b, ms: b / (ms * 1e-3) / 1e9 / peak  # noqa: E731
    return {'c_abi_step_ms': r['abi_ms'], 'c_abi_samples_per_s': r['samples'] / (r['abi_ms'] * 1e-3),
            'c_abi_step_hbm_frac_survey_bytes': frac(r['bytes'], r['abi_ms']),
            'fused_ms': r['fused_ms'], 'fused_with_count_ms': r['fused_with_count_ms'],
            'fused_hbm_frac_24B_per_px': frac(r['fused_bytes'], r['fused_ms']),
            'fused_hbm_frac_survey_40B_per_px': frac(r['fwd_bytes'] + r['bwd_bytes'], r['fused_ms']),
            'fwd_ms': r['fwd_ms'], 'bwd_ms': r['bwd_ms'],
            'fwd_hbm_frac': frac(r['fwd_bytes'], r['fwd_ms']), 'bwd_hbm_frac': frac(r['bwd_bytes'], r['bwd_ms'])}


def make_train_batch(B, H, W, n, T, seed):
    """A collated batch (utils/dataset.py:961-1020 layout) in pinned host memory."""
    ev, ts, sidx = fast_events(seed, B, H, W, n, T, 'cpu')
    ev['element_index'] = torch.zeros_like(ev['x'])
    g = torch.Generator().manual_seed(seed)
    images = torch.randint(0, 256, (2 * B, 1, H, W), generator=g).float()
    pin = (lambda t: t.pin_memory()) if torch.cuda.is_available() else (lambda t: t)
    return {'events': {k: pin(v) for k, v in ev.items()}, 'timestamps': pin(ts), 'sample_idx': pin(sidx),
            'images': pin(images), 'augmentation_params': None, 'size': B}


def encode_host_batch(batch):
    """The storage form of a collated host batch (utils/dataset.py:240-305 `encode_batch`, what
    scripts/prepare_batches.py writes and PreprocessedDataloader reads), in pinned memory."""
    ev = batch['events']
    B = int(batch['size'])
    per = batch['timestamps'].numel() // B
    pin = (lambda t: t.pin_memory()) if torch.cuda.is_available() else (lambda t: t)
    counts = torch.bincount(ev['sample_index'], minlength=B)           # one element per sample here
    return {'events': {'x': pin(ev['x'].to(torch.int16)), 'y': pin(ev['y'].to(torch.int16)),
                       'timestamp': pin(ev['timestamp']), 'polarity': pin(ev['polarity'] > 0),
                       'events_per_element': pin(counts)},
            'elements_per_sample': pin(torch.full((B,), per - 1, dtype=torch.uint8)),
            'timestamps': pin(batch['timestamps']), 'images': pin(batch['images'].to(torch.uint8)),
            'augmentation_params': None}


def bench_train_step(K, W, dev, world, per_gpu=32, encoded=False):
    """EV-FlowNet data-parallel training step (BASELINE configs[4] shapes: 256x256, 131072
    events/sample; `per_gpu` samples per GPU => global batch 256 at 8 GPUs): H2D of the shard,
    encode, cuDNN forward, fused loss, backward with overlapped NCCL allreduce, AdamW step."""
    import torch.distributed as dist
    import dvs_of_training_framework_b200 as dv
    from dvs_of_training_framework_b200 import training as tr
    from dvs_of_training_framework_b200.evflownet import Model
    c = CFG['cfg2']
    rank = dist.get_rank() if world > 1 else 0
    torch.manual_seed(0)
    # channels_last end to end: the glue kernels have (N, H, W, C) variants and the decoder inputs are
    # filled up to a multiple of 8 channels, so cuDNN runs without layout or padding copies
    # (15.9 ms/step against 21.1 ms in NCHW, scripts/exp_train_variants.py)
    model = Model(dev).to(memory_format=torch.channels_last)
    red = tr.GradientAllreducer(model.parameters())
    red.broadcast_parameters(model)
    # train_flownet.py:59-61; torch's single-kernel implementation of the same optimizer (the
    # gradient views carry the parameters' strides, training.GradientAllreducer)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, amsgrad=True, fused=True)
    evaluator = dv.init_losses((c['H'], c['W']), per_gpu, model, dev, 1)
    batch = make_train_batch(per_gpu, c['H'], c['W'], c['n'], c['T'], 1234 + 5 + rank)
    if encoded:             # 9 B/event + uint8 images cross PCIe, dataset.decode_batch expands on device
        batch = encode_host_batch(batch)
    timers = dv.timer.FakeTimer()
    model.train()
    # every step uploads its batch from pinned host memory; the upload of step k+1 runs on a copy
    # stream during step k (training.DevicePrefetcher), as a prefetching loader does
    import itertools
    feed = iter(tr.DevicePrefetcher(itertools.repeat(batch, W + K + 1), dev, True))

    def step():
        loss, _, _ = tr.process_minibatch(model, next(feed), timers, dev, True, evaluator, [0.5, 1, 1])
        loss.backward()
        red.wait()
        opt.step()
        red.zero_grad()
        return loss

    for _ in range(W):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    assert torch.isfinite(loss).item(), 'train step produced a non-finite loss'
    return dict(samples=per_gpu, step_ms=e0.elapsed_time(e1) / K, params=sum(p.numel() for p in model.parameters()))


def bench_e2e(name, K, W, dev, compact=False):
    """Public API, host buffers.  Every step: H2D of that step's event arrays from pinned host
    memory, `dv.quantize` (the reference-facing call), D2H of the event image into pinned memory
    -- all inside the timed region.  Steps alternate between two CUDA streams (user-level double
    buffering) so step k's D2H overlaps step k+1's H2D on the full-duplex link; a step's pinned
    output is only reused after its own D2H has completed.
    compact=True: the storage-form entry (utils/dataset.py:286-289 int16/bool/fp32, 9 B/event)."""
    import dvs_of_training_framework_b200 as dv
    c = CFG[name]
    B, H, Wd, n = c['B'], c['H'], c['W'], c['n']
    ev, ts, sidx = fast_events(1234 + int(name[3:]), B, H, Wd, n, c['T'], dev)
    if compact:
        host = {'x': ev['x'].to(torch.int16), 'y': ev['y'].to(torch.int16), 'timestamp': ev['timestamp'],
                'polarity': (ev['polarity'] > 0), 'eps': torch.full((B,), n, dtype=torch.int64, device=dev)}
    else:
        host = {k: v for k, v in ev.items()}
    host = {k: v.cpu().pin_memory() for k, v in host.items()}
    ts_h, sidx_h = ts.cpu().pin_memory(), sidx.cpu().pin_memory()
    He, We = dv.extend_size((H, Wd))
    outs = [torch.empty((B, 4, He, We), dtype=torch.float32).pin_memory() for _ in range(2)]
    h2d = sum(v.numel() * v.element_size() for v in host.values()) + ts_h.numel() * 4 + sidx_h.numel() * 8
    d2h = outs[0].numel() * 4
    streams = [torch.cuda.Stream(dev) for _ in range(2)]
    done = [None, None]

    def step(i):
        k = i & 1
        if done[k] is not None:
            done[k].synchronize()          # the pinned output of step i-2 has landed
        with torch.cuda.stream(streams[k]):
            d = {kk: v.to(dev, non_blocking=True) for kk, v in host.items()}
            t_d = ts_h.to(dev, non_blocking=True)
            if compact:
                img = dv.compute_event_image_compact(d['x'], d['y'], d['timestamp'], d['polarity'],
                                                     d['eps'], t_d, (He, We))
            else:
                img = dv.quantize(d, t_d, sidx_h.to(dev, non_blocking=True), (H, Wd))
            outs[k].copy_(img, non_blocking=True)
            done[k] = torch.cuda.Event()
            done[k].record()

    for i in range(2 * ((W + 1) // 2)):
        step(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(K):
        step(i)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / K
    for o in outs[:min(K, 2)]:
        assert int(o[:, :2].sum(dtype=torch.float64)) == B * n
    return dict(events=B * n, step_s=dt, h2d=h2d, d2h=d2h)


def cpu_encoder_pass(name, threads=None):
    """One pass of the oracle's torch-CPU port of the reference encoder on the workload."""
    from oracle import encoder_oracle_torch as eot
    import dvs_of_training_framework_b200 as dv
    c = CFG[name]
    # all the host threads this process may use -- torchrun exports OMP_NUM_THREADS=1, which would
    # time the reference on one core
    torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
    ev, ts, sidx = fast_events(1234 + int(name[3:]), c['B'], c['H'], c['W'], c['n'], c['T'], 'cpu')
    size = dv.extend_size((c['H'], c['W']))
    return (lambda: eot.compute_event_image(ev, ts, sidx, size)), c['B'] * c['n']


def cpu_baseline(name, budget_s=12.0):
    fn, N = cpu_encoder_pass(name)
    fn()
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < budget_s:
        fn()
        reps += 1
    dt = (time.perf_counter() - t0) / reps
    return {'value': N / dt / 1e6, 'unit': 'Mevents/s', 'cores': torch.get_num_threads(),
            'kind': 'port', 'sample': f'{reps} full passes of {name} ({N} events) in {budget_s:.0f} s, '
            'oracle/encoder_oracle_torch.py (the reference ATen op sequence, CPU)',
            'host_cpus': len(os.sched_getaffinity(0))}


def run_reference(args, rank):
    if rank != 0:
        return
    fn, N = cpu_encoder_pass(HEADLINE)
    for _ in range(args.warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = (time.perf_counter() - t0) / args.steps
    v = N / dt / 1e6
    c = CFG[HEADLINE]
    line = {'impl': 'reference', 'metric': 'encoded Mevents/s', 'value': v, 'unit': 'Mevents/s',
            'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
            'data': 'synthetic',
            'config': {'workload': f'{HEADLINE}: MVSEC-shaped 256x256 events, batch {c["B"]}, '
                                   f'{c["n"]} events/sample -> (B,4,256,256) event image'},
            'cpu_baseline': {'value': v, 'unit': 'Mevents/s', 'cores': torch.get_num_threads(),
                             'kind': 'port', 'sample': f'{args.steps} full passes of {HEADLINE}; the '
                             'reference is Python+torch_scatter and cannot travel: oracle torch-CPU port'},
            'e2e': {'value': v, 'unit': 'Mevents/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    emit(line)


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner, cuDNN warnings) write to fd 1; the driver wants exactly
    ONE JSON line there.  Route fd 1 to stderr for the run and keep the real one for emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + '\n').encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--quick', action='store_true', help='headline workload only (for ncu runs)')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' and not args.quick else args.warmup
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: the product path has no CPU fallback')
    import dvs_of_training_framework_b200 as dv
    dv._lib.load()
    dv._lib.Status.mode = 'deferred'
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    K, W = args.steps, args.warmup

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sync_all()
    with Clocks(local) as clk:
        head = bench_encoder(HEADLINE, K, W, dev, seed_off=rank)
        sync_all()
        loss = bench_loss(HEADLINE, K, W, dev)
        sync_all()
    clocks = clk.summary()
    head_ms = max_over_ranks(head['step_ms'])
    kern_ms = max_over_ranks(head['kernel_ms'])
    loss_ms = max_over_ranks(loss['abi_ms'])
    loss_py_ms = max_over_ranks(loss['step_ms'])
    e2e = bench_e2e(HEADLINE, max(3, K // 5), 3, dev)
    e2e_s = max_over_ranks(e2e['step_s'])
    e2e_c = bench_e2e(HEADLINE, max(3, K // 5), 3, dev, compact=True)
    e2e_c_s = max_over_ranks(e2e_c['step_s'])
    sync_all()
    train = None
    if not args.quick:
        train = bench_train_step(max(3, min(K, 10)), 3, dev, world)
        train['step_ms'] = max_over_ranks(train['step_ms'])
        torch.cuda.empty_cache()
        sync_all()
        train_enc = bench_train_step(max(3, min(K, 10)), 3, dev, world, encoded=True)
        train['encoded_step_ms'] = max_over_ranks(train_enc['step_ms'])
        torch.cuda.empty_cache()
    sync_all()
    extra = {}
    if not args.quick:
        for name in ('cfg3', 'cfg4'):
            k2 = max(3, min(K, 10))
            r = bench_encoder(name, k2, 3, dev, seed_off=rank)
            r['step_ms'], r['kernel_ms'] = max_over_ranks(r['step_ms']), max_over_ranks(r['kernel_ms'])
            extra[name] = r
            torch.cuda.empty_cache()
            ll = bench_loss(name, k2, 3, dev)
            ll['step_ms'] = max_over_ranks(ll['step_ms'])
            extra[name + '_loss'] = ll
            torch.cuda.empty_cache()
    codec = bench_codec(HEADLINE, max(3, min(K, 10)), 3, dev) if (rank == 0 and not args.quick) else None
    torch.cuda.empty_cache()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    c = CFG[HEADLINE]
    achieved = head['bytes'] / (kern_ms * 1e-3) / 1e9
    line = {
        'metric': 'encoded Mevents/s', 'value': world * head['events'] / (head_ms * 1e-3) / 1e6,
        'unit': 'Mevents/s', 'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': head_ms,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic',
        'config': {'workload': f'{HEADLINE}: MVSEC-shaped 256x256 events, batch {c["B"]}/GPU, '
                               f'{c["n"]} events/sample -> (B,4,256,256) event image; int64 SoA '
                               'events (36 B/event) resident in HBM',
                   'l2': '256 MiB memset between steps (outside the event pairs)',
                   'validation': 'deferred (device status word, no host sync per call)',
                   'parallelism': f'replicas x{world} (sharded by sample, no data-path collective)'},
        'clocks': clocks,
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                     'frac': achieved / peak, 'traffic': TRAFFIC.get(HEADLINE), 'kernel': 'encode_vec4_kernel',
                     'kernel_ms': kern_ms, 'algorithmic_bytes': head['bytes'], 'peak_source': peak_src},
        'e2e': {'value': world * e2e['events'] / e2e_s / 1e6, 'unit': 'Mevents/s',
                'h2d_bytes_per_step': e2e['h2d'], 'd2h_bytes_per_step': e2e['d2h'],
                'api': 'dv.quantize(events, timestamps, sample_idx, imsize) from pinned host tensors; '
                       'two streams, step k D2H overlaps step k+1 H2D',
                'compact_storage_form': {'value': world * e2e_c['events'] / e2e_c_s / 1e6, 'unit': 'Mevents/s',
                                         'h2d_bytes_per_step': e2e_c['h2d'], 'd2h_bytes_per_step': e2e_c['d2h'],
                                         'api': 'dv.compute_event_image_compact (int16/bool/fp32 storage form, 9 B/event)'}},
        'gpu_launches': K * head['launches_per_step'],
        'warp_loss': {'metric': 'warp-loss fwd+bwd samples/s', 'value': world * loss['samples'] / (loss_ms * 1e-3),
                      'unit': 'samples/s', 'ms_per_step': loss_ms, 'workload': f'{HEADLINE}: 4 scales, 256x256, B={c["B"]}',
                      'roofline': {'bound': 'hbm', 'achieved': loss['bytes'] / (loss_ms * 1e-3) / 1e9,
                                   'peak': peak, 'unit': 'GB/s', 'frac': loss['bytes'] / (loss_ms * 1e-3) / 1e9 / peak,
                                   'note': 'whole training pass through the C ABI, device-resident inputs: dvsof_multiscale_loss_fused (one prelude launch: timestamp match + chained pyramid + border count; then the fused terms-and-gradient sweep, PDL-chained) + dvsof_scale_gradients; bytes = SURVEY 8d figure (separate fwd + bwd + pyramid)'},
                      'python_api': {'value': world * loss['samples'] / (loss_py_ms * 1e-3), 'unit': 'samples/s',
                                     'ms_per_step': loss_py_ms,
                                     'note': 'training.combined_loss(...) + loss.backward() through torch.autograd (same kernels; host-bound: Python + autograd engine per step)'},
                      'kernels': fused_kernels(loss, peak)},
        'workloads': {},
    }
    if train is not None:
        line['train_step'] = {'metric': 'train-step samples/s', 'value': world * train['samples'] / (train['step_ms'] * 1e-3),
                              'unit': 'samples/s', 'ms_per_step': train['step_ms'],
                              'workload': f'EV-FlowNet ({train["params"]} params, fp32/TF32 cuDNN convolutions; upsample+reflect-pad and ReLU+BatchNorm through libdvsof), 256x256, 131072 events/sample, '
                                          f'{train["samples"]} samples/GPU (global {world * train["samples"]}), AdamW(amsgrad, fused), '
                                          'H2D of every batch from pinned memory, prefetched one step ahead on a copy stream; flat-bucket NCCL allreduce overlapped with backward',
                              'storage_form_feed': {'value': world * train['samples'] / (train['encoded_step_ms'] * 1e-3), 'unit': 'samples/s',
                                                    'ms_per_step': train['encoded_step_ms'],
                                                    'note': 'the same step fed with batches in the storage form of utils/dataset.py encode_batch (9 B/event, uint8 images) and dataset.decode_batch on the device: a quarter of the upload'}}
    for name, r in extra.items():
        if name.endswith('_loss'):
            line['workloads'][name] = {'samples_per_s': world * r['samples'] / (r['abi_ms'] * 1e-3),
                                       'ms_per_step': r['abi_ms'],
                                       'hbm_frac_step': r['bytes'] / (r['abi_ms'] * 1e-3) / 1e9 / peak,
                                       'python_api_ms_per_step': r['step_ms'],
                                       'kernels': fused_kernels(r, peak)}
        else:
            line['workloads'][name] = {'mevents_per_s': world * r['events'] / (r['step_ms'] * 1e-3) / 1e6,
                                       'ms_per_step': r['step_ms'], 'kernel_ms': r['kernel_ms'],
                                       'hbm_gbs_kernel': r['bytes'] / (r['kernel_ms'] * 1e-3) / 1e9,
                                       'hbm_frac_kernel': r['bytes'] / (r['kernel_ms'] * 1e-3) / 1e9 / peak,
                                       # the whole dvsof_encode_event_image call (prepare + the path the
                                       # library picks: direct scatter, or bin + reduce for dense streams)
                                       'hbm_frac_step': r['bytes'] / (r['step_ms'] * 1e-3) / 1e9 / peak,
                                       'note': 'kernel_* = the direct scatter kernel alone; ms_per_step / '
                                               'hbm_frac_step = the library call on the path it selects'}
    if codec is not None:
        line['workloads']['codec_' + HEADLINE] = {
            'decode_mevents_per_s': codec['events'] / (codec['decode_ms'] * 1e-3) / 1e6, 'decode_ms': codec['decode_ms'],
            'decode_hbm_frac': codec['bytes'] / (codec['decode_ms'] * 1e-3) / 1e9 / peak,
            'encode_mevents_per_s': codec['events'] / (codec['encode_ms'] * 1e-3) / 1e6, 'encode_ms': codec['encode_ms'],
            'encode_hbm_frac': codec['bytes'] / (codec['encode_ms'] * 1e-3) / 1e9 / peak,
            'note': 'dvsof_decode_events / dvsof_encode_events (utils/dataset.py decode_batch / encode_batch, event part), 45 B/event, one GPU (rank 0)'}
    line['cpu_baseline'] = cpu_baseline(HEADLINE) if world == 1 and not args.quick else None
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
