#!/usr/bin/env python
"""bench.py -- hot-path benchmark of dvs_of_training_framework_b200 on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    (N > 1: launched by torchrun, one rank per GPU)

Headline metric (BASELINE.json): encoded ("voxelized") Mevents/s on the largest single-GPU
configuration, configs[2] shapes -- MVSEC-shaped 256x256 events, batch 32, 5 M events/sample
(SURVEY.md 8d cfg3).  A step = one pass of the encoder over one synthetic batch, inputs resident
in HBM, L2 flushed between steps.  The same JSON line also carries the warp-loss fwd+bwd samples/s
(`warp_loss`), the other configurations (`workloads`: cfg1, cfg2, cfg2 with the contended event
distribution, cfg4), the end-to-end number through the public Python API from pinned host buffers
(`e2e`), the roofline of the dominant kernel and the CPU baselines (oracle ports, timed on this
box's host cores).
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
HEADLINE = 'cfg3'
DESCR = {'cfg1': 'DAVIS-shaped 260x346 events, batch 1, 32768 events/sample -> (1,4,272,352)',
         'cfg2': 'MVSEC-shaped 256x256 events, batch 8, 131072 events/sample -> (8,4,256,256)',
         'cfg3': 'MVSEC-shaped 256x256 events at high rate, batch 32, 5 M events/sample -> (32,4,256,256)',
         'cfg4': 'DSEC-shaped 480x640 events, batch 16, 1 M events/sample -> (16,4,480,640)'}


def traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` on `workload`, from the
    committed `ncu --set full` captures (profiles/traffic.json, written by scripts/ncu_traffic.py
    from the .ncu-rep files named there); None when that pair was not captured."""
    f = ROOT / 'profiles' / 'traffic.json'
    if not f.exists():
        return None
    return json.loads(f.read_text()).get(workload, {}).get(kernel, {}).get('dram_bytes')


def peaks():
    f = ROOT / 'MEASURED_PEAKS.json'
    if f.exists():
        return float(json.loads(f.read_text())['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


def fast_events(seed, B, H, W, n, T, device, contended=False):
    """Synthetic events of SURVEY 8d generated on `device` (same distribution as
    tests/helpers.synth_events; torch device RNG so cfg3's 160 M events take milliseconds).
    contended: 80 % of the events fall on a random 10 % of the pixels (SURVEY 8d's second set)."""
    g = torch.Generator(device=device).manual_seed(seed)
    N = B * n
    x = torch.randint(0, W, (N,), generator=g, device=device)
    y = torch.randint(0, H, (N,), generator=g, device=device)
    if contended:
        hot = torch.randperm(H * W, generator=g, device=device)[:max(1, H * W // 10)]
        pick = hot[torch.randint(0, hot.numel(), (N,), generator=g, device=device)]
        use = torch.rand(N, generator=g, device=device) < 0.8
        x = torch.where(use, pick % W, x)
        y = torch.where(use, pick // W, y)
    ev = {'x': x, 'y': y,
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


def graph_time(fn, dev, iters=100, reps=3):
    """Warm figure: `iters` back-to-back calls of `fn` captured in ONE CUDA graph and
    replayed -- no host launch cost, no L2 flush between iterations (inputs that fit the 126 MB L2
    stay resident).  Returns ms per call (best of `reps` replays)."""
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        for _ in range(2):
            fn()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(iters):
            fn()
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        best = ms if best is None else min(best, ms)
    return best


def bench_encoder(name, K, W, dev, seed_off=0, contended=False):
    """The encoder through the C ABI on workload `name`: the whole call (a step), the dominant
    kernel(s) alone with CUDA events on the launching stream, and the graph-captured warm figure."""
    import dvs_of_training_framework_b200 as dv
    L = dv._lib
    c = CFG[name]
    B, H, Wd, n = c['B'], c['H'], c['W'], c['n']
    He, We = dv.extend_size((H, Wd))
    ev, ts, _ = fast_events(1234 + int(name[3:]) + seed_off, B, H, Wd, n, c['T'], dev, contended)
    N = B * n
    out = torch.empty((B, 4, He, We), dtype=torch.float32, device=dev)
    ws = torch.empty(L.load().dvsof_encode_workspace_bytes(N, B, He, We), dtype=torch.uint8, device=dev)
    status = L.status_for(dev).word
    sp = lambda: L.stream_ptr(dev)  # noqa: E731  (the current stream: differs under graph capture)
    flush = L2Flusher(dev)
    evp = [L.ptr(ev[k]) for k in ('x', 'y', 'timestamp', 'polarity', 'sample_index')]
    binned = bool(L.load().dvsof_encode_path(N, B, He, We, ws.numel()))

    def prepare():
        L.call('dvsof_encode_prepare', L.ptr(ev['timestamp']), L.ptr(ev['sample_index']), N,
               L.ptr(ts), B, He, We, L.ptr(out), L.ptr(ws), sp())

    def scatter():
        L.call('dvsof_encode_scatter', *evp, N, B, He, We, L.ptr(out), L.ptr(ws), L.ptr(status), sp())

    def bin_pass(which):
        L.call('dvsof_encode_binned_pass', *evp, N, B, He, We, L.ptr(out), L.ptr(ws), ws.numel(),
               L.ptr(status), which, sp())

    def whole():
        L.call('dvsof_encode_event_image', *evp, N, L.ptr(ts), B, He, We, L.ptr(out),
               L.ptr(ws), ws.numel(), L.ptr(status), sp())

    def timed(pre, fn):
        """K event pairs around fn(); pre() (untimed, after the flush) sets up what fn needs."""
        for _ in range(W):
            flush(); pre(); fn()
        torch.cuda.synchronize()
        marks = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for e0, e1 in marks:
            flush(); pre()
            e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        return float(np.mean([e0.elapsed_time(e1) for e0, e1 in marks]))

    # (1) the step: one library call = prepare (zero + per-sample dt) -> the path the library picks
    step_ms = timed(lambda: None, whole)
    assert int(out[:, :2].sum(dtype=torch.float64).item()) == N, 'event count not conserved'
    # (2) the kernels alone (events on the launching stream around one launch)
    kern = {'scatter_ms': timed(prepare, scatter)}
    assert int(out[:, :2].sum(dtype=torch.float64).item()) == N, 'event count not conserved'
    if binned:
        kern['bin_ms'] = timed(prepare, lambda: bin_pass(1))
        kern['reduce_ms'] = timed(lambda: (prepare(), bin_pass(1)), lambda: bin_pass(2))
        assert int(out[:, :2].sum(dtype=torch.float64).item()) == N, 'event count not conserved (binned passes)'
    L.status_for(dev).check()
    # (3) warm: 100 graph-captured calls, no flush
    warm_ms = graph_time(whole, dev)
    return dict(workload=name, events=N, step_ms=step_ms, kernel_ms=kern['scatter_ms'], kernels=kern,
                binned=binned, warm_graph_ms=warm_ms, bytes=encoder_bytes(B, He, We, N),
                launches_per_step=3 if binned else 2, cells=B * 4 * He * We, contended=contended)


def bench_encoder_compact(name, K, W, dev):
    """The storage-form entry (utils/dataset.py:286-289: x, y int16, polarity bool, t fp32 = 9 B/event)
    through the C ABI on workload `name`; its own roofline: 9 N + 8 (B + 1) + 16 B H' W' bytes."""
    import dvs_of_training_framework_b200 as dv
    L = dv._lib
    c = CFG[name]
    B, H, Wd, n = c['B'], c['H'], c['W'], c['n']
    He, We = dv.extend_size((H, Wd))
    ev, ts, _ = fast_events(1234 + int(name[3:]), B, H, Wd, n, c['T'], dev)
    N = B * n
    x, y = ev['x'].to(torch.int16), ev['y'].to(torch.int16)
    p, t = (ev['polarity'] > 0).view(torch.uint8), ev['timestamp']
    begin = torch.arange(B + 1, device=dev, dtype=torch.int64) * n
    del ev
    out = torch.empty((B, 4, He, We), dtype=torch.float32, device=dev)
    ws = torch.empty(L.load().dvsof_encode_workspace_bytes(N, B, He, We), dtype=torch.uint8, device=dev)
    st = L.status_for(dev)
    binned = bool(L.load().dvsof_encode_path(N, B, He, We, ws.numel()))

    def whole():
        L.call('dvsof_encode_event_image_compact', L.ptr(x), L.ptr(y), L.ptr(t), L.ptr(p), L.ptr(begin), N,
               L.ptr(ts), B, He, We, L.ptr(out), L.ptr(ws), ws.numel(), L.ptr(st.word), L.stream_ptr(dev))
    ms = float(np.mean(time_steps(whole, K, W, L2Flusher(dev))))
    assert int(out[:, :2].sum(dtype=torch.float64).item()) == N, 'event count not conserved'
    st.check()
    return dict(events=N, step_ms=ms, warm_graph_ms=graph_time(whole, dev), binned=binned,
                bytes=9 * N + 8 * (B + 1) + 16 * B * He * We)


VOXEL_BINS = {'cfg1': 5, 'cfg2': 9, 'cfg3': 9, 'cfg4': 15}      # BASELINE.json configs: 5 / 9 / 15 bins


def bench_voxel(name, K, W, dev, cpu_budget_s=3.0):
    """EXTENSION (not in the reference, parity unpinned): the temporal-bilinear voxel grid of
    csrc/voxel.cuh through the C ABI on workload `name` with BASELINE.json's bin count; its own
    roofline: 36 N + 8 B + 4 C B H' W' bytes.  The CPU figure beside it is oracle/voxel_oracle.py
    (numpy, one thread) on a bounded sample."""
    import dvs_of_training_framework_b200 as dv
    L = dv._lib
    c = CFG[name]
    B, H, Wd, n, C = c['B'], c['H'], c['W'], c['n'], VOXEL_BINS[name]
    He, We = dv.extend_size((H, Wd))
    ev, ts, sidx = fast_events(4321 + int(name[3:]), B, H, Wd, n, c['T'], dev)
    N = B * n
    out = torch.empty((B, C, He, We), dtype=torch.float32, device=dev)
    ws = torch.empty(L.load().dvsof_voxel_workspace_bytes(N, B, He, We, C), dtype=torch.uint8, device=dev)
    st = L.status_for(dev)
    binned = bool(L.load().dvsof_voxel_path(N, B, He, We, C, ws.numel()))
    ptrs = [L.ptr(ev[k]) for k in ('x', 'y', 'timestamp', 'polarity', 'sample_index')]

    def whole():
        L.call('dvsof_encode_voxel_grid', *ptrs, N, L.ptr(ts), B, C, He, We, L.ptr(out), L.ptr(ws), ws.numel(),
               L.ptr(st.word), L.stream_ptr(dev))
    ms = float(np.mean(time_steps(whole, K, W, L2Flusher(dev))))
    net = float(out.sum(dtype=torch.float64).item())
    assert abs(net - float(ev['polarity'].sum().item())) <= 1e-4 * N, 'polarity sum not conserved'
    st.check()
    # CPU: the oracle on the first sample's events (bounded), checked against the device result
    from oracle import voxel_oracle as vo
    m = min(n, 500_000)
    sub = {k: v[:m].cpu().numpy() for k, v in ev.items()}
    t0, passes = time.perf_counter(), 0
    while passes == 0 or time.perf_counter() - t0 < cpu_budget_s:
        vo.voxel_grid(sub, ts[:2].cpu().numpy(), sidx[:2].cpu().numpy(), (He, We), C)
        passes += 1
    cpu_s = (time.perf_counter() - t0) / passes
    return dict(events=N, step_ms=ms, warm_graph_ms=graph_time(whole, dev), binned=binned, bins=C,
                bytes=36 * N + 8 * B + 4 * C * B * He * We,
                cpu={'value': m / cpu_s / 1e6, 'unit': 'Mevents/s', 'cores': 1, 'kind': 'port',
                     'sample': f'{passes} passes over the first {m} events of sample 0 in {cpu_budget_s:.0f} s, '
                               'oracle/voxel_oracle.py (numpy)'})


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

    def abi_step_graph():                         # the same calls on whatever stream is current (capture)
        spg = L.stream_ptr(dev)
        L.call('dvsof_multiscale_loss_fused', plan.bwd, fr, coef, base + 4 * o[S], base + 4 * o[S + 2],
               base + 4 * o[S + 3], base + 4 * o[S + 1], plan.ws_ptr, L.ptr(st.word), spg)
        L.call('dvsof_scale_gradients', plan.bwd, L.ptr(one), spg)
    warm_ms = graph_time(abi_step_graph, dev)
    return dict(workload=name, samples=B, step_ms=step_ms, abi_ms=abi_ms, bytes=loss_bytes(B, H, Wd),
                fwd_ms=fwd_ms, bwd_ms=bwd_ms, fwd_bytes=16 * px, bwd_bytes=24 * px,
                fused_ms=fused_ms, fused_with_count_ms=fusedc_ms, fused_bytes=24 * px,
                warm_graph_ms=warm_ms, launches_per_step=4, pixels=px)


def fused_kernels(r, peak):
    """Kernel-level loss figures: the fused training sweep (24 B/pixel moved; SURVEY 8d counts the
    same work as 16 + 24 = 40 B/pixel for separate forward and backward passes) and the generic
    forward / backward kernels behind Losses.__call__."""
    frac = lambda b, ms: b / (ms * 1e-3) / 1e9 / peak  # noqa: E731
    return {'c_abi_step_ms': r['abi_ms'], 'c_abi_samples_per_s': r['samples'] / (r['abi_ms'] * 1e-3),
            'c_abi_step_hbm_frac_survey_bytes': frac(r['bytes'], r['abi_ms']),
            'c_abi_step_warm_graph_ms': r['warm_graph_ms'],
            'fused_ms': r['fused_ms'], 'fused_with_count_ms': r['fused_with_count_ms'],
            'fused_hbm_frac_24B_per_px': frac(r['fused_bytes'], r['fused_ms']),
            'fused_hbm_frac_survey_40B_per_px': frac(r['fwd_bytes'] + r['bwd_bytes'], r['fused_ms']),
            'fwd_ms': r['fwd_ms'], 'bwd_ms': r['bwd_ms'],
            'fwd_hbm_frac': frac(r['fwd_bytes'], r['fwd_ms']), 'bwd_hbm_frac': frac(r['bwd_bytes'], r['bwd_ms'])}


def make_train_samples(B, H, W, n, T, seed):
    """B per-sample tuples (events dict, timestamps, images) as DatasetImpl.__getitem__ returns them
    (utils/dataset.py:776-797): int64 x / y / polarity / element_index, float32 timestamp, two
    grey frames -- the input of `collate_wrapper`."""
    ev, ts, _ = fast_events(seed, B, H, W, n, T, 'cpu')
    g = torch.Generator().manual_seed(seed)
    images = torch.randint(0, 256, (2 * B, H, W), generator=g, dtype=torch.uint8).numpy()
    out = []
    for b in range(B):
        sl = slice(b * n, (b + 1) * n)
        events = {'x': ev['x'][sl].numpy(), 'y': ev['y'][sl].numpy(), 'timestamp': ev['timestamp'][sl].numpy(),
                  'polarity': ev['polarity'][sl].numpy(), 'element_index': np.zeros(n, dtype=np.int64)}
        out.append((events, ts[2 * b:2 * b + 2].numpy().astype(np.float64), images[2 * b:2 * b + 2].astype(np.float32)))
    return out


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
    # the batch as the loader's collate_fn hands it over: dv.collate_wrapper of per-sample tuples, in the
    # reference's int64 layout (36 B/event + float32 images) or, encoded=True, assembled directly in the
    # storage form (9 B/event + uint8 images; widened by dataset.decode_batch on the device)
    batch = dv.collate_wrapper(make_train_samples(per_gpu, c['H'], c['W'], c['n'], c['T'], 1234 + 5 + rank),
                               encoded=encoded)
    timers = dv.timer.FakeTimer()
    model.train()
    # every step uploads its batch from pinned host memory; the upload of step k+1 runs on a copy
    # stream during step k (training.DevicePrefetcher), as a prefetching loader does
    import itertools
    feed = iter(tr.DevicePrefetcher(itertools.repeat(batch, W + K + 4), dev, True))

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
    step_ms = e0.elapsed_time(e1) / K
    dp = None
    if world > 1 and not encoded:
        # data-parallel self-check, outside the timed region: the flat gradient the bucketed NCCL
        # allreduce leaves on rank 0 == the mean of every rank's local gradient (gathered to rank 0),
        # for the same weights and each rank's own shard
        def grads(reduce):
            red.enabled = reduce
            red.zero_grad()
            lo_, _, _ = tr.process_minibatch(model, next(feed), timers, dev, True, evaluator, [0.5, 1, 1])
            lo_.backward()
            red.wait()
            red.enabled = True
            return torch.cat([b['flat'] for b in red.buckets]).clone()
        local, reduced = grads(False), grads(True)
        gathered = [torch.empty_like(local) for _ in range(world)] if rank == 0 else None
        dist.gather(local, gathered, dst=0)
        if rank == 0:
            mean = torch.stack(gathered).mean(0)
            scale = float(mean.abs().max())
            err = float((reduced - mean).abs().max())
            dp = {'max_abs_err': err, 'grad_max_abs': scale, 'rel': err / max(scale, 1e-30),
                  'ok': bool(err <= 1e-5 * scale + 1e-9), 'elements': int(mean.numel()),
                  'what': 'allreduced flat gradient on rank 0 vs mean of the per-rank local gradients (dist.gather)'}
        red.zero_grad()
    return dict(samples=per_gpu, step_ms=step_ms, params=sum(p.numel() for p in model.parameters()), dp_check=dp)


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


REF_SAMPLES = 8        # the CPU arms run this many of the headline workload's samples per pass


def cpu_encoder_pass(name, threads=None, samples=None):
    """One pass of the oracle's torch-CPU port of the reference encoder on (a sample of) the workload:
    the first `samples` samples of it, same events per sample, same image size."""
    from oracle import encoder_oracle_torch as eot
    import dvs_of_training_framework_b200 as dv
    c = CFG[name]
    B = min(samples or c['B'], c['B'])
    # all the host threads this process may use -- torchrun exports OMP_NUM_THREADS=1, which would
    # time the reference on one core
    torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
    ev, ts, sidx = fast_events(1234 + int(name[3:]), B, c['H'], c['W'], c['n'], c['T'], 'cpu')
    size = dv.extend_size((c['H'], c['W']))
    return (lambda: eot.compute_event_image(ev, ts, sidx, size)), B * c['n'], B


def _timed_passes(fn, budget_s):
    fn()
    t0 = time.perf_counter()
    reps = 0
    while reps < 1 or time.perf_counter() - t0 < budget_s:
        fn()
        reps += 1
    return (time.perf_counter() - t0) / reps, reps


def cpu_baseline(name, budget_s=12.0):
    fn, N, B = cpu_encoder_pass(name, samples=REF_SAMPLES)
    dt, reps = _timed_passes(fn, budget_s)
    return {'value': N / dt / 1e6, 'unit': 'Mevents/s', 'cores': torch.get_num_threads(),
            'kind': 'port', 'sample': f'{reps} passes over {B} of the {CFG[name]["B"]} samples of {name} ({N} events '
            f'per pass) in {budget_s:.0f} s, oracle/encoder_oracle_torch.py (the reference ATen op sequence, CPU)',
            'host_cpus': len(os.sched_getaffinity(0))}


def cpu_baseline_loss(name, budget_s=8.0):
    """utils/loss.py:174-214 + autograd on the host cores: the oracle's
    functional restatement, forward + backward of the weighted loss, all threads and one thread."""
    from oracle import loss_oracle as lo
    from tests.helpers import synth_loss_inputs
    c = CFG[name]
    B = min(REF_SAMPLES, c['B'])
    inp = synth_loss_inputs(1234 + int(name[3:]), B, c['H'], c['W'], 3.0)

    def step():
        fl = [f.clone().requires_grad_() for f in inp['flows']]
        terms = lo.multiscale_loss(fl, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                                   inp['timestamps'], inp['sample_idx'])
        lo.combined(terms).backward()
    out = {}
    for tag, threads in (('all_threads', len(os.sched_getaffinity(0))), ('one_thread', 1)):
        torch.set_num_threads(threads)
        dt, reps = _timed_passes(step, budget_s)
        out[tag] = {'value': B / dt, 'unit': 'samples/s', 'cores': threads, 'kind': 'port',
                    'sample': f'{reps} fwd+bwd passes over {B} of the {c["B"]} samples of {name} in {budget_s:.0f} s, '
                              'oracle/loss_oracle.py (utils/loss.py restated, torch CPU + autograd)'}
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    return out


def bench_map_events(dev, budget_s=4.0):
    """The one native component (transformation.cpp:35-85): the rotation remap of a DAVIS event array.
    GPU: RotationLUT.apply on device-resident events (LUT built once per angle, as RandomRotation
    does per sample); CPU: the plain C++ restatement oracle/map_events_ref.cpp, single thread like
    the original (which builds its hash map per call, so the build is inside its time)."""
    import dvs_of_training_framework_b200 as dv
    from oracle import encoder_oracle as eo
    from oracle import map_events_native as mn
    H, W, n = 260, 346, 1 << 20
    rng = np.random.default_rng(7)
    ev = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), np.sort(rng.random(n)),
                   rng.choice([-1, 1], n), np.zeros(n)], 1).astype(np.float32)
    src, dst = eo.rotation_lut((H, W), 17.5)
    mn.build()
    dt_cpu, reps = _timed_passes(lambda: mn.port_map(ev, (H, W), src, dst), budget_s)
    ev_d = torch.from_numpy(ev).to(dev)
    lut = dv.transformation.RotationLUT((H, W), torch.from_numpy(src.view(np.int64)).to(dev),
                                        torch.from_numpy(dst.view(np.int64)).to(dev), device=dev)
    out = lut.apply(ev_d)
    want = mn.port_map(ev, (H, W), src, dst)
    assert out.shape == want.shape and np.array_equal(out.cpu().numpy(), want), 'map_events differs from the oracle'
    ms = float(np.mean(time_steps(lambda: lut.apply(ev_d), 10, 3, lambda: None)))
    return {'workload': f'{n} DAVIS events (260x346) through the 17.5 degree rotation LUT ({src.size} entries)',
            'gpu_mevents_per_s': n / (ms * 1e-3) / 1e6, 'gpu_ms': ms,
            'note': 'transformation.RotationLUT.apply: count + scan + write, one host sync for the output size',
            'cpu_baseline': {'value': n / dt_cpu / 1e6, 'unit': 'Mevents/s', 'cores': 1, 'kind': 'port',
                             'sample': f'{reps} calls in {budget_s:.0f} s, oracle/map_events_ref.cpp (transformation.cpp:35-85 restated), single thread'}}


def run_reference(args, rank):
    if rank != 0:
        return
    fn, N, B = cpu_encoder_pass(HEADLINE, samples=REF_SAMPLES)
    for _ in range(args.warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = (time.perf_counter() - t0) / args.steps
    v = N / dt / 1e6
    sample = (f'each step = {B} of the {CFG[HEADLINE]["B"]} samples of {HEADLINE} ({N} events, same events per sample '
              'and image size); the reference is Python + torch_scatter and cannot travel: oracle torch-CPU port')
    line = {'impl': 'reference', 'metric': 'encoded Mevents/s', 'value': v, 'unit': 'Mevents/s',
            'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
            'data': 'synthetic',
            'config': {'workload': f'{HEADLINE}: {DESCR[HEADLINE]}'},
            'cpu_baseline': {'value': v, 'unit': 'Mevents/s', 'cores': torch.get_num_threads(),
                             'kind': 'port', 'sample': sample},
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

    def reduce_enc(r):
        r['step_ms'] = max_over_ranks(r['step_ms'])
        r['warm_graph_ms'] = max_over_ranks(r['warm_graph_ms'])
        r['kernels'] = {k: max_over_ranks(v) for k, v in r['kernels'].items()}
        return r

    def reduce_loss(r):
        for k in ('step_ms', 'abi_ms', 'fused_ms', 'warm_graph_ms'):
            r[k] = max_over_ranks(r[k])
        return r

    sync_all()
    with Clocks(local) as clk:
        head = reduce_enc(bench_encoder(HEADLINE, K, W, dev, seed_off=rank))
        torch.cuda.empty_cache()
        sync_all()
        loss = reduce_loss(bench_loss(HEADLINE, K, W, dev))
        sync_all()
    clocks = clk.summary()
    e2e_k = max(20, K)
    e2e = bench_e2e(HEADLINE, e2e_k, 3, dev)
    e2e_s = max_over_ranks(e2e['step_s'])
    torch.cuda.empty_cache()
    e2e_c = bench_e2e(HEADLINE, e2e_k, 3, dev, compact=True)
    e2e_c_s = max_over_ranks(e2e_c['step_s'])
    torch.cuda.empty_cache()
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
        k2 = max(3, min(K, 10))
        for name, kw in (('cfg1', {}), ('cfg2', {}), ('cfg2_contended', {'contended': True}), ('cfg4', {})):
            extra[name] = reduce_enc(bench_encoder(name[:4], k2, 3, dev, seed_off=rank, **kw))
            torch.cuda.empty_cache()
        for name in ('cfg1', 'cfg2', 'cfg4'):
            extra[name + '_loss'] = reduce_loss(bench_loss(name, k2, 3, dev))
            torch.cuda.empty_cache()
    compact = {}
    if not args.quick:
        for name in ('cfg2', 'cfg3'):
            r = bench_encoder_compact(name, max(3, min(K, 10)), 3, dev)
            r['step_ms'] = max_over_ranks(r['step_ms'])
            compact[name] = r
            torch.cuda.empty_cache()
    solo = rank == 0 and not args.quick
    voxel = {}
    if solo and world == 1:                       # extension lines: single-GPU runs only
        for name in ('cfg2', 'cfg3', 'cfg4'):
            voxel[name] = bench_voxel(name, max(3, min(K, 10)), 3, dev)
            torch.cuda.empty_cache()
    codec = bench_codec('cfg2', max(3, min(K, 10)), 3, dev) if solo else None
    mapev = bench_map_events(dev) if solo else None
    torch.cuda.empty_cache()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    frac = lambda nbytes, ms: nbytes / (ms * 1e-3) / 1e9 / peak  # noqa: E731

    def encoder_line(name, r):
        """Workload entry of the encoder: the call, the dominant kernel's roofline, traffic."""
        dom = 'encode_bin_kernel' if r['binned'] else 'encode_vec4_kernel'
        dom_ms = r['kernels']['bin_ms'] if r['binned'] else r['kernels']['scatter_ms']
        # algorithmic bytes of the dominant kernel's launch: the event stream (36 B/event) for the
        # counting-sort pass (the image is written by pass 2), stream + image for the direct scatter
        dom_bytes = 36 * r['events'] if r['binned'] else r['bytes']
        out = {'workload': DESCR[name[:4]] + ('; contended: 80 % of the events on 10 % of the pixels' if r['contended'] else ''),
               'mevents_per_s': world * r['events'] / (r['step_ms'] * 1e-3) / 1e6, 'ms_per_step': r['step_ms'],
               'hbm_frac_step': frac(r['bytes'], r['step_ms']), 'algorithmic_bytes_step': r['bytes'],
               'path': 'binned (counting sort into 6 B records + per-slab reduction in shared memory)' if r['binned']
                       else 'direct (RED.ADD / RED.MAX into the L2-resident image)',
               'warm_graph_ms_per_step': r['warm_graph_ms'],
               'roofline': {'bound': 'hbm', 'kernel': dom, 'kernel_ms': dom_ms, 'algorithmic_bytes': dom_bytes,
                            'achieved': dom_bytes / (dom_ms * 1e-3) / 1e9, 'peak': peak, 'unit': 'GB/s',
                            'frac': frac(dom_bytes, dom_ms), 'traffic': traffic(name[:4], dom)},
               'kernels_ms': r['kernels']}
        if r['binned']:
            out['roofline_pass2'] = {'kernel': 'encode_reduce_kernel', 'kernel_ms': r['kernels']['reduce_ms'],
                                     'algorithmic_bytes': 6 * r['events'] + 4 * r['cells'],
                                     'frac': frac(6 * r['events'] + 4 * r['cells'], r['kernels']['reduce_ms']),
                                     'traffic': traffic(name[:4], 'encode_reduce_kernel'),
                                     'note': 'reads the 6 B/event records pass 1 wrote, writes the image once'}
            out['direct_scatter_ms'] = r['kernels']['scatter_ms']
        else:
            # this design is bound by L2 atomics, not HBM: two REDs per event against the measured
            # ceiling for scattered REDs on this part (profiles/r01b_ubench_atomics.json: 190 G/s)
            out['red_rate_frac_of_measured_ceiling'] = 2 * r['events'] / (dom_ms * 1e-3) / 190e9
        return out

    def loss_line(name, r):
        c_ = CFG[name]
        return {'workload': f'{name}: 4 scales, {c_["H"]}x{c_["W"]}, B={c_["B"]}',
                'samples_per_s': world * r['samples'] / (r['step_ms'] * 1e-3), 'ms_per_step': r['step_ms'],
                'api': 'training.combined_loss(...) + loss.backward() (Python + torch.autograd around the C ABI)',
                'c_abi': {'samples_per_s': world * r['samples'] / (r['abi_ms'] * 1e-3), 'ms_per_step': r['abi_ms'],
                          'hbm_frac_step': frac(r['bytes'], r['abi_ms']),
                          'warm_graph_ms_per_step': r['warm_graph_ms']},
                'python_over_c_abi': r['step_ms'] / r['abi_ms'],
                'roofline': {'bound': 'hbm', 'kernel': 'warp_loss_ring_kernel<fused>', 'kernel_ms': r['fused_ms'],
                             'algorithmic_bytes': r['fused_bytes'], 'achieved': r['fused_bytes'] / (r['fused_ms'] * 1e-3) / 1e9,
                             'peak': peak, 'unit': 'GB/s', 'frac': frac(r['fused_bytes'], r['fused_ms']),
                             'traffic': traffic(name, 'warp_loss_ring_kernel'),
                             'note': '24 B/pixel moved by the fused sweep (16 read + 8 written); timed with its reduce kernel'},
                'kernels': fused_kernels(r, peak)}

    c = CFG[HEADLINE]
    hl = encoder_line(HEADLINE, head)
    ll = loss_line(HEADLINE, loss)
    line = {
        'metric': 'encoded Mevents/s', 'value': hl['mevents_per_s'], 'unit': 'Mevents/s', 'n_gpus': world,
        'steps': K, 'warmup': W, 'ms_per_step': head['step_ms'], 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': f'{HEADLINE}: {DESCR[HEADLINE]}; int64 SoA events (36 B/event) resident in HBM, per GPU',
                   'l2': '256 MiB memset between steps (outside the event pairs); the input (5.8 GB) exceeds the 126 MB L2 anyway',
                   'validation': 'deferred (status word in mapped pinned memory, no host sync per call)',
                   'parallelism': f'replicas x{world} (sharded by sample, no data-path collective)',
                   'path': hl['path']},
        'clocks': clocks,
        'roofline': dict(hl['roofline'], peak_source=peak_src, step_frac=hl['hbm_frac_step'],
                         note='dominant kernel of the step; step_frac = SURVEY 8d bytes of the whole call '
                              '(36 B/event + 16 B/cell) over the whole call'),
        'roofline_pass2': hl.get('roofline_pass2'),
        'warm_graph_ms_per_step': head['warm_graph_ms'],
        'e2e': {'value': world * e2e['events'] / e2e_s / 1e6, 'unit': 'Mevents/s', 'steps': e2e_k,
                'h2d_bytes_per_step': e2e['h2d'], 'd2h_bytes_per_step': e2e['d2h'],
                'api': 'dv.quantize(events, timestamps, sample_idx, imsize) from pinned host tensors; '
                       'two streams, step k D2H overlaps step k+1 H2D',
                'compact_storage_form': {'value': world * e2e_c['events'] / e2e_c_s / 1e6, 'unit': 'Mevents/s',
                                         'h2d_bytes_per_step': e2e_c['h2d'], 'd2h_bytes_per_step': e2e_c['d2h'],
                                         'api': 'dv.compute_event_image_compact (int16/bool/fp32 storage form, 9 B/event)'}},
        'gpu_launches': K * head['launches_per_step'],
        'warp_loss': dict(ll, metric='warp-loss fwd+bwd samples/s', value=ll['samples_per_s'], unit='samples/s',
                          note='value = the call train_flownet.py makes (Python API); c_abi = the same kernels '
                               'driven from the C ABI alone: one prelude launch (timestamp match + chained pyramid + '
                               'border count), the fused terms-and-gradient sweep (TMA-staged), its reduce kernel, '
                               'and dvsof_scale_gradients, PDL-chained; bytes = SURVEY 8d (separate fwd + bwd + pyramid)'),
        'workloads': {},
    }
    if train is not None:
        line['train_step'] = {'metric': 'train-step samples/s', 'value': world * train['samples'] / (train['encoded_step_ms'] * 1e-3),
                              'unit': 'samples/s', 'ms_per_step': train['encoded_step_ms'],
                              'workload': f'EV-FlowNet ({train["params"]} params, fp32/TF32 cuDNN convolutions; upsample+reflect-pad and ReLU+BatchNorm through libdvsof), 256x256, 131072 events/sample, '
                                          f'{train["samples"]} samples/GPU (global {world * train["samples"]}), AdamW(amsgrad, fused); every batch comes from '
                                          'dv.collate_wrapper(samples, encoded=True) in pinned memory (9 B/event, uint8 frames), is uploaded one step ahead on a copy '
                                          'stream and widened on the device (training.DevicePrefetcher); flat-bucket NCCL allreduce overlapped with backward',
                              'dp_check': train.get('dp_check'),
                              'int64_feed': {'value': world * train['samples'] / (train['step_ms'] * 1e-3), 'unit': 'samples/s',
                                             'ms_per_step': train['step_ms'],
                                             'note': 'the same step fed with dv.collate_wrapper(samples) in the reference layout (int64 events, 36 B/event, float32 frames): four times the upload'}}
    for name, r in extra.items():
        line['workloads'][name] = loss_line(name[:4], r) if name.endswith('_loss') else encoder_line(name, r)
    for name, r in compact.items():
        line['workloads'][name + '_compact'] = {
            'workload': DESCR[name] + '; storage form (int16 x, y / bool polarity / fp32 t = 9 B/event, per-sample prefix)',
            'mevents_per_s': world * r['events'] / (r['step_ms'] * 1e-3) / 1e6, 'ms_per_step': r['step_ms'],
            'warm_graph_ms_per_step': r['warm_graph_ms'], 'algorithmic_bytes_step': r['bytes'],
            'hbm_frac_step': frac(r['bytes'], r['step_ms']),
            'path': 'binned' if r['binned'] else 'direct',
            'note': 'dvsof_encode_event_image_compact; its own roofline (9 N + 8 (B + 1) + 16 B H W bytes), '
                    'never mixed with the 36 B/event form'}
    for name, r in voxel.items():
        line['workloads'][f"{name}_voxel_{r['bins']}bins"] = {
            'workload': DESCR[name].split(' -> ')[0] + f" -> temporal-bilinear voxel grid, {r['bins']} bins",
            'extension': 'not in the reference (its encoder is the 4-channel event image); parity unpinned, '
                         'checked against oracle/voxel_oracle.py (published definition, Zhu et al. CVPR 2019)',
            'mevents_per_s': r['events'] / (r['step_ms'] * 1e-3) / 1e6, 'ms_per_step': r['step_ms'],
            'warm_graph_ms_per_step': r['warm_graph_ms'], 'algorithmic_bytes_step': r['bytes'],
            'hbm_frac_step': frac(r['bytes'], r['step_ms']), 'path': 'binned' if r['binned'] else 'direct',
            'cpu_baseline': r['cpu'], 'note': 'dvsof_encode_voxel_grid, one GPU (rank 0); 36 N + 8 B + 4 C B H W bytes'}
    if codec is not None:
        line['workloads']['codec_cfg2'] = {
            'decode_mevents_per_s': codec['events'] / (codec['decode_ms'] * 1e-3) / 1e6, 'decode_ms': codec['decode_ms'],
            'decode_hbm_frac': frac(codec['bytes'], codec['decode_ms']),
            'encode_mevents_per_s': codec['events'] / (codec['encode_ms'] * 1e-3) / 1e6, 'encode_ms': codec['encode_ms'],
            'encode_hbm_frac': frac(codec['bytes'], codec['encode_ms']),
            'note': 'dvsof_decode_events / dvsof_encode_events (utils/dataset.py decode_batch / encode_batch, event part), 45 B/event, one GPU (rank 0)'}
    if mapev is not None:
        line['workloads']['map_events'] = mapev
    if world == 1 and not args.quick:
        line['cpu_baseline'] = cpu_baseline(HEADLINE)
        line['warp_loss']['cpu_baseline'] = cpu_baseline_loss(HEADLINE)
    else:
        line['cpu_baseline'] = None
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
