"""Experiment harness (GPU box): the fused sweep alone (`dvsof_warp_loss_fused`, ready-made levels,
counts given) and the whole C-ABI training pass, per config; one JSON line.  L2 flushed between
steps like bench.py."""
import ctypes
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402
from dvs_of_training_framework_b200.loss import _fill_frames, _scale_struct, interpolate  # noqa: E402
from tests.helpers import synth_loss_inputs  # noqa: E402


def run(name, K, dev):
    L = dv._lib
    c = bench.CFG[name]
    B, H, W = c['B'], c['H'], c['W']
    inp = synth_loss_inputs(1234 + int(name[3:]), B, H, W, 3.0, device=dev)
    fl = [f.contiguous() for f in inp['flows']]
    levels, cur = [], inp['images'].view(2 * B, H, W)
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
    coef = (ctypes.c_float * (3 * S))(*([0.5 / S] * S + [1.0 / S] * S + [1.0 / S] * S))
    loss1 = torch.empty(1, device=dev)
    sp = L.stream_ptr(dev)
    fus = lambda cp: L.call('dvsof_warp_loss_fused', sc, L.ptr(match), coef, L.ptr(terms), L.ptr(loss1),  # noqa: E731
                            L.ptr(n_bad), L.ptr(ws), cp, sp)
    fus(1)
    flush = bench.L2Flusher(dev)
    fused = float(np.mean(bench.time_steps(lambda: fus(0), K, 3, flush)))
    ev = dv.Losses(inp['shapes'], B, dev)
    plan = ev._plan(dev, B, 2 * B, H, W, fl)
    o = plan.offs
    buf = torch.empty(plan.total, dtype=torch.float32, device=dev)
    gbuf = torch.empty(plan.gtotal, dtype=torch.float32, device=dev)
    fr = _fill_frames(plan, inp['images'], inp['flow_ts'], inp['flow_sample_idx'], inp['timestamps'], inp['sample_idx'])
    plan.fill(plan.bwd, buf.data_ptr(), fl, gbuf.data_ptr())
    st = L.status_for(dev)
    one = torch.ones(1, device=dev)
    base = buf.data_ptr()

    def abi():
        L.call('dvsof_multiscale_loss_fused', plan.bwd, fr, coef, base + 4 * o[S], base + 4 * o[S + 2],
               base + 4 * o[S + 3], base + 4 * o[S + 1], plan.ws_ptr, L.ptr(st.word), sp)
        L.call('dvsof_scale_gradients', plan.bwd, L.ptr(one), sp)
    abi_ms = float(np.mean(bench.time_steps(abi, K, 3, flush)))
    px = sum(B * f.shape[-2] * f.shape[-1] for f in fl)
    peak, _ = bench.peaks()
    return {'fused_us': round(fused * 1e3, 1), 'fused_frac24': round(24 * px / fused / 1e6 / peak, 3),
            'abi_us': round(abi_ms * 1e3, 1), 'abi_frac': round(bench.loss_bytes(B, H, W) / abi_ms / 1e6 / peak, 3)}


if __name__ == '__main__':
    dev = torch.device('cuda', 0)
    dv._lib.Status.mode = 'deferred'
    K = int(os.environ.get('DVSOF_EXP_STEPS', 20))
    print(json.dumps({n: run(n, K, dev) for n in (sys.argv[1:] or ['cfg2', 'cfg3', 'cfg4'])}))
