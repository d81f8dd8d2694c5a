"""Evidence for "the gradient allreduce hides behind cuDNN's backward" (DESIGN.md section 6): the
data-parallel EV-FlowNet train step of bench.py under torch.profiler on rank 0, N >= 2 ranks under
torchrun.  From the kernel timeline: every NCCL allreduce kernel interval, how much of it overlaps
other (compute) kernels of the same step, and how long the step's compute stream is idle waiting for
the last bucket.  Prints one JSON object (rank 0)."""
import itertools
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402
from dvs_of_training_framework_b200 import training as tr  # noqa: E402
from dvs_of_training_framework_b200.evflownet import Model  # noqa: E402


def main():
    rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    dv._lib.Status.mode = 'deferred'
    c = bench.CFG['cfg2']
    torch.manual_seed(0)
    model = Model(dev).to(memory_format=torch.channels_last)
    red = tr.GradientAllreducer(model.parameters())
    red.broadcast_parameters(model)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, amsgrad=True, fused=True)
    evaluator = dv.init_losses((c['H'], c['W']), 32, model, dev, 1)
    batch = dv.collate_wrapper(bench.make_train_samples(32, c['H'], c['W'], c['n'], c['T'], 7 + rank), encoded=True)
    feed = iter(tr.DevicePrefetcher(itertools.repeat(batch, 40), dev, True))
    timers = dv.timer.FakeTimer()
    model.train()

    def step():
        loss, _, _ = tr.process_minibatch(model, next(feed), timers, dev, True, evaluator, [0.5, 1, 1])
        loss.backward()
        red.wait()
        opt.step()
        red.zero_grad()
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    import time
    t0 = time.perf_counter()
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    step_ms = (time.perf_counter() - t0) * 100
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(4):
            step()
        torch.cuda.synchronize()
    if rank == 0:
        ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time_total > 0]
        nccl = [(e.time_range.start, e.time_range.end) for e in ev if 'nccl' in e.name.lower()]
        comp = sorted((e.time_range.start, e.time_range.end) for e in ev if 'nccl' not in e.name.lower() and 'memcpy' not in e.name.lower())
        merged = []
        for a, b in comp:
            if merged and a <= merged[-1][1]:
                merged[-1][1] = max(merged[-1][1], b)
            else:
                merged.append([a, b])

        def covered(a, b):
            return sum(max(0, min(b, y) - max(a, x)) for x, y in merged)
        tot = sum(b - a for a, b in nccl)
        ov = sum(covered(a, b) for a, b in nccl)
        span = (max(b for _, b in comp) - min(a for a, _ in comp)) if comp else 0
        out = {'world': world, 'step_ms_unprofiled': step_ms, 'steps_profiled': 4, 'nccl_kernels': len(nccl), 'nccl_kernel_us_total': tot,
               'nccl_us_overlapped_by_compute_kernels': ov, 'overlap_fraction': ov / tot if tot else None,
               'compute_span_us': span, 'compute_busy_us': sum(b - a for a, b in merged),
               'top_kernels_us': sorted(((e.key[:60], round(e.device_time_total)) for e in prof.key_averages()
                                        if e.device_time_total > 0), key=lambda kv: -kv[1])[:12]}
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
