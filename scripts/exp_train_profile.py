"""Experiment harness (GPU box): where the EV-FlowNet train step spends its time (torch profiler,
top CUDA kernels, GPU busy fraction)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402


def main():
    import itertools
    import dvs_of_training_framework_b200 as dv
    from dvs_of_training_framework_b200 import training as tr
    from dvs_of_training_framework_b200.evflownet import Model
    dev = torch.device('cuda', 0)
    dv._lib.Status.mode = 'deferred'
    c = bench.CFG['cfg2']
    per_gpu = 32
    import os
    model = Model(dev)
    if os.environ.get('DVSOF_MODEL_FORMAT', 'nhwc') == 'nhwc':
        model = model.to(memory_format=torch.channels_last)
    red = tr.GradientAllreducer(model.parameters())
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, amsgrad=True)
    evaluator = dv.init_losses((c['H'], c['W']), per_gpu, model, dev, 1)
    batch = bench.make_train_batch(per_gpu, c['H'], c['W'], c['n'], c['T'], 1239)
    timers = dv.timer.FakeTimer()
    model.train()
    feed = iter(tr.DevicePrefetcher(itertools.repeat(batch, 40), dev, True))

    def step():
        loss, _, _ = tr.process_minibatch(model, next(feed), timers, dev, True, evaluator, [0.5, 1, 1])
        loss.backward()
        red.wait()
        opt.step()
        red.zero_grad()
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record()
    torch.cuda.synchronize()
    print('step_ms', e0.elapsed_time(e1) / 5)
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
        for _ in range(3):
            step()
        torch.cuda.synchronize()
    ka = prof.key_averages()
    tot = sum(k.device_time_total for k in ka if k.device_type == torch.autograd.DeviceType.CUDA)
    print('cuda kernel time per step (ms):', tot / 3 / 1000)
    print(ka.table(sort_by='device_time_total', row_limit=45, max_name_column_width=70))
    by_shape = prof.key_averages(group_by_input_shape=True)
    rows = [k for k in by_shape if k.key in ('aten::copy_', 'aten::contiguous', 'aten::clone', 'aten::constant_pad_nd', 'aten::cat')]
    rows.sort(key=lambda k: -k.device_time_total)
    for k in rows[:25]:
        print('SHAPE', k.key, round(k.device_time_total / 3000, 3), 'ms/step', k.count // 3, 'calls/step', k.input_shapes)


if __name__ == '__main__':
    main()
