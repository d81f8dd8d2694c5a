"""Experiment harness (GPU box): EV-FlowNet train-step time for memory-format / cuDNN settings."""
import itertools
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402


def run(fmt, benchmark, per_gpu=32, steps=8, fused_opt=False, no_bias=False):
    import dvs_of_training_framework_b200 as dv
    from dvs_of_training_framework_b200 import training as tr
    from dvs_of_training_framework_b200.evflownet import Model
    dev = torch.device('cuda', 0)
    torch.backends.cudnn.benchmark = benchmark
    c = bench.CFG['cfg2']
    torch.manual_seed(0)
    model = Model(dev)
    if fmt.startswith('nhwc'):
        model = model.to(memory_format=torch.channels_last)
    if fmt == 'nhwc_in':
        q = model.quantize
        model.quantize = lambda *a: q(*a).contiguous(memory_format=torch.channels_last)
    if no_bias:          # upper bound of folding the conv bias into the ReLU+BN kernels
        for m in model.modules():
            if m.__class__.__name__ == '_ConvReluBN':
                m.conv.bias = None
    red = tr.GradientAllreducer(model.parameters())
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5, amsgrad=True, fused=fused_opt)
    evaluator = dv.init_losses((c['H'], c['W']), per_gpu, model, dev, 1)
    batch = bench.make_train_batch(per_gpu, c['H'], c['W'], c['n'], c['T'], 1239)
    timers = dv.timer.FakeTimer()
    model.train()
    feed = iter(tr.DevicePrefetcher(itertools.repeat(batch, steps + 8), dev, True))

    def step():
        loss, _, _ = tr.process_minibatch(model, next(feed), timers, dev, True, evaluator, [0.5, 1, 1])
        loss.backward()
        red.wait()
        opt.step()
        red.zero_grad()
        return loss
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, float(loss)


def main():
    import dvs_of_training_framework_b200 as dv
    dv._lib.Status.mode = 'deferred'
    res = {}
    for fmt, fo in (('nhwc', False), ('nhwc', True), ('nhwc', False), ('nhwc', True)):
        ms, loss = run(fmt, False, fused_opt=fo)
        res[f'{fmt}{"+fused_adamw" if fo else ""}_{len(res)}'] = {'step_ms': round(ms, 2), 'loss': loss}
        torch.cuda.empty_cache()
    print(json.dumps(res))


if __name__ == '__main__':
    main()
