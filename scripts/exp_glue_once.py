"""Experiment harness (GPU box, for ncu): two train steps of the channels_last network."""
import itertools
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402
from dvs_of_training_framework_b200 import training as tr  # noqa: E402
from dvs_of_training_framework_b200.evflownet import Model  # noqa: E402

dv._lib.Status.mode = 'deferred'
dev = torch.device('cuda', 0)
c = bench.CFG['cfg2']
model = Model(dev).to(memory_format=torch.channels_last)
red = tr.GradientAllreducer(model.parameters())
opt = torch.optim.AdamW(model.parameters(), lr=1e-5, amsgrad=True, fused=True)
evaluator = dv.init_losses((c['H'], c['W']), 32, model, dev, 1)
batch = bench.make_train_batch(32, c['H'], c['W'], c['n'], c['T'], 1239)
timers = dv.timer.FakeTimer()
model.train()
feed = iter(tr.DevicePrefetcher(itertools.repeat(batch, 4), dev, True))
for _ in range(2):
    loss, _, _ = tr.process_minibatch(model, next(feed), timers, dev, True, evaluator, [0.5, 1, 1])
    loss.backward()
    red.wait()
    opt.step()
    red.zero_grad()
torch.cuda.synchronize()
print('ok', float(loss))
