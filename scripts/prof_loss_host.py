"""Host-side profile of one fused-loss training step (cProfile), GPU box."""
import cProfile
import pstats
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import dvs_of_training_framework_b200 as dv  # noqa: E402
from dvs_of_training_framework_b200 import training as tr  # noqa: E402
from tests.helpers import synth_loss_inputs  # noqa: E402

dv._lib.Status.mode = 'deferred'
dev = torch.device('cuda', 0)
inp = synth_loss_inputs(1, 8, 256, 256, 3.0, device=dev)
flows = [f.requires_grad_() for f in inp['flows']]
ev = dv.Losses(inp['shapes'], 8, dev)


def step():
    for f in flows:
        f.grad = None
    loss, _ = tr.combined_loss(ev, flows, inp['flow_ts'], inp['flow_sample_idx'], inp['images'],
                               inp['timestamps'], inp['sample_idx'], None, weights=[0.5, 1, 1])
    loss.backward()


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f'host enqueue {1e6 * (t1 - t0) / 200:.1f} us/step, with final sync {1e6 * (t2 - t0) / 200:.1f} us/step')
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(18)
