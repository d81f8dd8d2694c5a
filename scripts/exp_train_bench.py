"""Experiment harness (GPU box): bench.bench_train_step after other bench stages (interference check)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402

dv._lib.Status.mode = 'deferred'
dev = torch.device('cuda', 0)
for stage in sys.argv[1:] or ['train', 'enc', 'train', 'loss', 'train', 'e2e', 'train']:
    if stage == 'train':
        r = bench.bench_train_step(10, 3, dev, 1)
        print('train step_ms', round(r['step_ms'], 2), flush=True)
    elif stage == 'enc':
        bench.bench_encoder('cfg2', 20, 3, dev)
    elif stage == 'loss':
        bench.bench_loss('cfg2', 20, 3, dev)
    elif stage == 'e2e':
        bench.bench_e2e('cfg2', 5, 3, dev)
        bench.bench_e2e('cfg2', 5, 3, dev, compact=True)
    print('after', stage, 'reserved MB', torch.cuda.memory_reserved() >> 20, flush=True)
    torch.cuda.empty_cache()
