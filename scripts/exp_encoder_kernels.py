"""Kernel-level encoder timings (bench.bench_encoder) for the named workloads; one JSON line."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402

dev = torch.device('cuda', 0)
dv._lib.Status.mode = 'deferred'
out = {}
for name in sys.argv[1:] or ['cfg3', 'cfg4']:
    r = bench.bench_encoder(name, 10, 3, dev)
    out[name] = {'step_us': round(r['step_ms'] * 1e3, 1), **{k: round(v * 1e3, 1) for k, v in r['kernels'].items()},
                 'step_frac': round(r['bytes'] / r['step_ms'] / 1e6 / bench.peaks()[0], 3)}
    torch.cuda.empty_cache()
print(json.dumps(out))
