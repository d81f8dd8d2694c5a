"""Experiment harness (GPU box): scatter-kernel time against the number of events at the cfg2 image
size (B=8, 256x256): fixed cost a and slope b of t = a + b * N."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402

dv._lib.Status.mode = 'deferred'
dev = torch.device('cuda', 0)
res = {}
for n in [int(a) for a in sys.argv[1:]] or [8192, 32768, 65536, 131072, 262144, 524288, 1048576]:
    bench.CFG['cfg99'] = dict(B=8, H=256, W=256, n=n, T=0.032)
    r = bench.bench_encoder('cfg99', 30, 5, dev)
    res[8 * n] = {'kernel_us': round(r['kernel_ms'] * 1e3, 2), 'step_us': round(r['step_ms'] * 1e3, 2)}
    torch.cuda.empty_cache()
print(json.dumps(res))
