"""A few whole encoder calls on one workload (for ncu captures: prepare + the path the library picks)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'cfg3'
dev = torch.device('cuda', 0)
dv._lib.Status.mode = 'deferred'
c = bench.CFG[name]
ev, ts, sidx = bench.fast_events(1234 + int(name[3:]), c['B'], c['H'], c['W'], c['n'], c['T'], dev)
flush = bench.L2Flusher(dev)
for _ in range(4):
    flush()
    img = dv.quantize(ev, ts, sidx, (c['H'], c['W']))
torch.cuda.synchronize()
assert int(img[:, :2].sum(dtype=torch.float64)) == c['B'] * c['n']
print('ok', name, tuple(img.shape))
