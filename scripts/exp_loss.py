"""Experiment harness (GPU box): time the fused warp-loss step and its forward / backward
kernels on the named configs (default cfg2 cfg3 cfg4); used under ncu for the loss profiles."""
import json
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402


def main():
    dev = torch.device('cuda', 0)
    dv._lib.Status.mode = 'deferred'
    peak, _ = bench.peaks()
    res = {}
    for name in sys.argv[1:] or ['cfg2', 'cfg3', 'cfg4']:
        r = bench.bench_loss(name, int(os.environ.get('DVSOF_EXP_STEPS', 20)), 3, dev)
        res[name] = {'step_ms': round(r['step_ms'], 4), 'fwd_ms': round(r['fwd_ms'], 4),
                     'bwd_ms': round(r['bwd_ms'], 4),
                     'fwd_GBs': round(r['fwd_bytes'] / r['fwd_ms'] / 1e6, 1),
                     'bwd_GBs': round(r['bwd_bytes'] / r['bwd_ms'] / 1e6, 1),
                     'fwd_frac': round(r['fwd_bytes'] / r['fwd_ms'] / 1e6 / peak, 4),
                     'bwd_frac': round(r['bwd_bytes'] / r['bwd_ms'] / 1e6 / peak, 4),
                     'abi_ms': round(r['abi_ms'], 4), 'fused_ms': round(r['fused_ms'], 4),
                     'fused_with_count_ms': round(r['fused_with_count_ms'], 4),
                     'fused_GBs': round(r['fused_bytes'] / r['fused_ms'] / 1e6, 1),
                     'fused_frac': round(r['fused_bytes'] / r['fused_ms'] / 1e6 / peak, 4)}
        torch.cuda.empty_cache()
    print(json.dumps(res))


if __name__ == '__main__':
    main()
