"""Experiment harness (GPU box): the binned encoder path per pass (prepare + bin + reduce timed alone and as
the whole call) on cfg3 / cfg4, with an exactness check against a torch scatter on the GPU; one JSON
line.  Select a build with DVSOF_LIB, the compiler-scheduled pass 2 with DVSOF_REDUCE_OLD=1."""
import json
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import dvs_of_training_framework_b200 as dv  # noqa: E402
from exp_encoder import torch_ref  # noqa: E402

if __name__ == '__main__':
    dev = torch.device('cuda', 0)
    dv._lib.Status.mode = 'deferred'
    res = {'lib': os.environ.get('DVSOF_LIB', 'default'), 'old_reduce': bool(os.environ.get('DVSOF_REDUCE_OLD'))}
    for name in sys.argv[1:] or ['cfg3', 'cfg4']:
        c = bench.CFG[name]
        contended = bool(os.environ.get('DVSOF_EXP_CONTENDED'))       # 80 % of the events on 10 % of the pixels
        r = bench.bench_encoder(name, 10, 3, dev, contended=contended)
        ev, ts, sidx = bench.fast_events(1234 + int(name[3:]), c['B'], c['H'], c['W'], c['n'], c['T'], dev, contended)
        size = dv.extend_size((c['H'], c['W']))
        ok = bool(torch.equal(dv.compute_event_image(ev, ts, sidx, size), torch_ref(ev, ts, c['B'], *size)))
        res[name] = {'step_us': round(r['step_ms'] * 1e3, 1), 'bin_us': round(r['kernels'].get('bin_ms', 0) * 1e3, 1),
                     'reduce_us': round(r['kernels'].get('reduce_ms', 0) * 1e3, 1),
                     'direct_us': round(r['kernels']['scatter_ms'] * 1e3, 1), 'contended': contended, 'exact': ok}
        del ev
        torch.cuda.empty_cache()
    print(json.dumps(res))
