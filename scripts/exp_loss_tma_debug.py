"""One small multiscale loss call (forward + fused) against the oracle; prints OK or the error.
Used to bisect the staging modes: DVSOF_LOSS_NO_TMA / DVSOF_LOSS_TMA_MASK / DVSOF_LOSS_TMA_MINW."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import dvs_of_training_framework_b200 as dv  # noqa: E402
from oracle import loss_oracle as lo  # noqa: E402
from tests.helpers import synth_loss_inputs  # noqa: E402

B, H, W = (int(v) for v in (sys.argv[1:4] or (2, 64, 96)))
inp = synth_loss_inputs(3, B, H, W, 3.0)
tag = {k: os.environ.get(k) for k in ('DVSOF_LOSS_NO_TMA', 'DVSOF_LOSS_TMA_MASK', 'DVSOF_LOSS_TMA_MINW',
                                      'DVSOF_LOSS_SEG_ROWS') if os.environ.get(k)}
try:
    ev = dv.Losses(inp['shapes'], B, 'cuda')
    flows = [f.cuda().requires_grad_() for f in inp['flows']]
    args = [inp[k].cuda() for k in ('flow_ts', 'flow_sample_idx', 'images', 'timestamps', 'sample_idx')]
    loss, flat = ev.combined(flows, *args)
    loss.backward()
    torch.cuda.synchronize()
    of = [f.clone().requires_grad_() for f in inp['flows']]
    ot = lo.multiscale_loss(of, inp['flow_ts'], inp['flow_sample_idx'], inp['images'], inp['timestamps'],
                            inp['sample_idx'])
    lo.combined(ot).backward()
    ref = np.array([[float(v) for v in t] for t in ot]).reshape(-1)
    err = float(np.abs(flat.cpu().numpy() - ref).max())
    gerr = max(float((a.grad.cpu() - b.grad).abs().max()) for a, b in zip(flows, of))
    print('OK', (B, H, W), tag, 'terms err %.2e grad err %.2e' % (err, gerr))
except Exception as e:  # noqa: BLE001
    print('FAIL', (B, H, W), tag, str(e).splitlines()[0][:100])
