"""CPU restatement (numpy) of the reference's evaluation metric -- TEST INFRASTRUCTURE ONLY
(imported by tests/ and nothing else; the product path is csrc/eval.cu).

flow_error_dense follows utils/eval.py:6-50 step by step.  Pinned against the reference's own
function run in the build container: tests/golden/eval.npz (tests/golden/make_golden_eval.py)."""
import numpy as np


def flow_error_dense(flow_gt, flow_pred, event_img, is_car=False, is_dense=False):
    rows = 190 if is_car else flow_gt.shape[1]          # eval.py:16-19 (shape[1], as written there)
    ev = np.squeeze(event_img)[:rows]                    # eval.py:21
    gt, pred = flow_gt[:rows], flow_pred[:rows]          # eval.py:22-24
    keep = np.ones(ev.shape, dtype=bool) if is_dense else ev > 0                   # eval.py:26-29
    finite = ~np.isinf(gt[..., 0]) & ~np.isinf(gt[..., 1])                         # eval.py:32-34
    moving = np.sqrt(gt[..., 0] * gt[..., 0] + gt[..., 1] * gt[..., 1]) > 0        # eval.py:35
    keep = np.squeeze(keep & finite & moving)                                      # eval.py:36
    d = gt[keep] - pred[keep]                                                      # eval.py:38-41
    ee = np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1])
    n = ee.shape[0]
    aee = float(np.mean(ee)) if n else float('nan')
    return aee, float((ee < 3.).sum()) / float(n + 1e-5), n                        # eval.py:43-50
