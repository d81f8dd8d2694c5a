"""Golden vectors for the evaluation metric: the REFERENCE's own utils/eval.py:flow_error_dense
run in the build container (needs /root/reference and cv2).  Output (committed):
tests/golden/eval.npz.   python tests/golden/make_golden_eval.py"""
import importlib.util
from pathlib import Path

import numpy as np

REF = Path('/root/reference')
OUT = Path(__file__).parent


def main():
    spec = importlib.util.spec_from_file_location('ref_eval', REF / 'utils' / 'eval.py')
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    rng = np.random.default_rng(20260)
    out = {}
    cases = [('tall_f32', 200, 96, np.float32, np.float32, False, False),
             ('tall_car', 200, 96, np.float32, np.float32, True, False),     # 190-row crop
             ('gt64', 120, 70, np.float64, np.float32, False, False),
             ('dense_f64', 48, 56, np.float64, np.float64, False, True),
             ('small_empty', 9, 12, np.float32, np.float32, False, False)]
    for name, H, W, tg, tp, car, dense in cases:
        gt = (rng.standard_normal((H, W, 2)) * 4).astype(tg)
        gt[rng.random((H, W)) < 0.2] = 0                       # no GT motion
        gt[rng.random((H, W)) < 0.05, 0] = np.inf              # invalid GT
        gt[rng.random((H, W)) < 0.05, 1] = -np.inf
        pred = (gt + rng.standard_normal((H, W, 2)) * 2).astype(tp)
        pred[~np.isfinite(pred)] = 0
        cnt = rng.poisson(0.7, (H, W)).astype(np.uint64)
        if name == 'small_empty':
            cnt[:] = 0
        with np.errstate(all='ignore'):
            aee, pct, n = ref.flow_error_dense(gt, pred, cnt, is_car=car, is_dense=dense)
        out.update({f'{name}_gt': gt, f'{name}_pred': pred, f'{name}_cnt': cnt,
                    f'{name}_flags': np.array([car, dense]),
                    f'{name}_res': np.array([aee, pct, n], dtype=np.float64)})
        print(name, aee, pct, n)
    np.savez_compressed(OUT / 'eval.npz', **out)


if __name__ == '__main__':
    main()
