"""Golden vectors for the preprocessed-dataset index arithmetic, produced by the REFERENCE's own
utils/dataset.py: select_encoded_ranges (:67-115), select_quantized_ranges (:118-157) and
join_batches (:160-198) on seeded random run lengths and on the fixture of
tests/dataset/test_encoding.py:57-67.

    python tests/golden/make_golden_ranges.py        -> tests/golden/ranges.json (committed)"""
import json
import sys
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).parent
sys.path.insert(0, str(OUT))
from make_golden_codec import load_dataset_module  # noqa: E402


def main():
    ds = load_dataset_module()
    rng = np.random.default_rng(5)
    cases = [{'events_per_element': [1, 2, 1, 1, 0, 1, 1], 'elements_per_sample': [2, 1, 4]}]   # the fixture
    for _ in range(6):
        eps = rng.integers(1, 5, rng.integers(2, 9)).tolist()
        cases.append({'events_per_element': rng.integers(0, 50, sum(eps)).tolist(), 'elements_per_sample': eps})
    out = []
    for c in cases:
        epe = torch.tensor(c['events_per_element'], dtype=torch.long)
        eps = torch.tensor(c['elements_per_sample'], dtype=torch.uint8)
        cps = torch.tensor([4 * e for e in c['elements_per_sample']], dtype=torch.long)
        B = eps.numel()
        sel = []
        for b in range(B):
            for e in range(b + 1, B + 1):
                sel.append({'begin': b, 'end': e, 'encoded': ds.select_encoded_ranges(epe, eps, b, e),
                            'quantized': ds.select_quantized_ranges(cps, eps, b, e)})
        out.append(dict(c, channels_per_sample=cps.tolist(), selections=sel))
    (OUT / 'ranges.json').write_text(json.dumps(out))
    print(len(out), 'cases', sum(len(c['selections']) for c in out), 'selections')


if __name__ == '__main__':
    main()
