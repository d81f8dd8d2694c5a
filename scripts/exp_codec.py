"""Experiment harness (GPU box): batch codec throughput on the named configs."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402


def main():
    dev = torch.device('cuda', 0)
    peak, _ = bench.peaks()
    res = {}
    for name in sys.argv[1:] or ['cfg2', 'cfg4', 'cfg3']:
        r = bench.bench_codec(name, 10, 3, dev)
        res[name] = {'events': r['events'], 'decode_ms': round(r['decode_ms'], 4), 'encode_ms': round(r['encode_ms'], 4),
                     'decode_GBs': round(r['bytes'] / r['decode_ms'] / 1e6, 1), 'encode_GBs': round(r['bytes'] / r['encode_ms'] / 1e6, 1),
                     'decode_frac': round(r['bytes'] / r['decode_ms'] / 1e6 / peak, 4),
                     'encode_frac': round(r['bytes'] / r['encode_ms'] / 1e6 / peak, 4)}
        torch.cuda.empty_cache()
    print(json.dumps(res))


if __name__ == '__main__':
    main()
