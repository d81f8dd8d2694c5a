#!/usr/bin/env python
"""profiles/traffic.json from `ncu --set full` captures: per workload and kernel the DRAM bytes of
one launch (dram__bytes_read.sum + dram__bytes_write.sum) -- what bench.py reports as
`roofline.traffic` next to the algorithmic bytes.

    python scripts/ncu_traffic.py cfg3=gpurun_out/r02x/ncu_cfg3.ncu-rep cfg4=... [--out profiles/traffic.json]

Every argument is workload=report; the report is exported with `ncu -i ... --page raw --csv`.  Kernels
are keyed by their base name (no namespace, no template arguments); of several launches of a kernel
the LAST captured one is kept (the first ones see cold instruction caches).  Existing entries of the
output file for other workloads are kept."""
import csv
import io
import json
import re
import subprocess
import sys
from pathlib import Path

UNIT = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12, 'ns': 1e-3, 'us': 1, 'ms': 1e3, 's': 1e6}


def raw_rows(report):
    out = subprocess.run(['ncu', '-i', report, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    args = [a for a in sys.argv[1:] if '=' in a and not a.startswith('--')]
    out_path = Path(sys.argv[sys.argv.index('--out') + 1]) if '--out' in sys.argv else \
        Path(__file__).resolve().parent.parent / 'profiles' / 'traffic.json'
    table = json.loads(out_path.read_text()) if out_path.exists() else {}
    for a in args:
        workload, report = a.split('=', 1)
        hdr, units, rows = raw_rows(report)
        col = {h: i for i, h in enumerate(hdr)}
        entry = table.setdefault(workload, {})
        for r in rows:
            name = re.sub(r'^void\s+', '', r[col['Kernel Name']])
            name = re.sub(r'<.*', '', re.sub(r'\(.*', '', name)).split('::')[-1].strip()

            def val(metric):
                i = col[metric]
                return float(r[i].replace(',', '')) * UNIT.get(units[i], 1)
            rd, wr = val('dram__bytes_read.sum'), val('dram__bytes_write.sum')
            entry[name] = {'dram_bytes': rd + wr, 'dram_bytes_read': rd, 'dram_bytes_write': wr,
                           'duration_us_under_ncu': val('gpu__time_duration.sum'),
                           'grid': r[col['launch__grid_size']], 'source': str(report)}
    out_path.write_text(json.dumps(table, indent=1, sort_keys=True) + '\n')
    for w in table:
        for k, v in table[w].items():
            print(f'{w:6s} {k:28s} {v["dram_bytes"] / 1e6:10.2f} MB  ({v["duration_us_under_ncu"]:.1f} us under ncu)')


if __name__ == '__main__':
    main()
