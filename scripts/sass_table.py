#!/usr/bin/env python
"""One row per kernel of libdvsof.so: static SASS instruction count and the opcode roots that carry the
evidence the profiling recipe asks for (profiles/*_sass_opcodes.txt).  Usage: sass_table.py [lib.so]"""
import collections
import re
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent))
from sass_histogram import ROOT, demangle, kernels  # noqa: E402

COLS = ['REDG', 'ATOMG', 'ATOMS', 'UTMALDG', 'UBLKCP', 'SYNCS', 'ACQBULK', 'MUFU', 'FFMA2', 'FMUL2', 'FADD2', 'LDG', 'LDS', 'STS',
        'STG', 'ST', 'BAR', 'SHFL', 'DADD']

if __name__ == '__main__':
    lib = sys.argv[1] if len(sys.argv) > 1 else str(ROOT / 'dvs_of_training_framework_b200' / 'libdvsof.so')
    print('# cuobjdump -sass opcode counts per kernel of libdvsof.so (sm_100a), scripts/sass_table.py; static counts.')
    print('# TMA = UTMALDG (+ SYNCS mbarrier ops); PDL wait = ACQBULK; packed fp32x2 = FFMA2/FMUL2/FADD2; L2 atomics = REDG/ATOMG; shared atomics = ATOMS')
    print(f'{"kernel":70s} {"instr":>6s} ' + ' '.join(f'{c:>7s}' for c in COLS))
    rows = {}
    for name, ops in kernels(lib).items():
        d = re.sub(r'\(.*', '', demangle(name)).replace('void ', '').replace('dvsof::', '')
        rows[d] = ops
    for d in sorted(rows, key=lambda k: ('<' in k, k)):
        root = collections.Counter(o.split('.')[0] for o in rows[d])
        print(f'{d[:70]:70s} {len(rows[d]):6d} ' + ' '.join(f'{root.get(c, 0):7d}' for c in COLS))
