#!/usr/bin/env python
"""Opcode histogram per kernel of a built library (cuobjdump -sass), for profiles/sass_opcodes_*.txt.

    python scripts/sass_histogram.py [lib.so] [--kernel REGEX] [--out DIR]

Per kernel: instruction count, then counts per opcode root (FFMA2, MUFU, RED, ATOMS, UBLKCP,
UTMALDG, SYNCS, LDG, STG, ...) -- the evidence the profiling recipe asks for (TMA shows as
UTMALDG / UBLKCP, packed fp32x2 as FFMA2 / FMUL2 / FADD2, PDL as ACQBULK).
"""
import argparse
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def kernels(lib):
    out = subprocess.run(['cuobjdump', '-sass', str(lib)], capture_output=True, text=True, check=True).stdout
    name, rows = None, {}
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            name = m.group(1)
            rows[name] = []
            continue
        m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(.*?);', line)
        if m and name:
            ins = re.sub(r'^@!?U?P\w+\s+', '', m.group(1).strip())
            rows[name].append(ins.split()[0])
    return rows


def demangle(n):
    try:
        return subprocess.run(['c++filt', n], capture_output=True, text=True).stdout.strip() or n
    except OSError:
        return n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('lib', nargs='?', default=str(ROOT / 'dvs_of_training_framework_b200' / 'libdvsof.so'))
    ap.add_argument('--kernel', default='.')
    ap.add_argument('--out')
    a = ap.parse_args()
    for name, ops in kernels(a.lib).items():
        d = demangle(name)
        if not re.search(a.kernel, d):
            continue
        full = collections.Counter(ops)
        root = collections.Counter(o.split('.')[0] for o in ops)
        text = [f'# {d}', f'# instructions: {len(ops)}', '# opcode roots:']
        text += [f'{c:6d}  {o}' for o, c in root.most_common()]
        text += ['# full opcodes of interest:']
        text += [f'{c:6d}  {o}' for o, c in sorted(full.items())
                 if re.match(r'(RED|REDG|ATOM|ATOMS|ATOMG|UBLKCP|UTMA|SYNCS|ACQBULK|MUFU|F(FMA|MUL|ADD)2|LDG|STG|LDS|STS|BAR|CCTL|MEMBAR)', o)]
        body = '\n'.join(text) + '\n'
        if a.out:
            short = re.sub(r'[^A-Za-z0-9_]+', '_', re.sub(r'\(.*', '', d).replace('dvsof::', '').replace('void ', ''))[:80]
            Path(a.out).mkdir(parents=True, exist_ok=True)
            (Path(a.out) / f'sass_opcodes_{short}.txt').write_text(body)
        else:
            sys.stdout.write(body + '\n')


if __name__ == '__main__':
    main()
