"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line:
executed warp instructions and stall samples.  Usage: ncu_source_lines.py export.csv [top_n]"""
import csv
import sys
from collections import defaultdict


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    fname, hdr = None, None
    inst, samp, text = defaultdict(float), defaultdict(float), {}
    for r in rows:
        if not r:
            continue
        if r[0] in ('File Name', 'File Path'):
            fname = r[1].split('/')[-1]
            continue
        if r[0] == 'Line No':
            hdr = r
            ia, ii, isamp = hdr.index('Address'), hdr.index('Instructions Executed'), hdr.index('# Samples')
            continue
        if hdr is None or len(r) <= ii:
            continue
        if r[0].isdigit() and r[ia] == '-':          # a source line row carries its SASS rows' totals
            key = (fname, int(r[0]))
            text[key] = r[1].strip()
            try:
                inst[key] += float(r[ii])
                samp[key] += float(r[isamp])
            except ValueError:
                pass
    tot_i, tot_s = sum(inst.values()) or 1, sum(samp.values()) or 1
    print(f'total warp instructions {tot_i:.0f}, stall samples {tot_s:.0f}')
    for key in sorted(inst, key=lambda k: -inst[k])[:top]:
        print(f'{key[0]}:{key[1]:<5d} inst {100 * inst[key] / tot_i:5.1f}%  samples {100 * samp[key] / tot_s:5.1f}%  {text.get(key, "")[:110]}')


if __name__ == '__main__':
    main()
