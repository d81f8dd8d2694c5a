"""Per CUDA source line of an `ncu --page source --csv --print-source cuda,sass` export: executed warp
instructions, stall samples and the split of those samples by reason (top lines by samples).
Usage: ncu_source_stalls.py export.csv [top_n] [function-substring]"""
import csv
import sys
from collections import defaultdict

REASONS = ['stall_long_sb', 'stall_wait', 'stall_short_sb', 'stall_mio', 'stall_math', 'stall_not_selected',
           'stall_selected', 'stall_branch_resolving', 'stall_no_inst', 'stall_lg', 'stall_dispatch', 'stall_barrier']


def main():
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    want = sys.argv[3] if len(sys.argv) > 3 else ''
    fname, func, hdr = None, '', None
    inst, samp, text = defaultdict(float), defaultdict(float), {}
    by = defaultdict(lambda: defaultdict(float))
    for r in csv.reader(open(sys.argv[1])):
        if not r:
            continue
        if r[0] in ('File Name', 'File Path'):
            fname = r[1].split('/')[-1]
            continue
        if r[0] == 'Function Name':
            func = r[1]
            continue
        if r[0] == 'Line No':
            hdr = r
            ia, ii, isamp = hdr.index('Address'), hdr.index('Instructions Executed'), hdr.index('# Samples')
            cols = {k: hdr.index(k) for k in REASONS if k in hdr}
            continue
        if hdr is None or len(r) <= ii or want not in func:
            continue
        if r[0].isdigit() and r[ia] == '-':
            key = (fname, int(r[0]))
            text[key] = r[1].strip()
            try:
                inst[key] += float(r[ii])
                samp[key] += float(r[isamp])
                for k, c in cols.items():
                    by[key][k] += float(r[c])
            except ValueError:
                pass
    tot_i, tot_s = sum(inst.values()) or 1, sum(samp.values()) or 1
    tot_by = defaultdict(float)
    for key in by:
        for k, v in by[key].items():
            tot_by[k] += v
    print(f'total warp instructions {tot_i:.0f}, stall samples {tot_s:.0f}')
    print('by reason: ' + ', '.join(f'{k[6:]} {100 * v / tot_s:.1f}%' for k, v in sorted(tot_by.items(), key=lambda kv: -kv[1])))
    for key in sorted(samp, key=lambda k: -samp[k])[:top]:
        parts = ' '.join(f'{k[6:]}={100 * v / tot_s:.1f}' for k, v in sorted(by[key].items(), key=lambda kv: -kv[1])[:3] if v)
        print(f'{key[0]}:{key[1]:<5d} inst {100 * inst[key] / tot_i:5.1f}%  samples {100 * samp[key] / tot_s:5.1f}%  [{parts}]  {text.get(key, "")[:90]}')


if __name__ == '__main__':
    main()
