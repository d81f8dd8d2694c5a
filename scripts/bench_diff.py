"""Print the encoder / loss figures of one or more bench JSON lines side by side (A/B of kernel builds)."""
import json
import sys


def load(p):
    return json.loads([x for x in open(p).read().splitlines() if x.startswith('{')][-1])


def rows(d):
    r = {'cfg3 step ms': d['ms_per_step'], 'cfg3 bin ms': d['roofline']['kernel_ms'],
         'cfg3 reduce ms': d.get('roofline_pass2', {}).get('kernel_ms'), 'cfg3 step frac': d['roofline']['step_frac'],
         'cfg3 bin frac': d['roofline']['frac'], 'e2e Mev/s': d['e2e']['value']}
    for k, v in d.get('workloads', {}).items():
        if 'ms_per_step' in v:
            r[k + ' ms'] = v['ms_per_step']
        for kk, vv in (v.get('kernels_ms') or {}).items():
            r[f'{k} {kk}'] = vv
        if isinstance(v.get('c_abi'), dict):
            r[k + ' c_abi ms'] = v['c_abi']['ms_per_step']
    r['loss cfg3 c_abi ms'] = d['warp_loss']['c_abi']['ms_per_step']
    r['train ms'] = d.get('train_step', {}).get('ms_per_step')
    return r


if __name__ == '__main__':
    cols = [rows(load(p)) for p in sys.argv[1:]]
    for k in cols[0]:
        vals = [c.get(k) for c in cols]
        print(f'{k:42s}' + ''.join(f'{v:14.5f}' if isinstance(v, (int, float)) else f'{str(v):>14s}' for v in vals))
