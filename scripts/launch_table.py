"""Print mean duration (and the other collected metrics) per (kernel, grid) of an ncu --csv launch log."""
import collections
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
hdr = rows[hi]
d = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) != len(hdr):
        continue
    key = (r[hdr.index('Kernel Name')].split('(')[0][-30:], r[hdr.index('Grid Size')])
    d.setdefault(key, collections.defaultdict(list))[r[hdr.index('Metric Name')]].append(float(r[hdr.index('Metric Value')].replace(',', '')))
for k, v in d.items():
    print(k, {m.split('.')[0][-22:]: round(sum(x) / len(x), 1) for m, x in v.items()}, 'n=%d' % len(next(iter(v.values()))))
