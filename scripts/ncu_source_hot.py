"""Hottest source lines of an `ncu --set full --import-source on` report (stall samples per CUDA source line):
    ncu -i rep.ncu-rep --page source --print-source cuda,sass --csv > src.csv ; python scripts/ncu_source_hot.py src.csv [n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors='ignore')))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = collections.OrderedDict()
fname, h = '', None
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path':
        fname = r[1].split('/')[-1]
        continue
    if len(r) > 4 and r[0] == 'Line No':
        h = r
        ni, ii = h.index('# Samples'), h.index('Instructions Executed')
        names = [(c, i) for i, c in enumerate(h) if c.startswith('stall_') and 'Not' not in c]
        continue
    if h is None or len(r) <= ni:
        continue
    try:
        n = int(r[ni].replace(',', ''))
    except ValueError:
        continue
    key = (fname, r[0], r[1].strip())
    a = agg.setdefault(key, [0, 0, collections.Counter()])
    a[0] += n
    a[1] += int(r[ii].replace(',', '') or 0)
    for c, i in names:
        v = int(r[i].replace(',', '') or 0)
        if v:
            a[2][c[6:]] += v
tot = sum(a[0] for a in agg.values())
print('total samples', tot)
for (f, ln, src), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top_n]:
    top = ' '.join(f'{k}={v}' for k, v in a[2].most_common(3))
    print(f"{a[0]:6d} {100 * a[0] / tot:5.1f}% {f}:{ln:>4s} inst={a[1]:8d} {top} | {src[:100]}")
