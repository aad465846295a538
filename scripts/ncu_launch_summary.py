"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and mean time, share."""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
agg = collections.OrderedDict()
for r in rows[1:]:
    d = dict(zip(hdr, r))
    try:
        v = float(d['Metric Value'].replace(',', ''))
    except (KeyError, ValueError):
        continue
    k = d['Kernel Name'].split('(')[0][-70:]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(t for _, t in agg.values())
print(f'{"kernel":70s} {"n":>5s} {"total ms":>10s} {"mean us":>10s} {"share":>7s}')
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f'{k:70s} {n:5d} {t/1e6:10.3f} {t/n/1e3:10.1f} {100*t/tot:6.1f}%')
