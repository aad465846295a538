"""Per CUDA source line aggregation of `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name ...`:
samples and executed warp instructions per (file, line), plus totals per line range given as file:lo-hi arguments."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
cur_file = None
hdr = None
agg = collections.OrderedDict()
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]; continue
    if len(r) > 4 and r[0] == 'Line No':
        hdr = r; continue
    if hdr is None or len(r) < 10: continue
    if r[0] != '':      # a source line row: aggregated metrics of its SASS
        # quoted commas in the source text can shift columns: take from the right
        off = len(r) - len(hdr)
        def col(name):
            return r[hdr.index(name, 4) + off] if name in hdr[4:] else ''
        try:
            samp = int(col('# Samples') or 0); inst = int(col('Instructions Executed') or 0)
        except ValueError:
            continue
        a = agg.setdefault((cur_file, int(r[0])), [0, 0, r[1][:90]])
        a[0] += samp; a[1] += inst
tot_s = sum(a[0] for a in agg.values()); tot_i = sum(a[1] for a in agg.values())
print('total samples', tot_s, 'instructions', tot_i)
for spec in sys.argv[2:]:
    f, rng = spec.split(':'); lo, hi = map(int, rng.split('-'))
    s = sum(a[0] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    i = sum(a[1] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    print(f'{spec:40s} samples {100*s/tot_s:5.1f}%  inst {100*i/tot_i:5.1f}%')
if len(sys.argv) == 2:
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
        print(f'{f}:{l:<5d} samples {100*a[0]/tot_s:5.1f}% inst {100*a[1]/tot_i:5.1f}%  {a[2]}')
