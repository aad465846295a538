"""Warp instructions, samples and active lanes per line range of one source file, from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K`.
Usage: python scripts/ncu_phase_summary.py K.csv file.cuh name:lo-hi [name:lo-hi ...]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
fname = sys.argv[2]
cur = None; hdr = None; agg = {}
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) > 4 and r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) < 10 or r[0] == '': continue
    off = len(r) - len(hdr)
    g = lambda n: r[hdr.index(n, 4) + off]
    try:
        wi = int(g('Instructions Executed')); ti = int(g('Predicated-On Thread Instructions Executed')); sm = int(g('# Samples') or 0)
    except ValueError:
        continue
    a = agg.setdefault((cur, int(r[0])), [0, 0, 0]); a[0] += wi; a[1] += ti; a[2] += sm
W = sum(a[0] for a in agg.values()); T = sum(a[1] for a in agg.values()); S = sum(a[2] for a in agg.values())
print(f'all files: warp instructions {W}, samples {S}, active lanes {T / max(W, 1):.1f}')
for spec in sys.argv[3:]:
    name, rng = spec.split(':'); lo, hi = map(int, rng.split('-'))
    sel = [a for (f, l), a in agg.items() if f == fname and lo <= l <= hi]
    w = sum(a[0] for a in sel); t = sum(a[1] for a in sel); s = sum(a[2] for a in sel)
    print(f'{name:34s} lines {lo:4d}-{hi:4d}  instructions {100 * w / W:5.1f}%  samples {100 * s / S:5.1f}%  lanes {t / max(w, 1):4.1f}')
others = [(f, a) for (f, l), a in agg.items() if f != fname]
by = {}
for f, a in others:
    b = by.setdefault(f, [0, 0, 0]); b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
for f, b in sorted(by.items(), key=lambda x: -x[1][2])[:4]:
    print(f'{f:34s} (whole file)      instructions {100 * b[0] / W:5.1f}%  samples {100 * b[2] / S:5.1f}%  lanes {b[1] / max(b[0], 1):4.1f}')
