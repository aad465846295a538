"""Aggregate an `ncu --page source --csv` export by source line: samples, executed instructions, top stall reasons."""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot_s = 0
# this page is SASS-level when --page source has no source correlation; detect a "Source" column holding SASS text
agg = collections.OrderedDict()
opagg = collections.Counter(); opsamp = collections.Counter()
for r in rows[2:]:
    if len(r) < len(hdr): continue
    src = r[col['Source']].strip()
    try:
        samp = int(r[col['# Samples']] or 0); inst = int(r[col['Instructions Executed']] or 0)
    except ValueError:
        continue
    op = src.split()[0] if src else ''
    if op.startswith('@'): op = src.split()[1]
    op = op.split('.')[0]
    opagg[op] += inst; opsamp[op] += samp
    tot_s += samp
    st = {s: int(r[col[s]] or 0) for s in stalls}
    a = agg.setdefault(src, [0, 0, collections.Counter()]); a[0] += samp; a[1] += inst; a[2].update(st)
print('total samples', tot_s)
print('--- by opcode: instructions executed, samples')
ti = sum(opagg.values())
for op, n in opagg.most_common(25):
    print(f'{op:12s} {n:12d} {100*n/ti:5.1f}%   samples {100*opsamp[op]/max(tot_s,1):5.1f}%')
allst = collections.Counter()
for a in agg.values(): allst.update(a[2])
print('--- stall reasons (all samples)')
for k, v in allst.most_common(8): print(f'{k:28s} {100*v/max(sum(allst.values()),1):5.1f}%')
