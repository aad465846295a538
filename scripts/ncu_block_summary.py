"""Per basic block of an `ncu --page source --csv` export (SASS view): length, fp64 instructions, share of the samples,
executions and the top stall reasons. Usage: python scripts/ncu_block_summary.py export.csv [min_share]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; col = {h: i for i, h in enumerate(hdr)}
st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
def I(x):
    try: return int(x)
    except ValueError: return 0
out, launches = [], 0
for r in rows[2:]:
    if len(r) < len(hdr) or r[col['Source']].strip() == 'Source':
        launches += len(r) < len(hdr)
        continue
    out.append((r[col['Source']].strip(), I(r[col['# Samples']]), I(r[col['Instructions Executed']]), {k: I(r[col[k]]) for k in st}))
tot = sum(o[1] for o in out)
blk, acc = [], []
for i, o in enumerate(out):
    acc.append((i,) + o)
    w = o[0].split()
    if w and (w[0].startswith('BRA') or (w[0].startswith('@') and len(w) > 1 and w[1].startswith('BRA')) or w[0].startswith('EXIT')):
        blk.append(acc); acc = []
blk.append(acc)
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
print('total samples', tot)
for b in blk:
    if not b: continue
    s = sum(o[2] for o in b); n = sum(o[3] for o in b)
    if s < thr * tot: continue
    agg = {}
    for o in b:
        for k, v in o[4].items(): agg[k] = agg.get(k, 0) + v
    top = sorted(agg.items(), key=lambda x: -x[1])[:4]
    nd = sum(1 for o in b if o[1].split()[0] in ('DFMA', 'DMUL', 'DADD'))
    print(f'instr {b[0][0]:5d}-{b[-1][0]:5d} len {len(b):4d} fp64 {nd:3d} samples {100*s/tot:5.1f}% exec/instr {n/len(b)/1e6:7.3f}M  {[(k[6:], round(100*v/max(s,1))) for k, v in top]}  first: {b[0][1][:40]}')
