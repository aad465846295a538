"""Executed fp64 flops per launch (predicated-on thread-level DFMA x2 + DMUL + DADD) from
`ncu -i X.ncu-rep --page source --csv --print-source sass --kernel-name regex:K`. Usage: python scripts/ncu_fp64_flops.py K.csv"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr=None; ops=collections.Counter(); n_kern=0
for r in rows:
    if len(r)>4 and r[0]=='Address': hdr=r; n_kern+=1; continue
    if hdr is None or len(r)<10: continue
    try: ti=int(r[hdr.index('Predicated-On Thread Instructions Executed')])
    except: continue
    toks=r[1].split()
    if not toks: continue
    op=toks[0] if not toks[0].startswith('@') else toks[1]
    ops[op.split('.')[0]]+=ti
fl=2*ops['DFMA']+ops['DMUL']+ops['DADD']
print('launches',n_kern,'per launch: DFMA',ops['DFMA']//n_kern,'DMUL',ops['DMUL']//n_kern,'DADD',ops['DADD']//n_kern,'flops',fl//n_kern)
