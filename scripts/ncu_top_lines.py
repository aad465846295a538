"""Hottest source lines with their top stall reasons from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass
--kernel-name regex:K`. Usage: python scripts/ncu_top_lines.py K.csv [lo hi [n]] (lo-hi: line range of acino_sba_schur.cuh whose stall
reasons are summed)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
lo, hi = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (0, 10**9)
cur=None; hdr=None; out=[]
tot=0
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>4 and r[0]=='Line No': hdr=r; continue
    if hdr is None or len(r)<10 or r[0]=='': continue
    off=len(r)-len(hdr)
    g=lambda n: r[hdr.index(n,4)+off]
    try: s=int(g('# Samples') or 0)
    except: continue
    tot+=s
    st={h[6:]:int(r[i+off] or 0) for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h and i>4}
    out.append((s,cur,int(r[0]),r[1][:70],st))
phase=collections.Counter()
for s,f,l,src,st in out:
    if f=='acino_sba_schur.cuh' and lo<=l<=hi:
        for k,v in st.items(): phase[k]+=v
print('total',tot, 'range stalls', phase.most_common(8))
for s,f,l,src,st in sorted(out,key=lambda x:-x[0])[:int(sys.argv[4]) if len(sys.argv)>4 else 25]:
    top=sorted(st.items(),key=lambda x:-x[1])[:3]
    print(f'{f}:{l:4d} {100*s/tot:5.1f}% {src:70s} {top}')
