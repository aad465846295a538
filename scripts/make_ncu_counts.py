"""profiles/r02_ncu_counts.json from an ncu --set full capture of scripts/profile_target.py: per kernel the DRAM bytes
(dram__bytes_read.sum + dram__bytes_write.sum), the executed fp64 flops (thread-level DFMA x2 + DMUL + DADD of the SASS
page), fp64 pipe activity, registers and duration per launch, stamped with the hash of the CUDA sources (bench.py attaches
the counts only to a run of the same sources and workload). Kernels launched several times: mean over the launches
selected by --pick (default: all). Usage: python scripts/make_ncu_counts.py capture.ncu-rep [frames masked_frac outlier_frac]"""
import sys, os, csv, io, json, subprocess, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

rep = sys.argv[1]
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 119490
mf = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
of = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0


def ncu(*args):
    return subprocess.run(['ncu', '-i', rep, *args], capture_output=True, text=True).stdout


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return None


rows = list(csv.reader(io.StringIO(ncu('--page', 'raw', '--csv'))))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 'nsecond': 1e-6, 'usecond': 1e-3, 'msecond': 1.0}
agg = collections.OrderedDict()
for r in rows[2:]:
    name = r[col['Kernel Name']].split('(')[0].replace('void ', '').replace('acino::', '')
    name = name.replace('(int)', '').replace('(bool)', '').replace(', 1>', ', true>').replace(', 0>', ', false>')

    def val(metric):
        i = col[metric]
        return num(r[i]) * scale.get(units[i], 1)
    a = agg.setdefault(name, collections.defaultdict(list))
    a['dram'].append(val('dram__bytes_read.sum') + val('dram__bytes_write.sum'))
    a['ms'].append(val('gpu__time_duration.sum'))
    a['pipe'].append(num(r[col['sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']]))
    a['regs'].append(num(r[col['launch__registers_per_thread']]))
out = dict(workload=dict(frames=frames, masked_frac=mf, outlier_frac=of), kernel_source_sha=bench.kernel_source_sha(),
           capture=os.path.basename(rep), how='ncu --set full --clock-control none, scripts/profile_target.py; means over the captured launches', kernels={})
def norm(n):
    n = n.split('(')[0].replace('void ', '').replace('acino::', '') if '<' not in n else n[:n.index('>') + 1].replace('void ', '').replace('acino::', '')
    return n.replace('(int)', '').replace('(bool)', '').replace(', 1>', ', true>').replace(', 0>', ', false>')


# executed thread-level instructions per kernel NAME (the SASS page lists every captured launch under its full name)
src = list(csv.reader(io.StringIO(ncu('--page', 'source', '--csv', '--print-source', 'sass'))))
flops, n_launch = collections.defaultdict(collections.Counter), collections.Counter()
h, cur = None, None
for r in src:
    if len(r) >= 2 and r[0] == 'Kernel Name':
        cur = norm(r[1]); n_launch[cur] += 1; continue
    if len(r) > 4 and r[0] == 'Address':
        h = r; continue
    if h is None or cur is None or len(r) < 10:
        continue
    try:
        ti = int(r[h.index('Predicated-On Thread Instructions Executed')])
    except ValueError:
        continue
    t = r[1].split()
    if t:
        flops[cur][(t[0] if not t[0].startswith('@') else t[1]).split('.')[0]] += ti
for name, a in agg.items():
    ops, n_k = flops.get(name, collections.Counter()), n_launch.get(name, 0)
    fl = (2 * ops['DFMA'] + ops['DMUL'] + ops['DADD']) / max(n_k, 1)
    mean = lambda v: sum(v) / len(v)
    out['kernels'][name] = dict(launches=len(a['ms']), ms_under_ncu=mean(a['ms']), dram_bytes=int(mean(a['dram'])), fp64_flops=int(fl),
                                fp64_pipe_pct=round(mean(a['pipe']), 2), registers=int(a['regs'][0]))
json.dump(out, open(os.path.join(ROOT, 'profiles', 'r02_ncu_counts.json'), 'w'), indent=1)
print(json.dumps(out, indent=1))
