"""Key per-launch metrics from `ncu -i X.ncu-rep --page raw --csv` (stdin or file) as a compact table."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = [('gpu__time_duration.sum', 'time'), ('dram__bytes_read.sum', 'dram_rd'), ('dram__bytes_write.sum', 'dram_wr'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
        ('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'fp64pipe%'),
        ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue%'),
        ('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'lsu%'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'),
        ('launch__registers_per_thread', 'regs'), ('smsp__inst_executed.sum', 'warp_insts'),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smem_conflicts')]
idx = [(hdr.index(k), n) for k, n in want if k in hdr]
kn = hdr.index('Kernel Name')
print('kernel'.ljust(28), ' '.join(n.rjust(14) for _, n in idx))
for r in rows[2:]:
    name = r[kn].split('(')[0].replace('void ', '')[:28]
    cells = []
    for i, n in idx:
        v = r[i]
        try:
            v = f'{float(v):.4g}' + (units[i][:5] if units[i] not in ('%', '') else '')
        except ValueError:
            pass
        cells.append(v.rjust(14))
    print(name.ljust(28), ' '.join(cells))
