#!/bin/bash
# profiles/r02_ncu_full_raw.csv and profiles/r02_ncu_full_summary.txt from an `ncu --set full` capture of
# scripts/profile_target.py (run scripts/make_ncu_counts.py on the same capture first: the flop / DRAM lines come
# from profiles/r02_ncu_counts.json). Usage: scripts/make_ncu_summary.sh capture.ncu-rep "description of the run"
set -e
rep=$1; what=$2
out=profiles/r02_ncu_full_summary.txt
ncu -i "$rep" --page raw --csv > profiles/r02_ncu_full_raw.csv
sha=$(python -c "import json; print(json.load(open('profiles/r02_ncu_counts.json'))['kernel_source_sha'])")
{
  echo "ncu --set full --clock-control none --import-source on, python scripts/profile_target.py $what, sources $sha"
  echo
  python scripts/ncu_raw_summary.py profiles/r02_ncu_full_raw.csv
  echo
  echo "=== executed fp64 flops per launch (thread-level DFMA x2 + DMUL + DADD), DRAM bytes per launch: profiles/r02_ncu_counts.json"
  python - <<'PY'
import json
d = json.load(open('profiles/r02_ncu_counts.json'))
pk = json.load(open('MEASURED_PEAKS.json'))
hbm = pk['hbm_gbs']
fp64 = 36.0      # TFLOP/s: the DFMA loop bench.py measures on the same box (fp64_peak_tflops_measured)
for k, v in d['kernels'].items():
    ms = v['ms_under_ncu']
    tf = v['fp64_flops'] / ms / 1e9
    gb = v['dram_bytes'] / ms / 1e6
    print(f"{k:36s} flops {v['fp64_flops']:12d} {tf:6.1f} TFLOP/s = {100 * tf / fp64:4.1f} % of {fp64}  | DRAM {v['dram_bytes'] / 1e6:7.1f} MB = "
          f"{gb / 1e3:5.2f} TB/s = {100 * gb / hbm:4.1f} % of {hbm} GB/s | pipe {v['fp64_pipe_pct']:5.2f} % | regs {int(v['registers'])} | {ms * 1e3:6.1f} us under ncu")
PY
  for k in schur_accum schur_linearise schur_backsub; do
    echo
    echo "=== $k: instruction mix and stall reasons (all captured launches), then basic blocks >= 2 % of the samples"
    ncu -i "$rep" --page source --csv --print-source sass -k regex:$k > /tmp/src_$k.csv
    python scripts/ncu_source_summary.py /tmp/src_$k.csv > /tmp/ss_a.txt 2>&1 || true
    sed -n 1,40p /tmp/ss_a.txt
    python scripts/ncu_block_summary.py /tmp/src_$k.csv > /tmp/ss_b.txt 2>&1 || true
    sed -n 1,14p /tmp/ss_b.txt
  done
} > $out
echo wrote $out
