#!/bin/bash
# Development aid: time several builds of the library on the same box. Usage: scripts/variant_bench.sh build/a.so build/b.so ...
# (each is copied over acinoset_b200/libacino_b200.so in turn; the last one stays)
for v in "$@"; do
  cp "$v" acinoset_b200/libacino_b200.so
  python bench.py --no-cpu-baseline --e2e-solves 1 --steps 300 --warmup 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); r=d['roofline']
print('$v', 'step_ms %.4f' % d['ms_per_step'], 'K4_ms %.4f' % r['kernel_ms'], 'K6_ms %.4f' % r['backsub_kernel']['kernel_ms'], 'solve_ms %.2f' % d['solve']['ms'], 'iters', d['solve']['iterations'], 'cost_final %.10e' % d['solve']['cost_final'])
"
done
