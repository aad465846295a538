#!/bin/bash
# Development aid: per-kernel times of several builds of the library on the same box.
# Usage: scripts/variant_probe.sh build/a.so build/b.so ...   (each is copied over acinoset_b200/libacino_b200.so in turn;
# the original is restored at the end)
cp acinoset_b200/libacino_b200.so /tmp/libacino_orig.so
for v in "$@"; do
  cp "$v" acinoset_b200/libacino_b200.so
  timeout 200 python scripts/k4_probe.py 119490 100 2>&1 | python -c "
import json,sys
ls=[l for l in sys.stdin if l.startswith('{')]
if len(ls)<2: print('$v FAILED', ls); sys.exit(0)
a=json.loads(ls[0]); b=json.loads(ls[1]); k=b['kernels']
print('$v', 'step_wall %.4f' % b['ms_per_step_wall'], 'K4a %.4f K4b %.4f K5 %.4f K6 %.4f' % (k['linearise_ms'],k['accumulate_ms'],k['solve_ms'],k['backsub_ms']), 'solve_ms %.2f' % a['solve']['solve_ms'], 'iters', a['solve']['iterations'], 'cost_final %.12e' % a['solve']['cost_final'])
"
done
cp /tmp/libacino_orig.so acinoset_b200/libacino_b200.so
