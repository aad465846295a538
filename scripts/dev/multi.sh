#!/bin/bash
# usage: multi.sh N
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "1" ]; then TR="python"; fi
run() { port=$1; shift; if [ "$N" = "1" ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; fi; }
if [ "$N" != "1" ]; then
run 29601 bench.py --gpus $N --steps 20 --warmup 5 --no-c1 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
fi
run 29602 bench.py --gpus $N --frames $((2000000/N)) --masked-frac 0.3 --outlier-frac 0.05 --strong-frames 2000000 --steps 10 --warmup 3 --no-c1 --no-cpu-baseline > gpurun_out/r02_bench_c5_n$N.json 2> gpurun_out/r02_bench_c5_n$N.err
if [ -n "$WITH_C3" ]; then run 29603 scripts/tri_sweep.py > gpurun_out/r02_tri_sweep_c3_n$N.jsonl 2> gpurun_out/c3_n$N.err; fi
tail -c 300 gpurun_out/r02_bench_c5_n$N.err
python - <<PY
import json
for f in ("gpurun_out/r02_bench_n$N.json", "gpurun_out/r02_bench_c5_n$N.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        s = d.get("strong", {})
        print(f, d["value"], d["ms_per_step"], "strong", s.get("ms_per_step"), s.get("solve_ms"), s.get("e2e_solve_ms"), s.get("efficiency_vs_n1"), s.get("cost_final"))
    except Exception as e:
        print(f, "ERR", e)
PY
