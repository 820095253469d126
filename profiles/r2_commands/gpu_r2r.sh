#!/usr/bin/env bash
# Round-2 GPU call R: shared-memory stack levels of the walk kernel (L1 carve-out vs spills)
O=gpurun_out/r2r; mkdir -p $O
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M ms/step %.1f relmse %s" % (d["value"]/1e6, d["ms_per_step"], d["config"].get("relmse_vs_analytic")))
except Exception as e:
    print("parse failed", e)
PY
)"; tail -c 200 $O/$name.err; }
for lv in 4 8 12 16 20; do WOST_STACK_LEVELS=$lv run c3_lv$lv --steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-sub; done
for lv in 4 8 12 16 20; do WOST_STACK_LEVELS=$lv run c1_lv$lv --workload c1 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for lv in 4 8 12 16 20; do WOST_STACK_LEVELS=$lv run c4s_lv$lv --workload c4-small --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for lv in 8 12 20; do WOST_STACK_LEVELS=$lv run c4_lv$lv --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
