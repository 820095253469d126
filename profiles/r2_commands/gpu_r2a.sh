#!/usr/bin/env bash
# Round-2 GPU call A: parity suite + bench lines of every workload after the oriented-box traversal records.
O=gpurun_out/r2a; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("value %.1f M e2e %s ms/step %.1f frac %.3f" % (d["value"]/1e6, ("%.1f M" % (d["e2e"]["value"]/1e6)) if d.get("e2e") and d["e2e"].get("value") else None, d["ms_per_step"], d["roofline"]["frac"]))
except Exception as e:
    print("parse failed", e)
PY
)"; }
run c3_uniform --steps 5 --warmup 3
run c3_guided --steps 3 --warmup 2 --integrator guided --no-cpu-baseline
run c1_uniform --workload c1 --spp 256 --steps 2 --warmup 1 --no-cpu-baseline
run c1_guided --workload c1 --spp 256 --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
run c2_uniform --workload c2 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline
run c4small_uniform --workload c4-small --steps 2 --warmup 1 --no-cpu-baseline
run c4small_guided --workload c4-small --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
run c4_uniform --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
WOST_L2_PERSIST=0 run c4_uniform_nopersist --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
run c4_guided --workload c4 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt
