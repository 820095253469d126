#!/usr/bin/env bash
# Round-2 GPU call F: guided layer with hardware transcendentals + 4 CTAs/SM bounce kernel
O=gpurun_out/r2f; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M e2e %s ms/step %.1f relmse %s %s" % (d["value"]/1e6, d["e2e"]["value"] and round(d["e2e"]["value"]/1e6,1), d["ms_per_step"], d["config"].get("relmse_vs_analytic"), d.get("profile_ms_last_solve")))
except Exception as e:
    print("parse failed", e)
PY
)"; }
WOST_PROFILE=1 run c3_guided --steps 3 --warmup 1 --integrator guided --no-cpu-baseline
WOST_PROFILE=1 run c1_guided --workload c1 --spp 256 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
WOST_PROFILE=1 run c4small_guided --workload c4-small --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
WOST_PROFILE=1 run c2_guided --workload c2 --spp 128 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
