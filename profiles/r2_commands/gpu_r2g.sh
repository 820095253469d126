#!/usr/bin/env bash
# Round-2 GPU call G: full parity suite after the guided fast-math change + items-per-lane sweep for the small guided launches
O=gpurun_out/r2g; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $O/pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M ms/step %.1f relmse %s %s" % (d["value"]/1e6, d["ms_per_step"], d["config"].get("relmse_vs_analytic"), d.get("profile_ms_last_solve")))
except Exception as e:
    print("parse failed", e)
PY
)"; }
for ipl in 1 2 3 4 6 8; do WOST_G_ITEMS_PER_LANE=$ipl WOST_PROFILE=1 run c3_guided_ipl$ipl --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e; done
for ipl in 1 4; do WOST_G_ITEMS_PER_LANE=$ipl WOST_PROFILE=1 run c4s_guided_ipl$ipl --workload c4-small --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e; done
