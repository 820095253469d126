#!/usr/bin/env bash
# Round-2 GPU call O: bounce kernel at 5 CTAs/SM (96 registers, 2-level shared stack), smoke(), parity suite
O=gpurun_out/r2o; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/smoke.log
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
WOST_BENCH_DEBUG=1 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub > $O/e2e_debug.json 2> $O/e2e_debug.err; grep "e2e stages" $O/e2e_debug.err
