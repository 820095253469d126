#!/usr/bin/env bash
# Round-2 GPU call H: 2D candidate-mode traversal (bounds only + exact finish): parity suite, then throughput + refill-threshold sweep
O=gpurun_out/r2h; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    c=d["config"]
    print("value %.1f M ms/step %.1f relmse %s reruns/step %.4f %s" % (d["value"]/1e6, d["ms_per_step"], c.get("relmse_vs_analytic"), c.get("closest_point_candidate_reruns_per_walk_step",-1), d.get("profile_ms_last_solve")))
except Exception as e:
    print("parse failed", e)
PY
)"; }
run c3_uniform --steps 3 --warmup 2 --no-cpu-baseline --no-e2e --no-sub
for thr in 4 8 10 12 16; do WOST_REFILL_THRESHOLD=$thr run c3_uniform_refill$thr --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-sub; done
run c1_uniform --workload c1 --spp 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
run c2_uniform --workload c2 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
WOST_PROFILE=1 run c3_guided --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
WOST_PROFILE=1 run c1_guided --workload c1 --spp 256 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
