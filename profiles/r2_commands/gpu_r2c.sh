#!/usr/bin/env bash
# Round-2 GPU call C: new bench.py (sub-workloads, parity check, reference arm with the reference's CUDA binary), guided profile, 3D after OBB in the second pass
O=gpurun_out/r2c; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
t0=$(date +%s); timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench default rc=$? $(( $(date +%s) - t0 )) s"; tail -c 600 $O/bench_default.err
t0=$(date +%s); timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench reference rc=$? $(( $(date +%s) - t0 )) s"; tail -c 400 $O/bench_reference.err
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("value %.1f M ms/step %.1f second passes/step %.3f %s" % (d["value"]/1e6, d["ms_per_step"], d["config"].get("closest_point_second_passes_per_walk_step", -1), d.get("profile_ms_last_solve")))
except Exception as e:
    print("parse failed", e)
PY
)"; }
WOST_PROFILE=1 run c3_guided_profile --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
WOST_PROFILE=1 run c1_guided_profile --workload c1 --spp 256 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
run c4small_uniform --workload c4-small --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
WOST_PROFILE=1 run c4small_guided --workload c4-small --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
run c4_uniform --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
