#!/usr/bin/env bash
# Round-2 GPU call P: two walks per lane (k_walk_parked) vs one (k_walk_persistent): parity suite on the new default, then A/B + tail-threshold sweep
O=gpurun_out/r2p; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_walk.py tests/test_gpu_analytic.py tests/test_exec.py -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M ms/step %.1f relmse %s" % (d["value"]/1e6, d["ms_per_step"], d["config"].get("relmse_vs_analytic")))
except Exception as e:
    print("parse failed", e)
PY
)"; tail -c 200 $O/$name.err; }
C3="--steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-sub"
C1="--workload c1 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
C4S="--workload c4-small --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
WOST_WALK_KERNEL=persistent run c3_persistent $C3
run c3_parked $C3
for thr in 8 16 24; do WOST_REFILL_THRESHOLD=$thr run c3_parked_thr$thr $C3; done
WOST_WALK_KERNEL=persistent run c1_persistent $C1
run c1_parked $C1
for thr in 12 16 28; do WOST_REFILL_THRESHOLD=$thr run c1_parked_thr$thr $C1; done
WOST_WALK_KERNEL=persistent run c4s_persistent $C4S
run c4s_parked $C4S
for thr in 12 16 28; do WOST_REFILL_THRESHOLD=$thr run c4s_parked_thr$thr $C4S; done
run c4_parked --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
run c2_parked --workload c2 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e
