#!/usr/bin/env bash
# Round-2 GPU call W: parity suite after the far-field guard and the row-wise active list; c2 (far-field walks) and C3 throughput
O=gpurun_out/r2w; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M e2e %s ms/step %.1f" % (d["value"]/1e6, d["e2e"]["value"] and round(d["e2e"]["value"]/1e6,1), d["ms_per_step"]))
except Exception as e:
    print("parse failed", e)
PY
)"; tail -c 200 $O/$name.err; }
run c2_uniform --workload c2 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline
run c2_guided --workload c2 --spp 128 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
run c3_uniform --steps 5 --warmup 2 --no-cpu-baseline --no-sub
