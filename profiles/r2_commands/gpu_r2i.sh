#!/usr/bin/env bash
# Round-2 GPU call I: refill-threshold sweep of the walk kernels after the oriented boxes (queries are 3x shorter: the optimum may have moved)
O=gpurun_out/r2i; mkdir -p $O
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M ms/step %.1f" % (d["value"]/1e6, d["ms_per_step"]))
except Exception as e:
    print("parse failed", e)
PY
)"; }
for thr in 4 6 8 10 12 16; do WOST_REFILL_THRESHOLD=$thr run c3_refill$thr --steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-sub; done
for thr in 6 10 14; do WOST_REFILL_THRESHOLD=$thr run c1_refill$thr --workload c1 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for thr in 6 10 14; do WOST_REFILL_THRESHOLD=$thr run c4s_refill$thr --workload c4-small --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
