#!/usr/bin/env bash
# Round-2 GPU call J: refill-threshold sweep, upper range + guided kernels; B2 (reference LBVH on the host vs device query kernels)
O=gpurun_out/r2j; mkdir -p $O
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M ms/step %.1f %s" % (d["value"]/1e6, d["ms_per_step"], d.get("profile_ms_last_solve")))
except Exception as e:
    print("parse failed", e)
PY
)"; }
for thr in 14 18 22 26 32; do WOST_REFILL_THRESHOLD=$thr run c1_refill$thr --workload c1 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for thr in 14 18 22 26 32; do WOST_REFILL_THRESHOLD=$thr run c4s_refill$thr --workload c4-small --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for thr in 10 14; do WOST_REFILL_THRESHOLD=$thr run c3_refill$thr --steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-sub; done
for thr in 6 10 14 18 24; do WOST_REFILL_THRESHOLD=$thr WOST_PROFILE=1 run c3g_refill$thr --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e; done
for thr in 6 12 18 24; do WOST_REFILL_THRESHOLD=$thr WOST_PROFILE=1 run c4sg_refill$thr --workload c4-small --steps 2 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e; done
for thr in 16 24; do WOST_REFILL_THRESHOLD=$thr run c4_refill$thr --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
timeout 900 python tools/b2_queries.py > $O/b2.jsonl 2> $O/b2.err; echo "b2 rc=$?"; tail -3 $O/b2.err; python - <<'PY'
import json
for l in open('gpurun_out/r2j/b2.jsonl'):
    d=json.loads(l); c=d.get("cpu_baseline",{})
    print(d["workload"], d["mesh"], d["query"], "gpu %.1f M q/s e2e %.1f M cpu-ref %.3f M parity %s frac %.3f" % (d["value"]/1e6, d["e2e"]["value"]/1e6, c.get("value",0)/1e6, d.get("parity"), d["roofline"]["frac"]))
PY
