#!/usr/bin/env bash
# Round-2 GPU call S: spp-batch sweep (launches per solve vs accumulator footprint), then the final default bench line + reference arm
O=gpurun_out/r2s; mkdir -p $O
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M ms/step %.1f launches %d" % (d["value"]/1e6, d["ms_per_step"], d["gpu_launches"]))
except Exception as e:
    print("parse failed", e)
PY
)"; tail -c 200 $O/$name.err; }
for sb in 16 44 64 128 256; do run c3_sb$sb --spp-batch $sb --steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-sub; done
for sb in 16 64 128; do run c1_sb$sb --workload c1 --spp 128 --spp-batch $sb --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for sb in 32 64; do run c4s_sb$sb --workload c4-small --spp-batch $sb --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
for sb in 44 128; do run c3g_sb$sb --integrator guided --spp-batch $sb --steps 2 --warmup 1 --no-cpu-baseline --no-e2e; done
