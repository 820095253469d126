#!/usr/bin/env bash
# Round-2 GPU call L (--gpus N): strong scaling of the default bench line (C3 uniform + sub-workloads) and of C5 through the NCCL path inside libwost.so
N=${1:-8}; O=gpurun_out/r2l; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"
pick() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("N=%d value %.0f M e2e %s ms/step %.1f" % (d["n_gpus"], d["value"]/1e6, d["e2e"]["value"] and round(d["e2e"]["value"]/1e6), d["ms_per_step"]))
    for s in d.get("sub", []): print("   ", s["name"], "%.0f M e2e %.0f M ms/step %.1f" % (s["value"]/1e6, s["e2e"]["value"]/1e6, s["ms_per_step"]))
except Exception as e:
    print("parse failed", e)
PY
}
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N rc=$?"; pick $O/bench_n$N.json; grep -v "^\*\|OMP_NUM" $O/bench_n$N.err | tail -3
timeout 900 $TR bench.py --gpus $N --workload c5 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $O/c5_uniform_n$N.json 2> $O/c5_uniform_n$N.err; echo "c5 uniform n$N rc=$?"; pick $O/c5_uniform_n$N.json
timeout 900 $TR bench.py --gpus $N --workload c5 --integrator guided --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $O/c5_guided_n$N.json 2> $O/c5_guided_n$N.err; echo "c5 guided n$N rc=$?"; pick $O/c5_guided_n$N.json
