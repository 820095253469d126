#!/usr/bin/env bash
# Round-2 GPU call V (--gpus 8): default bench line with the final code at N = 8 (device-timed and e2e scaling)
O=gpurun_out/r2v; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
timeout 900 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench n8 rc=$?"
WOST_BENCH_DEBUG=1 timeout 900 $TR bench.py --gpus 8 --steps 3 --warmup 2 --no-sub > $O/bench_n8_debug.json 2> $O/bench_n8_debug.err; grep "e2e stages" $O/bench_n8_debug.err | tail -8
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2v/bench_n8.json').read().splitlines() if l.startswith('{')][-1])
print("N=%d value %.0f M e2e %.0f M ms/step %.1f e2e relmse %s" % (d["n_gpus"], d["value"]/1e6, d["e2e"]["value"]/1e6, d["ms_per_step"], d["e2e"].get("relmse_vs_analytic")))
for s in d.get("sub", []): print("   ", s["name"], "%.0f M e2e %.0f M" % (s["value"]/1e6, s["e2e"]["value"]/1e6))
PY
