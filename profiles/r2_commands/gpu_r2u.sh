#!/usr/bin/env bash
# Round-2 GPU call U (--gpus 4): final code on several GPUs -- 2-rank parity tests, multi-GPU elaina-exec, bench at N = 4
O=gpurun_out/r2u; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multi.py -q > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -4 $O/pytest_multi.log
timeout 900 python tools/exec_multi_check.py --gpus 4 > $O/exec_multi.json 2> $O/exec_multi.err; echo "exec multi rc=$?"; python -c "
import json; d=json.loads(open('$O/exec_multi.json').read().strip().splitlines()[-1]); print({k:(v.get('bit_identical'), v.get('walk_steps_equal')) for k,v in d.items() if isinstance(v,dict)})"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531"
timeout 900 $TR bench.py --gpus 4 --steps 5 --warmup 3 > $O/bench_n4.json 2> $O/bench_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2u/bench_n4.json').read().splitlines() if l.startswith('{')][-1])
print("N=%d value %.0f M e2e %.0f M ms/step %.1f e2e relmse %s" % (d["n_gpus"], d["value"]/1e6, d["e2e"]["value"]/1e6, d["ms_per_step"], d["e2e"].get("relmse_vs_analytic")))
for s in d.get("sub", []): print("   ", s["name"], "%.0f M e2e %.0f M" % (s["value"]/1e6, s["e2e"]["value"]/1e6))
PY
