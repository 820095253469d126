#!/usr/bin/env bash
# Round-2 GPU call T: final validation of the round -- parity suite, smoke(), default bench line, reference arm
O=gpurun_out/r2t; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
t0=$(date +%s); timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err; echo "bench default rc=$? $(( $(date +%s) - t0 )) s"; tail -c 300 $O/bench_default.err
t0=$(date +%s); timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench reference rc=$? $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2t/bench_default.json').read().splitlines() if l.startswith('{')][-1])
print("head %.0f M e2e %.0f M frac %.3f ms/step %.1f"%(d["value"]/1e6,d["e2e"]["value"]/1e6,d["roofline"]["frac"],d["ms_per_step"]), d.get("parity_check"), d["clocks"])
for s in d["sub"]: print("  ", s["name"], "%.0f M e2e %.0f M"%(s["value"]/1e6, s["e2e"]["value"]/1e6))
r=json.loads([l for l in open('gpurun_out/r2t/bench_reference.json').read().splitlines() if l.startswith('{')][-1])
print("reference arm %.2f M (cpu port, %d cores); reference_cuda %s" % (r["value"]/1e6, r["cpu_baseline"]["cores"], {k:r["reference_cuda"].get(k) for k in ("value","rc","unavailable")}))
PY
