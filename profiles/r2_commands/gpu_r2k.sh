#!/usr/bin/env bash
# Round-2 GPU call K: parity suite, default bench line (new thresholds), C4 as a slice set (uniform + guided)
O=gpurun_out/r2k; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
t0=$(date +%s); timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench default rc=$? $(( $(date +%s) - t0 )) s"; tail -c 300 $O/bench_default.err
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$? $(python - $O/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("value %.1f M e2e %s ms/step %.1f frac %.3f" % (d["value"]/1e6, d["e2e"]["value"] and round(d["e2e"]["value"]/1e6,1), d["ms_per_step"], d["roofline"]["frac"]))
except Exception as e:
    print("parse failed", e)
PY
)"; tail -c 300 $O/$name.err; }
run c4_slices_uniform --workload c4-slices --steps 2 --warmup 1 --no-cpu-baseline
run c4_slices_guided --workload c4-slices --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e
run c2_uniform --workload c2 --spp 128 --steps 2 --warmup 1 --no-cpu-baseline
python tools/b2_queries.py --workloads c1 c4-small > $O/b2.jsonl 2> $O/b2.err; grep silhouette $O/b2.jsonl | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['workload'], d['mesh'], d['query'], round(d['value']/1e6,1), 'M q/s', d.get('parity'))"
