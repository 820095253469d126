#!/usr/bin/env bash
# Round-2 GPU call N: parity suite (incl. degenerate meshes), relMSE at equal time uniform vs guided on C3 (analytic) and c1 (16 k-spp reference)
O=gpurun_out/r2n; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest.log
timeout 900 python tools/equal_time.py --workload c3 --every 16 --out /tmp/eq > $O/equal_time_c3.json 2> $O/equal_time_c3.err; echo "c3 rc=$?"; tail -2 $O/equal_time_c3.err
timeout 1500 python tools/equal_time.py --workload c1 --spp 1024 --ref-spp 16384 --every 64 --out /tmp/eq > $O/equal_time_c1.json 2> $O/equal_time_c1.err; echo "c1 rc=$?"; tail -2 $O/equal_time_c1.err
python - <<'PY'
import json
for n in ("c3","c1"):
    try:
        d=json.loads(open('gpurun_out/r2n/equal_time_%s.json'%n).read().strip().splitlines()[-1])
        print(n, "uniform %.0f ms relmse %.5f | guided %.0f ms relmse %.5f | equal time %s" % (d["uniform"]["duration_ms"], d["uniform"]["final_relmse"], d["guided"]["duration_ms"], d["guided"]["final_relmse"], d["equal_time"]))
    except Exception as e: print(n, "failed", e)
PY
