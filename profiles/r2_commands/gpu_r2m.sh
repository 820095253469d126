#!/usr/bin/env bash
# Round-2 GPU call M (1 GPU): final ncu evidence of the round -- launch list of the default bench command, --set full captures of the
# dominant kernels (summarised on the box: hot lines, lane bins, raw metrics; the .ncu-rep files stay there)
O=gpurun_out/r2m; mkdir -p $O
cap() { tag=$1; k=$2; skip=$3; shift 3; timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o /tmp/$tag "$@" > $O/$tag.log 2>&1; echo "$tag rc=$?"
  ncu -i /tmp/$tag.ncu-rep --page source --csv --print-source sass,cuda 2>/dev/null > /tmp/$tag.src.csv
  python tools/ncu_lines.py 60 < /tmp/$tag.src.csv > $O/${tag}_hot_lines.txt
  python tools/ncu_bins.py < /tmp/$tag.src.csv > $O/${tag}_lane_bins.txt
  ncu -i /tmp/$tag.ncu-rep --page raw --csv > $O/${tag}_raw.csv 2>/dev/null; cat $O/${tag}_lane_bins.txt; }
C3="python bench.py --spp 16 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-sub"
C4="python bench.py --workload c4-small --spp 8 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
C1="python bench.py --workload c1 --spp 16 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
G="python bench.py --spp 16 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e"
$C3 > $O/c3_plain.json 2>&1 && cap c3_walk k_walk_persistent 1 $C3
$C4 > $O/c4s_plain.json 2>&1 && cap c4s_walk k_walk_persistent 1 $C4
$C1 > $O/c1_plain.json 2>&1 && cap c1_walk k_walk_persistent 1 $C1
$G > $O/g_plain.json 2>&1 && cap g_bounce k_g_bounce 60 $G
FULL="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub"
$FULL > $O/c3_full_plain.json 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv $FULL > $O/c3_launch.log 2>&1; echo "launch list rc=$?"
ls -la $O | head -30
# e2e stage breakdown + C5 at N = 1 (denominators of the 8-GPU efficiency)
WOST_BENCH_DEBUG=1 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-sub > $O/e2e_debug.json 2> $O/e2e_debug.err; grep "e2e stages" $O/e2e_debug.err
pick() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
    print("N=%d value %.0f M ms/step %.1f" % (d["n_gpus"], d["value"]/1e6, d["ms_per_step"]))
except Exception as e:
    print("parse failed", e)
PY
}
timeout 900 python bench.py --workload c5 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $O/c5_uniform_n1.json 2> $O/c5_uniform_n1.err; echo "c5 uniform n1 rc=$?"; pick $O/c5_uniform_n1.json
timeout 900 python bench.py --workload c5 --integrator guided --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $O/c5_guided_n1.json 2> $O/c5_guided_n1.err; echo "c5 guided n1 rc=$?"; pick $O/c5_guided_n1.json
