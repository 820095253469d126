#!/usr/bin/env bash
# Round-2 GPU call B: ncu captures of the walk kernels after the oriented-box records (each after a clean run of the same command).
O=gpurun_out/r2b; mkdir -p $O
C3="python bench.py --spp 16 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
C4="python bench.py --workload c4-small --spp 8 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$C3 > $O/c3_plain.json 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_walk_persistent -s 1 -c 1 -o $O/c3_walk $C3 > $O/c3_ncu.log 2>&1; echo "c3 ncu rc=$?"
$C4 > $O/c4s_plain.json 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_walk_persistent -s 1 -c 1 -o $O/c4s_walk $C4 > $O/c4s_ncu.log 2>&1; echo "c4s ncu rc=$?"
FULL="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$FULL > $O/c3_full_plain.json 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv $FULL > $O/c3_launch.log 2>&1; echo "launch list rc=$?"
ls -la $O
