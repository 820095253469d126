#!/usr/bin/env bash
# Round-2 GPU call D (1 GPU): ncu source-level captures of the guided kernels (after a clean run of the same command);
# reports are summarised on the box (hot lines + raw metrics) because gpurun_out/ is capped at 64 MiB
O=gpurun_out/r2d; mkdir -p $O
G="python bench.py --spp 16 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e"
$G > $O/g_plain.json 2>&1 || { echo "plain run failed"; tail -5 $O/g_plain.json; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_guided.csv $G > $O/g_launch.log 2>&1; echo "launch list rc=$?"
cap() { k=$1; skip=$2; timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o /tmp/$k $G > $O/$k.log 2>&1; echo "$k rc=$?"
  ncu -i /tmp/$k.ncu-rep --page source --csv --print-source sass,cuda 2>/dev/null | python tools/ncu_lines.py 60 > $O/${k}_hot_lines.txt
  ncu -i /tmp/$k.ncu-rep --page raw --csv > $O/${k}_raw.csv 2>/dev/null; }
cap k_g_bounce 60
cap k_g_arrive 60
cap k_g_train 2
cap k_g_finish 20
ls -la $O
