#!/usr/bin/env bash
# Round-2 GPU call D (1 GPU): ncu source-level captures of the guided kernels (after a clean run of the same command)
O=gpurun_out/r2d; mkdir -p $O
G="python bench.py --spp 16 --steps 1 --warmup 1 --integrator guided --no-cpu-baseline --no-e2e"
$G > $O/g_plain.json 2>&1 || { echo "plain run failed"; tail -5 $O/g_plain.json; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_guided.csv $G > $O/g_launch.log 2>&1; echo "launch list rc=$?"
# guiding-phase bounce launch (large): skip the training-phase launches
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_g_bounce -s 60 -c 1 -o $O/g_bounce $G > $O/g_bounce.log 2>&1; echo "bounce rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_g_arrive -s 60 -c 1 -o $O/g_arrive $G > $O/g_arrive.log 2>&1; echo "arrive rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_g_train -s 2 -c 1 -o $O/g_train $G > $O/g_train.log 2>&1; echo "train rc=$?"
ls -la $O
