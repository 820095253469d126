#!/usr/bin/env bash
# Round-2 GPU call Q: ncu capture of k_walk_parked on C3 (lanes per trip, instruction mix)
O=gpurun_out/r2q; mkdir -p $O
C3="python bench.py --spp 16 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-sub"
$C3 > $O/c3_plain.json 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_walk_parked -s 1 -c 1 -o /tmp/c3p $C3 > $O/c3p.log 2>&1; echo "rc=$?"
ncu -i /tmp/c3p.ncu-rep --page source --csv --print-source sass,cuda 2>/dev/null > /tmp/c3p.src.csv
python tools/ncu_lines.py 70 < /tmp/c3p.src.csv > $O/c3p_hot_lines.txt
python tools/ncu_bins.py < /tmp/c3p.src.csv > $O/c3p_lane_bins.txt; cat $O/c3p_lane_bins.txt
ncu -i /tmp/c3p.ncu-rep --page raw --csv > $O/c3p_raw.csv 2>/dev/null
grep -n "integrator.cu\|geom.cuh:120" $O/c3p_hot_lines.txt | head -30
