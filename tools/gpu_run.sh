#!/usr/bin/env bash
# One runner for the GPU box (replaces the per-experiment tools/gpu_r2?.sh of round 2, kept under profiles/r2_commands/ as the
# provenance of the round-2 profiles). Run through gpurun from the repo root:  gpurun --timeout 1500 -- 'tools/gpu_run.sh <what> [tag]'
#   tests          pytest -m gpu                                         -> gpurun_out/<tag>_gpu_tests.log
#   bench          bench.py (reference arm first, like the driver)       -> gpurun_out/<tag>_bench{,_reference}.json
#   launches       ncu launch list of a short bench run                  -> gpurun_out/<tag>_launches.csv
#   ncu <workload> ncu --set full of the dominant walk kernel            -> gpurun_out/<tag>_<workload>.ncu-rep
#   l2             tools/l2_peak (32 and 8 MiB)                          -> gpurun_out/<tag>_l2_peak_*.json
set -u
what=${1:-tests}; shift || true
mkdir -p gpurun_out
case "$what" in
  tests)    tag=${1:-r3}; python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_gpu_tests.log 2>&1; rc=$?; tail -4 gpurun_out/${tag}_gpu_tests.log; exit $rc ;;
  bench)    tag=${1:-r3}
            python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
            python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; rc=$?
            cut -c1-600 gpurun_out/${tag}_bench.json; exit $rc ;;
  launches) tag=${1:-r3}
            python bench.py --steps 2 --warmup 1 --no-sub --no-cpu-baseline > gpurun_out/${tag}_plain.json 2>&1 || exit 1
            ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
                python bench.py --steps 2 --warmup 1 --no-sub --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu.log 2>&1 ;;
  ncu)      wl=${1:-c3}; tag=${2:-r3}
            python bench.py --workload $wl --spp 16 --steps 1 --warmup 1 --no-sub --no-cpu-baseline --no-e2e > gpurun_out/${tag}_${wl}_plain.json 2>&1 || exit 1
            ncu --set full --clock-control none --import-source on -k regex:k_walk_persistent -c 1 -o gpurun_out/${tag}_${wl} -f \
                python bench.py --workload $wl --spp 16 --steps 1 --warmup 1 --no-sub --no-cpu-baseline --no-e2e > gpurun_out/${tag}_${wl}_ncu.log 2>&1 ;;
  l2)       tag=${1:-r3}; tools/l2_peak 32 > gpurun_out/${tag}_l2_peak_32.json; tools/l2_peak 8 > gpurun_out/${tag}_l2_peak_8.json; cat gpurun_out/${tag}_l2_peak_32.json ;;
  *)        echo "usage: tools/gpu_run.sh tests|bench|launches|ncu <workload>|l2 [tag]"; exit 2 ;;
esac
