#!/usr/bin/env bash
# Round-2 GPU call E (--gpus 2): NCCL inside libwost.so -- 2-rank parity tests, bench at N=2 (uniform headline + guided sub-workloads), multi-GPU elaina-exec
O=gpurun_out/r2e; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -15 $O/pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
NCCL_DEBUG=INFO timeout 900 $TR bench.py --gpus 2 --steps 3 --warmup 2 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"; grep -c "NCCL INFO" $O/bench_n2.err; grep -E "NVLS|Connected all|ncclCommInitRank.*nranks" $O/bench_n2.err | head -6; tail -c 300 $O/bench_n2.json
timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu-baseline > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?"
timeout 900 python tools/exec_multi_check.py --gpus 2 > $O/exec_multi.json 2> $O/exec_multi.err; echo "exec multi rc=$?"; cat $O/exec_multi.json | cut -c1-1500
