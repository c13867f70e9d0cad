#!/bin/bash
# 2 GPUs: the partial deferral on the multi-rank merge path: bench at N=2, two multi-GPU tests
mkdir -p gpurun_out/y
O=gpurun_out/y
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29517 bench.py --gpus 2 --steps 128 --warmup 16 > $O/bench_own_2gpu.json 2> $O/bench_own_2gpu.err; echo "own2 rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own_2gpu.json | cut -c1-1500 | tee -a $O/summary.txt
timeout 300 python -m pytest tests/test_multigpu.py -q -k "failed_worker and 2 or elastic" > $O/pytest_multigpu.log 2>&1; echo "multigpu rc=$?" | tee -a $O/summary.txt
tail -4 $O/pytest_multigpu.log | cut -c1-300 | tee -a $O/summary.txt
grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $O/bench_own_2gpu.err | tail -5
exit 0
