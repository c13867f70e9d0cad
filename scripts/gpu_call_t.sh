#!/bin/bash
# 8 GPUs, lean: bench at N=8 with the gap table, the 8-GPU-only tests, dynamic parallelism 2 -> 8
N=8
mkdir -p gpurun_out/t
O=gpurun_out/t
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
KUBEML_BENCH_GAPS=1 timeout 600 $TR --master-port 29517 bench.py --gpus $N --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | cut -c1-1800 | tee -a $O/summary.txt
timeout 600 python -m pytest tests/test_multigpu.py -q -k "failed_worker and 8 or two_jobs or stalled" > $O/pytest_multigpu.log 2>&1; echo "multigpu tests rc=$?" | tee -a $O/summary.txt
tail -4 $O/pytest_multigpu.log | tee -a $O/summary.txt
timeout 600 $TR --master-port 29533 bench/dynamic_bench.py 10 53760 > $O/dynamic.json 2> $O/dynamic.err; echo "dynamic rc=$?" | tee -a $O/summary.txt
tail -1 $O/dynamic.json | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
