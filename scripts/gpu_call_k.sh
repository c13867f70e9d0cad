#!/bin/bash
# 4 GPUs: multi-GPU tests incl. two concurrent jobs, bench at N=4, dynamic parallelism 2->4 with pre-warmed joins
N=4
mkdir -p gpurun_out/k
O=gpurun_out/k
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -x -q > $O/pytest_multigpu.log 2>&1; echo "multigpu tests rc=$?" | tee -a $O/summary.txt
tail -6 $O/pytest_multigpu.log | tee -a $O/summary.txt
timeout 600 $TR --master-port 29517 bench.py --gpus $N --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | cut -c1-900 | tee -a $O/summary.txt
timeout 600 $TR --master-port 29533 bench/dynamic_bench.py 6 30720 > $O/dynamic.json 2> $O/dynamic.err; echo "dynamic rc=$?" | tee -a $O/summary.txt
tail -1 $O/dynamic.json | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 8; done
exit 0
