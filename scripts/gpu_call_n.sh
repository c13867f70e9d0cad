#!/bin/bash
# 2 GPUs: full GPU test suite (new torch-oracle tests), bench at N=2, experiment grid through the harness
mkdir -p gpurun_out/n
O=gpurun_out/n
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -6 $O/pytest_gpu.log | tee -a $O/summary.txt
timeout 600 $TR --master-port 29517 bench.py --gpus 2 --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | cut -c1-1300 | tee -a $O/summary.txt
timeout 900 $TR --master-port 29537 bench/grid_experiment.py 8 30 > $O/grid.json 2> $O/grid.err; echo "grid rc=$?" | tee -a $O/summary.txt
tail -1 $O/grid.json | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 8; done
exit 0
