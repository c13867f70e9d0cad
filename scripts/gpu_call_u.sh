#!/bin/bash
# 2 GPUs: the threaded two-GPU test first, then the full GPU suite, then a 1-GPU bench of the final tree
mkdir -p gpurun_out/u
O=gpurun_out/u
timeout 300 python -m pytest tests/test_multigpu.py -q -k "worker_threads" > $O/pytest_threaded.log 2>&1; echo "threaded rc=$?" | tee -a $O/summary.txt
tail -25 $O/pytest_threaded.log | cut -c1-300 | tee -a $O/summary.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -6 $O/pytest_gpu.log | cut -c1-300 | tee -a $O/summary.txt
timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_own_1gpu.json 2> $O/bench_own_1gpu.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own_1gpu.json | cut -c1-1500 | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
