#!/bin/bash
# 1 GPU, last call of the round: whole GPU suite and the default bench line on the final tree
mkdir -p gpurun_out/z
O=gpurun_out/z
timeout 600 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -4 $O/pytest_gpu.log | cut -c1-300 | tee -a $O/summary.txt
timeout 300 python bench.py > $O/bench_own_1gpu.json 2> $O/bench_own_1gpu.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own_1gpu.json | cut -c1-1400 | tee -a $O/summary.txt
exit 0
