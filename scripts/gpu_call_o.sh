#!/bin/bash
# 1 GPU: full GPU suite (no -x), bench with the per-round gap table, A/B of the round prefetch
mkdir -p gpurun_out/o
O=gpurun_out/o
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -8 $O/pytest_gpu.log | tee -a $O/summary.txt
KUBEML_BENCH_GAPS=1 timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | cut -c1-1800 | tee -a $O/summary.txt
KUBEML_ROUND_PREFETCH=0 KUBEML_BENCH_GAPS=1 timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_noprefetch.json 2> $O/bench_noprefetch.err; echo "noprefetch rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_noprefetch.json | cut -c1-1800 | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
