#!/bin/bash
# 1 GPU: the partial-deferral end-to-end test on weight movement, then the whole GPU suite
mkdir -p gpurun_out/w
O=gpurun_out/w
timeout 600 python -m pytest tests/test_gpu_e2e.py -q -k "partial_deferral" > $O/pytest_new.log 2>&1; echo "new test rc=$?" | tee -a $O/summary.txt
tail -25 $O/pytest_new.log | cut -c1-400 | tee -a $O/summary.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -5 $O/pytest_gpu.log | cut -c1-300 | tee -a $O/summary.txt
exit 0
