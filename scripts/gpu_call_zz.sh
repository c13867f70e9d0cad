#!/bin/bash
mkdir -p gpurun_out/zz
KUBEML_NVTX=1 timeout 200 python -m pytest tests/test_gpu_e2e.py -q > gpurun_out/zz/pytest_e2e_nvtx.log 2>&1; echo "e2e with NVTX ranges rc=$?" | tee gpurun_out/zz/summary.txt
tail -3 gpurun_out/zz/pytest_e2e_nvtx.log | cut -c1-200 | tee -a gpurun_out/zz/summary.txt
exit 0
