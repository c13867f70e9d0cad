#!/bin/bash
# round-2 GPU call B (1 GPU): fused conv+BN kernels - numerics, full GPU suite, bench A/B, kernel list
mkdir -p gpurun_out/b
O=gpurun_out/b
timeout 600 python -m pytest tests/test_gpu_layers.py -x -q -k "fused" > $O/pytest_fused.log 2>&1; echo "fused tests rc=$?" | tee -a $O/summary.txt
tail -15 $O/pytest_fused.log | tee -a $O/summary.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -5 $O/pytest_gpu.log | tee -a $O/summary.txt
timeout 300 python bench.py --gpus 1 --steps 64 --warmup 16 > $O/bench_fused.json 2> $O/bench_fused.err; echo "bench fused rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_fused.json | tee -a $O/summary.txt
KUBEML_FUSE_BN=0 timeout 300 python bench.py --gpus 1 --steps 64 --warmup 16 > $O/bench_unfused.json 2> $O/bench_unfused.err; echo "bench unfused rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_unfused.json | tee -a $O/summary.txt
timeout 300 python bench/step_kernels.py resnet34 3 > $O/step_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/k.csv python bench/step_kernels.py resnet34 3 > $O/ncu.log 2>&1
python bench/step_kernels.py --summarize $O/k.csv 3 > $O/resnet34_step_kernels.txt 2>&1
head -30 $O/resnet34_step_kernels.txt | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; tail -n 5 $f; done
exit 0
