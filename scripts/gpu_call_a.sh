#!/bin/bash
# round-2 GPU call A (1 GPU): tests, both bench arms, stock baselines, loss parity
mkdir -p gpurun_out/a
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/a/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/a/summary.txt
tail -3 gpurun_out/a/pytest_gpu.log | tee -a gpurun_out/a/summary.txt
timeout 300 python bench.py --gpus 1 --steps 64 --warmup 16 > gpurun_out/a/bench_own.json 2> gpurun_out/a/bench_own.err; echo "own rc=$?" | tee -a gpurun_out/a/summary.txt
tail -1 gpurun_out/a/bench_own.json | tee -a gpurun_out/a/summary.txt
timeout 400 python bench.py --impl reference --gpus 1 --steps 24 --warmup 8 > gpurun_out/a/bench_ref.json 2> gpurun_out/a/bench_ref.err; echo "ref rc=$?" | tee -a gpurun_out/a/summary.txt
tail -1 gpurun_out/a/bench_ref.json | tee -a gpurun_out/a/summary.txt
for m in fp32-eager bf16-eager fp32-graph bf16-graph; do
  timeout 300 python bench.py --impl ref-torch --ref-mode $m --gpus 1 --steps 64 --warmup 16 > gpurun_out/a/reftorch_$m.json 2> gpurun_out/a/reftorch_$m.err; echo "ref-torch $m rc=$?" | tee -a gpurun_out/a/summary.txt
  tail -1 gpurun_out/a/reftorch_$m.json | tee -a gpurun_out/a/summary.txt
done
timeout 400 python bench/loss_parity.py --impl native --gpus 1 --epochs 16 --steps-per-epoch 32 > gpurun_out/a/parity_native_1gpu.json 2> gpurun_out/a/parity_native.err; echo "parity native rc=$?" | tee -a gpurun_out/a/summary.txt
tail -1 gpurun_out/a/parity_native_1gpu.json | tee -a gpurun_out/a/summary.txt
timeout 400 python bench/loss_parity.py --impl ref-torch --gpus 1 --epochs 16 --steps-per-epoch 32 > gpurun_out/a/parity_reftorch_1gpu.json 2> gpurun_out/a/parity_ref.err; echo "parity ref rc=$?" | tee -a gpurun_out/a/summary.txt
tail -1 gpurun_out/a/parity_reftorch_1gpu.json | tee -a gpurun_out/a/summary.txt
tail -5 gpurun_out/a/*.err
