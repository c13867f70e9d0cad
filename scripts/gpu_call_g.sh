#!/bin/bash
# 2-GPU call: merge fixed cost, new 1-GPU tests (strided conv, trace, resume), bench
mkdir -p gpurun_out/g
O=gpurun_out/g
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench/merge_bench.py --check --iters 30 --sizes lenet,1M,resnet34 > $O/merge_bench.jsonl 2> $O/merge_bench.err; echo "merge bench rc=$?" | tee -a $O/summary.txt
cat $O/merge_bench.jsonl | cut -c1-420 | tee -a $O/summary.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -6 $O/pytest_gpu.log | tee -a $O/summary.txt
timeout 300 python bench.py --gpus 1 --steps 128 --warmup 16 > $O/bench_1gpu.json 2> $O/bench_1gpu.err; echo "bench rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_1gpu.json | cut -c1-600 | tee -a $O/summary.txt
KUBEML_FUSE_LAST_STEP=0 timeout 300 python bench.py --gpus 1 --steps 128 --warmup 16 > $O/bench_1gpu_nolast.json 2> $O/bench_1gpu_nolast.err
tail -1 $O/bench_1gpu_nolast.json | cut -c1-300 | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
