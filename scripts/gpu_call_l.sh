#!/bin/bash
mkdir -p gpurun_out/l
O=gpurun_out/l
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -5 $O/pytest_gpu.log | tee -a $O/summary.txt
timeout 300 python bench.py --gpus 1 --steps 128 --warmup 16 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench.json | cut -c1-1200 | tee -a $O/summary.txt
timeout 400 python bench.py --model vgg16 --gpus 1 --steps 64 --warmup 16 > $O/vgg_own.json 2> $O/vgg_own.err; echo "vgg own rc=$?" | tee -a $O/summary.txt
tail -1 $O/vgg_own.json | cut -c1-900 | tee -a $O/summary.txt
KUBEML_FUSE_LAST_STEP=0 timeout 400 python bench.py --model vgg16 --gpus 1 --steps 64 --warmup 16 > $O/vgg_nolast.json 2> $O/vgg_nolast.err
tail -1 $O/vgg_nolast.json | cut -c1-300 | tee -a $O/summary.txt
timeout 300 python bench/cifar_resnet_bench.py 2>&1 | tail -2 | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; tail -n 5 $f; done
exit 0
