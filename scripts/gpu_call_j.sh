#!/bin/bash
# 1 GPU: VGG-16 (config 3) own vs stock, CIFAR ResNet-32, dynamic-parallelism smoke is multi-GPU (not here)
mkdir -p gpurun_out/j
O=gpurun_out/j
timeout 400 python bench.py --model vgg16 --gpus 1 --steps 64 --warmup 16 > $O/vgg_own.json 2> $O/vgg_own.err; echo "vgg own rc=$?" | tee -a $O/summary.txt
tail -1 $O/vgg_own.json | cut -c1-700 | tee -a $O/summary.txt
for m in fp32-eager bf16-graph; do
  timeout 400 python bench.py --model vgg16 --impl ref-torch --ref-mode $m --gpus 1 --steps 64 --warmup 16 > $O/vgg_reftorch_$m.json 2> $O/vgg_reftorch_$m.err; echo "vgg ref-torch $m rc=$?" | tee -a $O/summary.txt
  tail -1 $O/vgg_reftorch_$m.json | cut -c1-500 | tee -a $O/summary.txt
done
timeout 300 python bench/cifar_resnet_bench.py > $O/cifar_resnet32.txt 2> $O/cifar.err; echo "cifar rc=$?" | tee -a $O/summary.txt
tail -5 $O/cifar_resnet32.txt | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; tail -n 5 $f; done
exit 0
