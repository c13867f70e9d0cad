#!/bin/bash
# final multi-GPU call: tests, bench, dynamic parallelism, merge sweep (N GPUs)
N=${1:-8}
mkdir -p gpurun_out/m$N
O=gpurun_out/m$N
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -x -q > $O/pytest_multigpu.log 2>&1; echo "multigpu tests rc=$?" | tee -a $O/summary.txt
tail -4 $O/pytest_multigpu.log | tee -a $O/summary.txt
timeout 600 $TR --master-port 29517 bench.py --gpus $N --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | cut -c1-1300 | tee -a $O/summary.txt
timeout 600 $TR --master-port 29533 bench/dynamic_bench.py 10 53760 > $O/dynamic.json 2> $O/dynamic.err; echo "dynamic rc=$?" | tee -a $O/summary.txt
tail -1 $O/dynamic.json | tee -a $O/summary.txt
timeout 600 $TR --master-port 29531 bench/merge_bench.py --iters 30 --sizes lenet,resnet34 > $O/merge_bench.jsonl 2> $O/merge_bench.err; echo "merge bench rc=$?" | tee -a $O/summary.txt
cat $O/merge_bench.jsonl | tee -a $O/summary.txt
timeout 600 $TR --master-port 29519 bench.py --model vgg16 --gpus $N --steps 64 --warmup 16 > $O/bench_vgg.json 2> $O/bench_vgg.err; echo "vgg rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_vgg.json | cut -c1-700 | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
