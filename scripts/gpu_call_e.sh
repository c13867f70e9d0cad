#!/bin/bash
# round-2 GPU call E (N GPUs): multi-GPU tests, merge sweep, bench own/reference/stock at N
N=${1:-2}
mkdir -p gpurun_out/e$N
O=gpurun_out/e$N
timeout 1200 python -m pytest tests/test_multigpu.py -x -q > $O/pytest_multigpu.log 2>&1; echo "multigpu tests rc=$?" | tee -a $O/summary.txt
tail -15 $O/pytest_multigpu.log | tee -a $O/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench/merge_bench.py --iters 20 > $O/merge_bench.jsonl 2> $O/merge_bench.err; echo "merge bench rc=$?" | tee -a $O/summary.txt
cat $O/merge_bench.jsonl | tee -a $O/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; tail -n 8 $f; done
exit 0
