#!/bin/bash
# round-2 GPU call F (8 GPUs): the numbers that need the whole box
N=${1:-8}
mkdir -p gpurun_out/f$N
O=gpurun_out/f$N
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -x -q > $O/pytest_multigpu.log 2>&1; echo "multigpu tests rc=$?" | tee -a $O/summary.txt
tail -4 $O/pytest_multigpu.log | tee -a $O/summary.txt
timeout 600 $TR --master-port 29531 bench/merge_bench.py --iters 20 > $O/merge_bench.jsonl 2> $O/merge_bench.err; echo "merge bench rc=$?" | tee -a $O/summary.txt
cat $O/merge_bench.jsonl | tee -a $O/summary.txt
timeout 600 $TR --master-port 29517 bench.py --gpus $N --steps 128 --warmup 16 > $O/bench_own.json 2> $O/bench_own.err; echo "own rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_own.json | tee -a $O/summary.txt
timeout 900 $TR --master-port 29519 bench.py --impl reference --gpus $N --steps 24 --warmup 8 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" | tee -a $O/summary.txt
tail -1 $O/bench_ref.json | tee -a $O/summary.txt
timeout 600 $TR --master-port 29521 bench.py --impl ref-torch --ref-mode bf16-graph --gpus $N --steps 128 --warmup 16 > $O/reftorch_bf16graph.json 2> $O/reftorch.err; echo "ref-torch rc=$?" | tee -a $O/summary.txt
tail -1 $O/reftorch_bf16graph.json | tee -a $O/summary.txt
timeout 600 $TR --master-port 29523 bench.py --impl ddp --ref-mode bf16-eager --gpus $N --steps 64 --warmup 16 > $O/ddp_bf16.json 2> $O/ddp.err; echo "ddp rc=$?" | tee -a $O/summary.txt
tail -1 $O/ddp_bf16.json | tee -a $O/summary.txt
timeout 600 $TR --master-port 29533 bench/loss_parity.py --impl native --gpus $N --epochs 16 --steps-per-epoch 32 > $O/parity_native.json 2> $O/parity_native.err; echo "parity native rc=$?" | tee -a $O/summary.txt
tail -1 $O/parity_native.json | tee -a $O/summary.txt
timeout 600 $TR --master-port 29535 bench/loss_parity.py --impl ref-torch --gpus $N --epochs 16 --steps-per-epoch 32 > $O/parity_reftorch.json 2> $O/parity_ref.err; echo "parity ref rc=$?" | tee -a $O/summary.txt
tail -1 $O/parity_reftorch.json | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
