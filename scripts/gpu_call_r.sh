#!/bin/bash
# 2 GPUs: full GPU suite incl. the 2-GPU tests, bench at N=1 and N=2 with the gap tables, experiment grid
mkdir -p gpurun_out/r
O=gpurun_out/r
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -6 $O/pytest_gpu.log | tee -a $O/summary.txt
show() {
  python - "$1" <<'PY' | tee -a gpurun_out/r/summary.txt
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print({k: d.get(k) for k in ("n_gpus", "value", "ms_per_step", "merge_ms_per_sync", "merge_wait_ms_per_sync", "gpu_launches", "samples_ms_per_step", "round_timeline_us")}, d["e2e"]["value"], d["clocks"])
PY
}
KUBEML_BENCH_GAPS=1 timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_own_1gpu.json 2> $O/bench_own_1gpu.err; echo "own1 rc=$?" | tee -a $O/summary.txt
show $O/bench_own_1gpu.json
KUBEML_BENCH_GAPS=1 timeout 600 $TR --master-port 29517 bench.py --gpus 2 --steps 128 --warmup 16 > $O/bench_own_2gpu.json 2> $O/bench_own_2gpu.err; echo "own2 rc=$?" | tee -a $O/summary.txt
show $O/bench_own_2gpu.json
KUBEML_ROUND_LOOKAHEAD=0 timeout 600 $TR --master-port 29519 bench.py --gpus 2 --steps 128 --warmup 16 > $O/bench_nolookahead_2gpu.json 2> $O/bench_nolookahead_2gpu.err; echo "nolookahead2 rc=$?" | tee -a $O/summary.txt
show $O/bench_nolookahead_2gpu.json
timeout 900 $TR --master-port 29537 bench/grid_experiment.py 8 30 > $O/grid.json 2> $O/grid.err; echo "grid rc=$?" | tee -a $O/summary.txt
tail -1 $O/grid.json | cut -c1-3000 | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 8; done
exit 0
