#!/bin/bash
# 1 GPU, final: full GPU suite, ResNet-34 and VGG-16 bench, ncu --set full of the hot kernels of the current build
mkdir -p gpurun_out/s
O=gpurun_out/s
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -5 $O/pytest_gpu.log | tee -a $O/summary.txt
show() {
  python - "$1" <<'PY' | tee -a gpurun_out/s/summary.txt
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print({k: d.get(k) for k in ("metric", "n_gpus", "value", "ms_per_step", "merge_ms_per_sync", "gpu_launches", "samples_ms_per_step", "round_timeline_us")}, d["e2e"]["value"], d["clocks"])
PY
}
KUBEML_BENCH_GAPS=1 timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_own_1gpu.json 2> $O/bench_own_1gpu.err; echo "own rc=$?" | tee -a $O/summary.txt
show $O/bench_own_1gpu.json
timeout 600 python bench.py --model vgg16 --steps 64 --warmup 16 > $O/bench_vgg16_1gpu.json 2> $O/bench_vgg16_1gpu.err; echo "vgg rc=$?" | tee -a $O/summary.txt
show $O/bench_vgg16_1gpu.json
timeout 300 python bench/ncu_target.py 1 > $O/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_tcgen05|kavg" -c 12 -o /tmp/prof_hot python bench/ncu_target.py 1 > $O/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $O/summary.txt
ncu -i /tmp/prof_hot.ncu-rep --page raw --csv > $O/prof_hot_raw.csv 2> $O/raw.err
wc -c $O/*.csv | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
