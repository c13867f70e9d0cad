#!/bin/bash
# 1 GPU: partial deferral of the round's last SGD step (n_fuse): tests, full suite, bench A/B
mkdir -p gpurun_out/v
O=gpurun_out/v
timeout 600 python -m pytest tests/test_gpu_native.py tests/test_gpu_e2e.py -q -k "fuses_only_the_head or partial_deferral or fused" > $O/pytest_new.log 2>&1; echo "new tests rc=$?" | tee -a $O/summary.txt
tail -30 $O/pytest_new.log | cut -c1-400 | tee -a $O/summary.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -5 $O/pytest_gpu.log | cut -c1-300 | tee -a $O/summary.txt
show() {
  python - "$1" <<'PY' | tee -a gpurun_out/v/summary.txt
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print({k: d.get(k) for k in ("value", "ms_per_step", "merge_ms_per_sync", "gpu_launches", "samples_ms_per_step")}, d["config"]["fused_last_sgd_steps"], d["e2e"]["value"], d["clocks"])
PY
}
timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_own_1gpu.json 2> $O/bench_own_1gpu.err; echo "own rc=$?" | tee -a $O/summary.txt
show $O/bench_own_1gpu.json
KUBEML_PARTIAL_DEFER=0 timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_fulldefer_1gpu.json 2> $O/bench_fulldefer_1gpu.err; echo "fulldefer rc=$?" | tee -a $O/summary.txt
show $O/bench_fulldefer_1gpu.json
KUBEML_FUSE_LAST_STEP=0 timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_nofuse_1gpu.json 2> $O/bench_nofuse_1gpu.err; echo "nofuse rc=$?" | tee -a $O/summary.txt
show $O/bench_nofuse_1gpu.json
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
