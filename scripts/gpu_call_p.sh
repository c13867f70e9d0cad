#!/bin/bash
# 1 GPU: full GPU suite, bench with the round look-ahead / kernel-posted loss / stable batch buffer, and each switched off
mkdir -p gpurun_out/p
O=gpurun_out/p
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -6 $O/pytest_gpu.log | tee -a $O/summary.txt
run() { # name, env...
  name=$1; shift
  env KUBEML_BENCH_GAPS=1 "$@" timeout 600 python bench.py --steps 128 --warmup 16 > $O/bench_$name.json 2> $O/bench_$name.err; echo "$name rc=$?" | tee -a $O/summary.txt
  python - "$O/bench_$name.json" <<'PY' | tee -a $O/summary.txt
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print({k: d.get(k) for k in ("value", "ms_per_step", "merge_ms_per_sync", "gpu_launches", "samples_ms_per_step", "round_timeline_us")}, d["e2e"]["value"], d["clocks"])
PY
}
run own A=1
run nolookahead KUBEML_ROUND_LOOKAHEAD=0
run losscopy KUBEML_LOSS_POST=copy
run nostable KUBEML_STABLE_BATCH=0
for f in $O/*.err; do echo "== $f"; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 6; done
exit 0
