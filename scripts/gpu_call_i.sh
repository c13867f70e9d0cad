#!/bin/bash
# quick 1-GPU check: fused tests + bench
mkdir -p gpurun_out/i
O=gpurun_out/i
timeout 900 python -m pytest tests/test_gpu_layers.py tests/test_gpu_native.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -4 $O/pytest.log | tee -a $O/summary.txt
for v in "A=1" "KUBEML_FUSE_LAST_STEP=0" ${EXTRA_VARIANTS}; do
  env $v timeout 300 python bench.py --gpus 1 --steps 128 --warmup 16 > $O/bench_$v.json 2> $O/bench_$v.err
  echo "$v: $(python -c "import json; d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['samples_ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'])" 2>&1)" | tee -a $O/summary.txt
done
for f in $O/*.err; do echo "== $f"; tail -n 4 $f; done
exit 0
