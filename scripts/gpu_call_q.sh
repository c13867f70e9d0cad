#!/bin/bash
mkdir -p gpurun_out/q
O=gpurun_out/q
export KUBEML_FUSED_CAP_SCALE=2
timeout 120 python bench/probe_fused_plan.py 2>&1 | grep -E "layer|stem" | tee -a $O/summary.txt
timeout 300 python -m pytest tests/test_gpu_layers.py -x -q -k "fused or barrier" > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -3 $O/pytest.log | tee -a $O/summary.txt
timeout 300 python bench.py --gpus 1 --steps 128 --warmup 16 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/summary.txt
python -c "import json; d=json.loads(open('$O/bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['samples_ms_per_step'], d['gpu_launches'])" | tee -a $O/summary.txt
tail -3 $O/bench.err
timeout 120 python bench/step_timeline.py 50 | tee -a $O/summary.txt
exit 0
