#!/bin/bash
# round-2 GPU call C (1 GPU): A/B of the step on one box - round-1 tree vs this tree with the switches
mkdir -p gpurun_out/c
O=gpurun_out/c
B="--gpus 1 --steps 128 --warmup 16"
run() { # name, dir, env...
  name=$1; dir=$2; shift 2
  ( cd $dir && env "$@" MASTER_PORT=0 timeout 300 python bench.py $B ) > $O/$name.json 2> $O/$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.loads(open('$O/$name.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d.get('merge_ms_per_sync'), d['gpu_launches'], d['config'].get('fused_last_sgd_steps'))" 2>&1)" | tee -a $O/summary.txt
}
run r1_tree build/wt_r1 X=1
run new_nofuse_nolast . KUBEML_FUSE_BN=0 KUBEML_FUSE_LAST_STEP=0
run new_nofuse_last . KUBEML_FUSE_BN=0 KUBEML_FUSE_LAST_STEP=1
run new_fuse_nolast . KUBEML_FUSE_BN=1 KUBEML_FUSE_LAST_STEP=0
run new_fuse_last . KUBEML_FUSE_BN=1 KUBEML_FUSE_LAST_STEP=1
run r1_tree_again build/wt_r1 X=1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt
tail -5 $O/pytest_gpu.log | tee -a $O/summary.txt
for f in $O/*.err; do echo "== $f"; tail -n 3 $f; done
exit 0
