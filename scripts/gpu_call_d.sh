#!/bin/bash
# per-kernel lists (ncu, one metric): round-1 tree vs this tree unfused vs fused
mkdir -p gpurun_out/d
O=$PWD/gpurun_out/d
list() { # name dir env
  name=$1; dir=$2; shift 2
  ( cd $dir && env "$@" timeout 300 python bench/step_kernels.py resnet34 3 > $O/$name.plain.log 2>&1 && \
    env "$@" timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/$name.csv python bench/step_kernels.py resnet34 3 > $O/$name.ncu.log 2>&1 ; \
    python bench/step_kernels.py --summarize $O/$name.csv 3 > $O/$name.txt 2>&1 )
  echo "== $name"; head -12 $O/$name.txt
}
list r1 build/wt_r1 X=1
list new_unfused . KUBEML_FUSE_BN=0
list new_fused . KUBEML_FUSE_BN=1
exit 0
