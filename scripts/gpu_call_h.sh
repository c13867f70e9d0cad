#!/bin/bash
# ncu --set full of the hot kernels (one short program, 1 GPU); the report is summarised on the box (the
# .ncu-rep itself is too large to bring back) into raw / source CSV pages
mkdir -p gpurun_out/h
O=gpurun_out/h
rm -f $O/*.ncu-rep
timeout 300 python bench/ncu_target.py 1 > $O/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"gemm_tcgen05|gemm2sm|kavg|bn_bwd_kernel|sgd_kernel" -c 16 -o /tmp/prof_hot python bench/ncu_target.py 1 > $O/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $O/summary.txt
ncu -i /tmp/prof_hot.ncu-rep --page raw --csv > $O/prof_hot_raw.csv 2> $O/raw.err
ncu -i /tmp/prof_hot.ncu-rep --page source --csv --kernel-name regex:"gemm_tcgen05_kernel" > $O/prof_hot_source_gemm.csv 2> $O/src.err
ls -la $O /tmp/prof_hot.ncu-rep
wc -c $O/*.csv
exit 0
