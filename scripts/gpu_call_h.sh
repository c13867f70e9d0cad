#!/bin/bash
# ncu --set full of the hot kernels (one short program, 1 GPU) + launch list of the current step
mkdir -p gpurun_out/h
O=gpurun_out/h
timeout 300 python bench/ncu_target.py 1 > $O/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -c 60 -o $O/prof_hot python bench/ncu_target.py 1 > $O/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $O/summary.txt
timeout 300 python bench/step_kernels.py resnet34 3 > $O/step_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/k.csv python bench/step_kernels.py resnet34 3 > $O/ncu_list.log 2>&1
python bench/step_kernels.py --summarize $O/k.csv 3 > $O/resnet34_step_kernels.txt 2>&1
head -30 $O/resnet34_step_kernels.txt | tee -a $O/summary.txt
ls -la $O
exit 0
