#!/bin/bash
mkdir -p gpurun_out/x
O=gpurun_out/x
timeout 400 python bench/probe_partial_defer.py 0.0001 > $O/probe.json 2> $O/probe.err; echo "probe rc=$?" | tee -a $O/summary.txt
tail -1 $O/probe.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for n,i in d['runs'].items(): print(n, i)
for k,row in d['relative_difference_of_movement'].items():
    print(k, 'norms', d['movement_norm'][k]); print('   ', ' '.join(f'{p}={v}' for p,v in row.items()))
" | tee -a $O/summary.txt
grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $O/probe.err | tail -5
exit 0
