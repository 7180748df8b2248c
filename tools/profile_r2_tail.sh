#!/bin/bash
# Refresh of the FCN launch list / share and the tail-kernel ncu capture after the per-anchor solver tuning (under gpurun).
set -u
O=gpurun_out
FCN="python bench.py --extra 0 --strong-anchors 0 --no-cpu-baseline --steps 1 --warmup 1"
$FCN > $O/r2b_fcn_plain.log 2>&1 || { echo "plain FCN bench failed"; tail -5 $O/r2b_fcn_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2b_fcn_launches.csv $FCN > $O/r2b_fcn_ncu.log 2>&1
ncu --set full --clock-control none -k regex:"center_gram|tridiag_fused_small|bisect_kernel|perturb_kernel" -s 4 -c 4 -o $O/r2b_fcn_tail $FCN > $O/r2b_tail_ncu.log 2>&1
[ -f $O/r2b_fcn_tail.ncu-rep ] && ncu -i $O/r2b_fcn_tail.ncu-rep --page raw --csv > $O/r2b_fcn_tail_raw.csv 2>/dev/null
rm -f $O/r2b_fcn_tail.ncu-rep
ls -la $O | grep r2b
