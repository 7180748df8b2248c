#!/bin/bash
# Round-2 evidence run on one B200 (under gpurun): launch lists + ncu --set full captures with tensor-pipe counters.
# Every profiled command is first run plain (must exit 0); numbers printed under ncu are never bench values.
set -u
O=gpurun_out
FCN="python bench.py --extra 0 --strong-anchors 0 --no-cpu-baseline --steps 1 --warmup 1"
TM="sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor_subpipe_hmma.sum,sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_elapsed"
$FCN > $O/r2_fcn_plain.log 2>&1 || { echo "plain FCN bench failed"; tail -5 $O/r2_fcn_plain.log; exit 1; }
# launch list of one warm-up + one timed step (+ the e2e steps)
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2_fcn_launches.csv $FCN > $O/r2_fcn_ncu.log 2>&1
# gemm launches: warm-up step = 7 anchor + 2 x 7 chain = 21, then 7 anchor layers of the timed step -> launch 28 = first layer, 29 = hidden
ncu --set full --metrics $TM --clock-control none --import-source on -k regex:gemm_bf16x3_kernel -s 28 -c 2 -o $O/r2_gemm $FCN > $O/r2_gemm_ncu.log 2>&1
ncu --set full --metrics $TM --clock-control none --import-source on -k regex:"center_gram|tridiag_fused_small|bisect_kernel|perturb_kernel" -s 4 -c 4 -o $O/r2_fcn_tail $FCN > $O/r2_tail_ncu.log 2>&1
# U-Net (NCSN++ forward, batch 512): launch list + the convolution / prep / attention kernels
UN="python tools/bench_unet.py --model ncsnpp --batch 512 --steps 1 --warmup 1"
$UN > $O/r2_unet_plain.log 2>&1 || { echo "plain U-Net bench failed"; tail -5 $O/r2_unet_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 800 --csv --log-file $O/r2_unet_launches.csv $UN > $O/r2_unet_ncu.log 2>&1
ncu --set full --metrics $TM --clock-control none --import-source on -k regex:conv_fp16x3_kernel -s 170 -c 12 -o $O/r2_conv $UN > $O/r2_conv_ncu.log 2>&1
ncu --set full --metrics $TM --clock-control none -k regex:"conv_prep|attention_kernel|gn_from_partials" -s 300 -c 8 -o $O/r2_unet_misc $UN > $O/r2_misc_ncu.log 2>&1
# the reports themselves exceed what gpurun copies back: keep their raw (all metrics per launch) and source pages as CSV
for r in r2_gemm r2_fcn_tail r2_conv r2_unet_misc; do
  [ -f $O/$r.ncu-rep ] || continue
  ncu -i $O/$r.ncu-rep --page raw --csv > $O/${r}_raw.csv 2>/dev/null
  case $r in r2_gemm|r2_conv) ncu -i $O/$r.ncu-rep --page source --csv > $O/${r}_source.csv 2>/dev/null; gzip -f $O/${r}_source.csv;; esac
  rm -f $O/$r.ncu-rep
done
ls -la $O/
