#!/bin/bash
# ncu evidence for the panel tridiagonalisation kernel (under gpurun, one GPU): plain run first, then one --set full capture
# of tridiag_panel_kernel at n = 12288 (the reference's own K = 16768 on 64x64x3 images) and n = 4096; raw pages kept as CSV.
set -u
O=gpurun_out
for n in 4096 12288; do
  python tools/bench_spectrum_variant.py $n 3 2 > $O/r2_eig_panel_plain_$n.json 2> $O/r2_eig_panel_plain_$n.err || { echo "plain run failed ($n)"; exit 1; }
  cat $O/r2_eig_panel_plain_$n.json
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:tridiag_panel_kernel -c 1 -o $O/r2_eig_panel_$n \
      python tools/bench_spectrum_variant.py $n 3 1 > $O/r2_eig_panel_ncu_$n.log 2>&1
  [ -f $O/r2_eig_panel_$n.ncu-rep ] || { echo "no report ($n)"; tail -5 $O/r2_eig_panel_ncu_$n.log; continue; }
  ncu -i $O/r2_eig_panel_$n.ncu-rep --page raw --csv > $O/r2_eig_panel_${n}_raw.csv 2>/dev/null
  ncu -i $O/r2_eig_panel_$n.ncu-rep --page source --csv > $O/r2_eig_panel_${n}_source.csv 2>/dev/null
  gzip -f $O/r2_eig_panel_${n}_source.csv
  rm -f $O/r2_eig_panel_$n.ncu-rep
done
ls -la $O/ | grep eig_panel
