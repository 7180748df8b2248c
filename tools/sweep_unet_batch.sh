#!/bin/bash
# U-Net forward throughput against the forward batch (micro-batching experiment): tools/sweep_unet_batch.sh
for m in ncsnpp beatgans; do
  for b in 64 128 256 512 1024; do
    timeout 200 python tools/bench_unet.py --model $m --batch $b --steps 3 --warmup 2 2>/dev/null | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['model'], d.get('batch'), round(d['ms_per_forward'],2), round(d['score_evals_per_s'],1))"
  done
done
