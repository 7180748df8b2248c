# Headline bench at several device-batch sizes (anchors per chunk): prints value, ms/step, stage times, SM clock.
for c in 128 148 256 296 512; do
  python bench.py --no-cpu-baseline --extra 0 --strong-anchors 0 --chunk $c 2>/dev/null | C=$c python -c "import sys,json,os; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(os.environ['C'], round(d['value']), round(d['ms_per_step'],1), {k: round(v,1) for k,v in d['stages_ms_per_step'].items()}, d['clocks']['sm_mhz'])"
done
