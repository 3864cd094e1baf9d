# mixed workload (8 GiB) for a few split thresholds (DEFL_SPLIT_TAIL_PCT: share of 1 / resident warps from which a member is split)
for t in ${1:-40 50 60}; do
  DEFL_SPLIT_TAIL_PCT=$t DEFL_SPLIT_TRACE=1 python bench.py --workload mixed --mixed-gib 8 --no-e2e --no-cpu --steps 3 --warmup 2 > gpurun_out/ts_$t.json 2> gpurun_out/ts_$t.err
  python -c "import json; d=json.loads(open('gpurun_out/ts_$t.json').read().strip().splitlines()[-1]); print('tail pct $t', round(d['value'],1), round(d['ms_per_step'],2))"
  grep "batch:" gpurun_out/ts_$t.err | tail -1
done
