# bench the headline kernel (config 2, device resident) with each library variant named on the command line
# (build_variants/<name>.so)
mkdir -p gpurun_out
for v in "$@"; do
  DEFL_B200_LIB=$PWD/build_variants/$v.so python bench.py --no-extra --no-e2e --no-cpu --steps 5 --warmup 3 > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python -c "import json,sys; d=json.loads(open('gpurun_out/ab_$v.json').read().strip().splitlines()[-1]); print('$v', round(d['value'],1), round(d['ms_per_step'],2))"
done
if [ -n "$PARITY" ]; then timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_reference_callers.py -m gpu -x -q 2>&1 | tail -3; fi
