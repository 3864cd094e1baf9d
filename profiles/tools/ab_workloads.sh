# usage: ab_workloads.sh "<batch variants>" "<single/mixed variants>"
mkdir -p gpurun_out
pr() { python -c "import json,sys; d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$2', round(d['value'],1), round(d['ms_per_step'],2))"; }
for v in $1; do
  DEFL_B200_LIB=$PWD/build_variants/$v.so python bench.py --no-extra --no-e2e --no-cpu --steps 5 --warmup 3 > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  pr gpurun_out/ab_$v.json "batch $v"
done
for v in $2; do
  DEFL_B200_LIB=$PWD/build_variants/$v.so python bench.py --workload single --no-e2e --no-cpu --steps 3 --warmup 3 > gpurun_out/abs_$v.json 2> gpurun_out/abs_$v.err
  pr gpurun_out/abs_$v.json "single $v"
  DEFL_B200_LIB=$PWD/build_variants/$v.so python bench.py --workload mixed --mixed-gib 4 --no-e2e --no-cpu --steps 3 --warmup 3 > gpurun_out/abm_$v.json 2> gpurun_out/abm_$v.err
  pr gpurun_out/abm_$v.json "mixed $v"
done
