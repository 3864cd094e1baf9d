# the round's evidence in one call: GPU tests, the bench line, the launch list and one ncu --set full capture of the
# batch kernel at the bench launch size (each profiler pass only after the plain run exited 0)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/re_gputests.txt 2>&1; tail -3 gpurun_out/re_gputests.txt
python bench.py > gpurun_out/re_bench.json 2> gpurun_out/re_bench.err || exit 1
tail -c 400 gpurun_out/re_bench.json
python bench.py --profile --steps 2 --warmup 3 > gpurun_out/re_profile_plain.json 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/re_launches.csv \
  python bench.py --profile --steps 2 --warmup 3 > gpurun_out/re_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:inflate_warp_kernel -s 3 -c 1 -f -o gpurun_out/re_ncu_full \
  python bench.py --profile --steps 2 --warmup 3 > gpurun_out/re_ncu_full.log 2>&1
ls -la gpurun_out | tail -8
