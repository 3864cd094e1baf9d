# ncu --set full (with source) of one launch of inflate_warp_kernel for each library variant named on the command line
set -x
mkdir -p gpurun_out
for v in "$@"; do
  DEFL_B200_LIB=$PWD/build_variants/$v.so ncu --set full --clock-control none --import-source on \
    -k regex:inflate_warp_kernel -s 2 -c 1 -f -o gpurun_out/ncu_$v \
    python bench.py --profile --streams 16000 --unique 2048 --steps 2 --warmup 1 > gpurun_out/ncu_$v.log 2>&1
  tail -2 gpurun_out/ncu_$v.log
done
ls -la gpurun_out/
