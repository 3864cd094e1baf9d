# ncu --set full (with source) of the replay kernel (C pass) of one 1 GiB pigz-style stream; library variant in $1
mkdir -p gpurun_out
DEFL_B200_LIB=$PWD/build_variants/$1.so ncu --set full --clock-control none --import-source on -k regex:split_replay_kernel -s 1 -c 1 -f -o gpurun_out/ncu_replay \
  python profiles/tools/single_probe.py 1073741824 pigz > gpurun_out/ncu_replay.log 2>&1
tail -3 gpurun_out/ncu_replay.log
