# phase times of the split decoder on the 1 GiB pigz-style stream for a few search granules (library variant in $1)
for g in ${2:-32768 49152 65536}; do
  echo "== granule $g"
  DEFL_B200_LIB=$PWD/build_variants/$1.so DEFL_SPLIT_GRANULE=$g DEFL_SPLIT_TRACE=1 python profiles/tools/single_probe.py 1073741824 pigz 2>&1 | tail -7 | grep -v "batch:"
done
