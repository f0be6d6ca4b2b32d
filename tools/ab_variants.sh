#!/bin/bash
# Same-box A/B of library variants under variants/*.so (box-to-box spread is larger than most kernel changes).
# usage: bash tools/ab_variants.sh ROUNDS name1 name2 ...   -> gpurun_out/ab_<name>_<round>.json
mkdir -p gpurun_out
rounds=$1; shift
for r in $(seq 1 $rounds); do
  for v in "$@"; do
    ARTIS_GPU_LIB=variants/$v.so python bench.py --steps 3 --warmup 3 --cpu-baseline off > gpurun_out/ab_${v}_$r.json 2> gpurun_out/ab_${v}_$r.err
  done
done
