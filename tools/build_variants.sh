#!/bin/bash
# builds tuning variants of libartis_gpu.so into variants/ (git-ignored, travels to the GPU box): name=flags ...
#   bash tools/build_variants.sh k2="-DARTIS_KAPPA_CHUNK_TILES=2" h2="-DARTIS_HEAVY_MINBLOCKS=2"
cd "$(dirname "$0")/.."
mkdir -p variants
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  ( /usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared \
      -split-compile 4 -DARTIS_NO_VPKT $flags -I include artis_b200/csrc/transport.cu -o variants/libartis_gpu_$name.so -ldl \
      > variants/build_$name.log 2>&1 && echo "built $name" || echo "FAILED $name" ) &
done
wait
