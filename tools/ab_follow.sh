#!/bin/bash
# one-off A/B: gated role-following against the committed build, both workloads, plus the parity tests on the candidate
mkdir -p gpurun_out
bash tools/ab_variants.sh 2 base f1 f4 f16
for v in base f1 f4 f16; do
  ARTIS_GPU_LIB=variants/$v.so python bench.py --workload nebular --steps 5 --warmup 3 --cpu-baseline off > gpurun_out/ab_neb_$v.json 2> gpurun_out/ab_neb_$v.err
done
cp variants/f4.so artis_b200/libartis_gpu.so
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_follow.log 2>&1
