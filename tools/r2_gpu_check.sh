#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the bench lines of both arms on the named workload, the launch list.
#   gpurun --timeout 1500 -- bash tools/r2_gpu_check.sh [tag]
TAG=${1:-check}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt 2>&1
nproc >> $O/gpu.txt
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
( time python bench.py --impl reference ) > $O/bench_reference.json 2> $O/bench_reference.err
( time python bench.py ) > $O/bench.json 2> $O/bench.err
( time python bench.py --workload nebular ) > $O/bench_nebular.json 2> $O/bench_nebular.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
  python bench.py --steps 2 --warmup 1 --cpu-baseline off > $O/ncu_bench.log 2>&1
tail -3 $O/pytest_gpu.log; tail -2 $O/smoke.log; cut -c1-400 $O/bench.json
