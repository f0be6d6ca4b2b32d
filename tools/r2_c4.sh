#!/bin/bash
# One gpurun call for config 4: the rows-on-demand GPU tests, the classic bench (dense tables: regression check of the
# accessor change), then the 3-D workload.
#   gpurun --timeout 1500 -- bash tools/r2_c4.sh [tag] [packets]
TAG=${1:-c4}
PK=${2:-12500000}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,memory.used --format=csv > $O/gpu.txt 2>&1
free -g >> $O/gpu.txt
( time timeout 600 python -m pytest tests/test_rows_on_demand.py tests/test_normalise.py tests/test_gpu_parity.py -m gpu -x -q -s ) > $O/pytest_rows.log 2>&1
echo "pytest rc=$?" >> $O/pytest_rows.log
( time timeout 300 python bench.py --steps 3 --warmup 3 --cpu-baseline off ) > $O/bench_classic.json 2> $O/bench_classic.err
( time ARTIS_HANG_S=60 timeout 1000 python bench.py --workload 3d --packets $PK --steps 2 --warmup 1 ) > $O/bench_3d.json 2> $O/bench_3d.err
echo "3d rc=$?" >> $O/bench_3d.err
nvidia-smi --query-gpu=memory.used --format=csv >> $O/gpu.txt 2>&1
tail -n 4 $O/pytest_rows.log; cut -c1-400 $O/bench_classic.json; tail -n 5 $O/bench_3d.err; cut -c1-600 $O/bench_3d.json
