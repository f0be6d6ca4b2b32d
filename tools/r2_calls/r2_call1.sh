#!/bin/bash
# One gpurun call after the container was re-created: GPU parity tests, smoke, the classic bench (both arms), and the first
# lines of the two new workloads (config 5: virtual packets, warp-cooperative ray trace against the one-lane build; config 3:
# kilonova with 1.1e6 lines), each with its reference arm.   gpurun --timeout 2400 -- bash tools/r2_calls/r2_call1.sh [tag]
TAG=${1:-call1}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt 2>&1
nproc >> $O/gpu.txt
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
export ARTIS_HANG_S=60
( time python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off ) > $O/bench_vpkt.json 2> $O/bench_vpkt.err
( time ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_vpkt_serial.so python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off ) > $O/bench_vpkt_serial.json 2> $O/bench_vpkt_serial.err
( time python bench.py --workload vpkt --impl reference --steps 1 --warmup 1 ) > $O/bench_vpkt_reference.json 2> $O/bench_vpkt_reference.err
( time python bench.py --workload kilonova --steps 2 --warmup 1 --cpu-baseline off ) > $O/bench_kilonova.json 2> $O/bench_kilonova.err
( time python bench.py --workload kilonova --impl reference --steps 1 --warmup 1 ) > $O/bench_kilonova_reference.json 2> $O/bench_kilonova_reference.err
unset ARTIS_HANG_S
( time python bench.py --impl reference ) > $O/bench_reference.json 2> $O/bench_reference.err
( time python bench.py ) > $O/bench.json 2> $O/bench.err
tail -n 3 $O/pytest_gpu.log; tail -n 2 $O/smoke.log
for f in bench bench_reference bench_vpkt bench_vpkt_serial bench_vpkt_reference bench_kilonova bench_kilonova_reference; do
  echo "== $f"; cut -c1-400 $O/$f.json; tail -n 4 $O/$f.err | cut -c1-300
done
