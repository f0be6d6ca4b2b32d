#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the bench lines of both arms on the named workloads, the launch list and one
# full ncu capture of the transport kernel (single-kernel scheduler: ncu serialises kernels and the two-kernel pair needs
# both resident, so the pair cannot run under the profiler).
#   gpurun --timeout 1800 -- bash tools/r2_gpu_check.sh [tag]
TAG=${1:-check}
O=gpurun_out/$TAG
mkdir -p $O ncu_tmp
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt 2>&1
nproc >> $O/gpu.txt
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
( time python bench.py --impl reference ) > $O/bench_reference.json 2> $O/bench_reference.err
( time python bench.py ) > $O/bench.json 2> $O/bench.err
( time python bench.py --scheduler-kernels 1 --cpu-baseline off ) > $O/bench_single_kernel.json 2> $O/bench_single_kernel.err
( time python bench.py --workload nebular ) > $O/bench_nebular.json 2> $O/bench_nebular.err
( time python bench.py --workload nebular --impl reference --steps 2 --warmup 1 ) > $O/bench_nebular_reference.json 2> $O/bench_nebular_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
  python bench.py --steps 2 --warmup 1 --cpu-baseline off --scheduler-kernels 1 > $O/ncu_bench.log 2>&1
ARTIS_HANG_S=120 ncu --set full --clock-control none --import-source on -k regex:k_transport_persistent -s 3 -c 1 -f -o ncu_tmp/p \
  python bench.py --steps 1 --warmup 1 --cpu-baseline off --scheduler-kernels 1 > $O/ncu_full.log 2>&1
python tools/ncu_summary.py ncu_tmp/p.ncu-rep artis_b200/libartis_gpu.so k_transport_persistent \
  > $O/persistent_ncu_summary.txt 2> $O/ncu_summary.err
tail -n 3 $O/pytest_gpu.log; tail -n 2 $O/smoke.log; cut -c1-300 $O/bench.json; head -16 $O/persistent_ncu_summary.txt
