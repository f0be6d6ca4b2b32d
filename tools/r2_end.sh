#!/bin/bash
# Round-end evidence, one gpurun call on one GPU: GPU parity tests, smoke, every workload's bench line with its reference arm,
# the launch lists, and full ncu captures of the two kernels that carry the time (the every-role persistent kernel -- ncu
# serialises kernels, the co-resident pair cannot run under it -- and k_vpkt_trace).
#   gpurun --timeout 2400 -- bash tools/r2_end.sh [tag]
TAG=${1:-end}
O=gpurun_out/$TAG
mkdir -p $O ncu_tmp
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt 2>&1
nproc >> $O/gpu.txt
( time python -m pytest tests -m gpu -x -q -rs ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
( time python bench.py --impl reference ) > $O/bench_reference.json 2> $O/bench_reference.err
( time python bench.py ) > $O/bench.json 2> $O/bench.err
for w in nebular kilonova vpkt; do
  ( time python bench.py --workload $w --impl reference --steps 2 --warmup 1 ) > $O/bench_${w}_reference.json 2> $O/bench_${w}_reference.err
  ( time python bench.py --workload $w --cpu-baseline off ) > $O/bench_$w.json 2> $O/bench_$w.err
done
( time python bench.py --workload 3d --steps 2 --warmup 3 --cpu-baseline off ) > $O/bench_3d.json 2> $O/bench_3d.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
  python bench.py --steps 2 --warmup 1 --cpu-baseline off --scheduler-kernels 1 > $O/ncu_bench.log 2>&1
ARTIS_BENCH_SCHED=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_vpkt.csv \
  python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off > $O/ncu_bench_vpkt.log 2>&1
ARTIS_HANG_S=120 ncu --set full --clock-control none --import-source on -k regex:k_transport_persistent -s 3 -c 1 -f -o ncu_tmp/p \
  python bench.py --steps 1 --warmup 1 --cpu-baseline off --scheduler-kernels 1 > $O/ncu_full.log 2>&1
python tools/ncu_summary.py ncu_tmp/p.ncu-rep artis_b200/libartis_gpu_classic.so k_transport_persistent \
  > $O/persistent_ncu_summary.txt 2> $O/ncu_summary_p.err
ARTIS_BENCH_SCHED=1 ARTIS_HANG_S=120 ncu --set full --clock-control none --import-source on -k regex:k_vpkt_trace -c 1 -f -o ncu_tmp/v \
  python bench.py --workload vpkt --steps 1 --warmup 0 --cpu-baseline off > $O/ncu_vpkt.log 2>&1
python tools/ncu_summary.py ncu_tmp/v.ncu-rep artis_b200/libartis_gpu_vpkt.so k_vpkt_trace > $O/vpkt_trace_ncu_summary.txt 2> $O/ncu_summary_v.err
tail -n 4 $O/pytest_gpu.log; tail -n 2 $O/smoke.log
for f in bench bench_reference bench_nebular bench_nebular_reference bench_kilonova bench_kilonova_reference bench_vpkt bench_vpkt_reference bench_3d; do
  echo "== $f"; cut -c1-330 $O/$f.json; tail -n 2 $O/$f.err | cut -c1-200
done
head -24 $O/persistent_ncu_summary.txt; head -24 $O/vpkt_trace_ncu_summary.txt
