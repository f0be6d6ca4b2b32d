#!/bin/bash
# gpurun call: ncu evidence of the final build. (1) full capture of k_vpkt_trace on the config-5 workload, (2) full capture of
# the every-role persistent kernel on the classic workload (ncu serialises kernels, so the co-resident pair cannot run under
# it: single-kernel scheduler), (3) the launch list of one classic bench run and of one config-5 run.
TAG=${1:-call5}
O=gpurun_out/$TAG
mkdir -p $O ncu_tmp
export ARTIS_BENCH_SCHED=1 ARTIS_HANG_S=120
ncu --set full --clock-control none --import-source on -k regex:k_vpkt_trace -c 1 -f -o ncu_tmp/v \
  python bench.py --workload vpkt --steps 1 --warmup 0 --cpu-baseline off > $O/ncu_vpkt.log 2>&1
python tools/ncu_summary.py ncu_tmp/v.ncu-rep artis_b200/libartis_gpu_vpkt.so k_vpkt_trace > $O/vpkt_trace_ncu_summary.txt 2> $O/ncu_summary_v.err
ncu --set full --clock-control none --import-source on -k regex:k_transport_persistent -s 3 -c 1 -f -o ncu_tmp/p \
  python bench.py --steps 1 --warmup 1 --cpu-baseline off > $O/ncu_full.log 2>&1
python tools/ncu_summary.py ncu_tmp/p.ncu-rep artis_b200/libartis_gpu_classic.so k_transport_persistent > $O/persistent_ncu_summary.txt 2> $O/ncu_summary_p.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
  python bench.py --steps 2 --warmup 1 --cpu-baseline off > $O/ncu_bench.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_vpkt.csv \
  python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off > $O/ncu_bench_vpkt.log 2>&1
cp ncu_tmp/v.ncu-rep $O/ 2>/dev/null
head -50 $O/vpkt_trace_ncu_summary.txt
