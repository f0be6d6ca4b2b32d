#!/bin/bash
# one full ncu capture of the every-role persistent kernel (single-kernel scheduler: ncu serialises kernels, the two-kernel
# pair needs both resident) on the classic workload at $1 packets, summarised on the box
P=${1:-300000}; TAG=${2:-ncu}
mkdir -p gpurun_out/$TAG ncu_tmp
export ARTIS_BENCH_SCHED=1 ARTIS_HANG_S=60
ncu --set full --clock-control none --import-source on -k regex:k_transport_persistent -s 3 -c 1 -f -o ncu_tmp/p \
  python bench.py --steps 1 --warmup 1 --packets $P --cpu-baseline off > gpurun_out/$TAG/ncu_full.log 2>&1
python tools/ncu_summary.py ncu_tmp/p.ncu-rep ${ARTIS_GPU_LIB:-artis_b200/libartis_gpu.so} k_transport_persistent \
  > gpurun_out/$TAG/persistent_ncu_summary.txt 2> gpurun_out/$TAG/ncu_summary.err
head -40 gpurun_out/$TAG/persistent_ncu_summary.txt
