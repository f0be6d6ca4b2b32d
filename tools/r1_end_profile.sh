#!/bin/bash
# Round-end GPU job: all GPU parity tests, smoke, both bench arms for both workloads, the ncu launch list and one full
# capture per workload, summarised on the box (the .ncu-rep files are too large to bring back).
# Run from the repo root under gpurun:  gpurun --timeout 2400 -- 'bash tools/r1_end_profile.sh'
mkdir -p gpurun_out ncu_tmp
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r1_end.json 2> gpurun_out/bench_r1_end.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1_end_reference.json 2> gpurun_out/bench_r1_end_reference.err
python bench.py --workload nebular --steps 5 --warmup 3 > gpurun_out/bench_r1_end_nebular.json 2> gpurun_out/bench_r1_end_nebular.err
python bench.py --workload nebular --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1_end_nebular_reference.json 2> gpurun_out/bench_r1_end_nebular_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r1_end_launches.csv \
  python bench.py --steps 2 --warmup 3 --cpu-baseline off > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_transport_persistent -s 3 -c 1 -f -o ncu_tmp/persistent \
  python bench.py --steps 1 --warmup 3 --cpu-baseline off > gpurun_out/ncu_full.log 2>&1
python tools/ncu_summary.py ncu_tmp/persistent.ncu-rep artis_b200/libartis_gpu.so k_transport_persistent \
  > gpurun_out/r1_end_persistent_ncu_summary.txt 2> gpurun_out/ncu_summary.err
ncu --set full --clock-control none --import-source on -k regex:k_transport_persistent -s 3 -c 1 -f -o ncu_tmp/nebular \
  python bench.py --workload nebular --steps 1 --warmup 3 --cpu-baseline off > gpurun_out/ncu_full_nebular.log 2>&1
python tools/ncu_summary.py ncu_tmp/nebular.ncu-rep artis_b200/libartis_gpu.so k_transport_persistent \
  > gpurun_out/r1_end_nebular_ncu_summary.txt 2>> gpurun_out/ncu_summary.err
