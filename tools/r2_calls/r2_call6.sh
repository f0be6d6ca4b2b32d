#!/bin/bash
# gpurun call: GPU tests; config 5 with the lock-step request tracer against the thread-per-request one; config 3 with the
# warp-per-row table kernels; config 4 with cooperative cooling rows on demand against a lane building them alone.
TAG=${1:-call6}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -n 5 $O/pytest_gpu.log
export ARTIS_HANG_S=60
show() {
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$1", round(d["ms_per_step"],1), "ms value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]), "Mcyc", c["warp_Mcycles_by_kind"][:6], "upload", c["s_per_timestep_parts_ms"], "req", c.get("vpkt_requests_per_step"), "rows", c.get("rows_on_demand"))
except Exception as e:
    print("$1 FAILED", e, open("$O/bench_$1.err").read()[-600:])
PY
}
for m in default ls4 ls2 t0; do
  LIBENV="ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_vpkt_$m.so"; [ "$m" = "default" ] && LIBENV="ARTIS_X=0"
  ( env $LIBENV timeout 600 python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off > $O/bench_vpkt_$m.json 2> $O/bench_vpkt_$m.err )
  show vpkt_$m
done
ARTIS_BENCH_SCHED=1 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum \
  --clock-control none -k regex:k_vpkt_trace -c 1 --csv --log-file $O/vpkt_trace_quick.csv python bench.py --workload vpkt --steps 1 --warmup 0 --cpu-baseline off > $O/ncu_quick.log 2>&1
cut -d, -f5,13,15 $O/vpkt_trace_quick.csv | tail -6
( timeout 600 python bench.py --workload kilonova --steps 3 --warmup 2 --cpu-baseline off > $O/bench_kilonova.json 2> $O/bench_kilonova.err ); show kilonova
( timeout 900 python bench.py --workload 3d --steps 1 --warmup 1 --cpu-baseline off > $O/bench_3d.json 2> $O/bench_3d.err ); show 3d
( ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_rows_nocoopcool.so timeout 900 python bench.py --workload 3d --steps 1 --warmup 1 --cpu-baseline off > $O/bench_3d_nocoopcool.json 2> $O/bench_3d_nocoopcool.err ); show 3d_nocoopcool
