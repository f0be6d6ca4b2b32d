#!/bin/bash
# gpurun call: GPU parity tests (with the bench-workload ensembles), then config 5 A/B on one box: the threshold between the
# warp-cooperative and the lane-parallel virtual-packet ray trace, and the same timestep with virtual packets switched off.
TAG=${1:-call2}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -n 5 $O/pytest_gpu.log
export ARTIS_HANG_S=60
run() { # name, env...
  local name=$1; shift
  ( env "$@" timeout 600 python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off > $O/bench_$name.json 2> $O/bench_$name.err )
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$name.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$name", round(d["ms_per_step"],1), "ms", "Mcyc", c["warp_Mcycles_by_kind"][:6], c["warp_Mcycles_refill_push_idle"], "busy", c["mean_busy_lanes_by_kind"][:6])
except Exception as e:
    print("$name FAILED", e, open("$O/bench_$name.err").read()[-500:])
PY
}
run vpkt_default ARTIS_X=0
run vpkt_off ARTIS_BENCH_VPKT_OFF=1
for m in $VARIANTS; do
  run vpkt_$m ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_vpkt_$m.so
done
