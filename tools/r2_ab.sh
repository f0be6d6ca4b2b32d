#!/bin/bash
# one gpurun call: GPU tests, then the classic bench under several scheduler geometries (same box)
TAG=${1:-ab}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -n 4 $O/pytest_gpu.log
run() { # name, env..., -- bench args
  local name=$1; shift
  ( env "$@" timeout 300 python bench.py --steps 3 --warmup 2 --cpu-baseline off > $O/bench_$name.json 2> $O/bench_$name.err ) 
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$name.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$name", round(d["ms_per_step"],1), "ms", "geo", c.get("scheduler_geometry_kernels_blocks_single_big_jump"), "Mcyc", c["warp_Mcycles_by_kind"], c["warp_Mcycles_refill_push_idle"], "busy", c["mean_busy_lanes_by_kind"])
except Exception as e:
    print("$name FAILED", e, open("$O/bench_$name.err").read()[-500:])
PY
}
export ARTIS_HANG_S=5
# RUNS: space-separated name:lib:geometry (lib "-" = the shipped library, geometry "-" = the library's default)
for spec in $RUNS; do
  IFS=: read name lib geo <<< "$spec"
  envs="ARTIS_X=0"
  [ "$lib" != "-" ] && envs="$envs ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_$lib.so"
  [ "$geo" != "-" ] && envs="$envs ARTIS_GEOMETRY=$geo"
  [ "$name" = "single" ] && envs="$envs ARTIS_BENCH_SCHED=1"
  run $name $envs
done
