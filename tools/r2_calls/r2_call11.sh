#!/bin/bash
# gpurun call: the classic bench with the lane-parallel opacity sums from 1 / 8 / 16 owners on, against the always-cooperative default
TAG=${1:-call11}
O=gpurun_out/$TAG
mkdir -p $O
export ARTIS_HANG_S=60
show() {
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$1", round(d["ms_per_step"],1), "ms value %.4g" % d["value"], "Mcyc", c["warp_Mcycles_by_kind"][:6], "thin begin/kappa", c.get("thin_role_Mcycles_begin_kappa"), c["warp_Mcycles_refill_push_idle"], "busy", c["mean_busy_lanes_by_kind"][:6], "steps/evals", c.get("rpkt_steps_and_kappa_evals_per_step"), "iters", [round(x/1e6,2) for x in c["warp_iters_by_kind"][:6]])
except Exception as e:
    print("$1 FAILED", e, open("$O/bench_$1.err").read()[-600:])
PY
}
( timeout 600 python bench.py --steps 3 --warmup 2 --cpu-baseline off > $O/bench_default.json 2> $O/bench_default.err ); show default
for m in $VARIANTS; do
  ( ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_classic_$m.so timeout 600 python bench.py --steps 3 --warmup 2 --cpu-baseline off > $O/bench_$m.json 2> $O/bench_$m.err ); show $m
done
