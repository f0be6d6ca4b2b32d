#!/bin/bash
TAG=${1:-call12}
O=gpurun_out/$TAG
mkdir -p $O
export ARTIS_HANG_S=60
for m in $VARIANTS; do
  LIBENV="ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_classic_$m.so"; [ "$m" = "default" ] && LIBENV="ARTIS_X=0"
  ( env $LIBENV timeout 600 python bench.py --steps 3 --warmup 2 --cpu-baseline off > $O/bench_$m.json 2> $O/bench_$m.err )
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$m.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$m", round(d["ms_per_step"],1), "ms", "thin", c["warp_Mcycles_by_kind"][1], "begin/kappa", c.get("thin_role_Mcycles_begin_kappa"), "kprof [prologue, wait, terms, issue, owner, reduce]", c.get("kappa_profile_Mcycles"), "evals", c.get("rpkt_steps_and_kappa_evals_per_step"), "iters", round(c["warp_iters_by_kind"][1]/1e6,2), "geo", c["scheduler_geometry_kernels_blocks_single_big_jump"], "Mcyc", c["warp_Mcycles_by_kind"][:6], c["warp_Mcycles_refill_push_idle"], "busy", c["mean_busy_lanes_by_kind"][:6])
except Exception as e:
    print("$m FAILED", e, open("$O/bench_$m.err").read()[-600:])
PY
done
