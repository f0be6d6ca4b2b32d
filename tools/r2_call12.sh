#!/bin/bash
TAG=${1:-call12}
O=gpurun_out/$TAG
mkdir -p $O
export ARTIS_HANG_S=60
for m in $VARIANTS; do
  ( ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_classic_$m.so timeout 600 python bench.py --steps 2 --warmup 1 --cpu-baseline off > $O/bench_$m.json 2> $O/bench_$m.err )
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$m.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$m", round(d["ms_per_step"],1), "ms", "thin", c["warp_Mcycles_by_kind"][1], "begin/kappa", c.get("thin_role_Mcycles_begin_kappa"), "kprof [prologue, wait, terms, issue, owner, reduce]", c.get("kappa_profile_Mcycles"), "evals", c.get("rpkt_steps_and_kappa_evals_per_step"), "iters", round(c["warp_iters_by_kind"][1]/1e6,2))
except Exception as e:
    print("$m FAILED", e, open("$O/bench_$m.err").read()[-600:])
PY
done
