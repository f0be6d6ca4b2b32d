#!/bin/bash
# gpurun call: GPU tests, then the classic and config-5 lines with the per-cell free-free sum, the rank-mapped live ranges and
# the thin role's phase counters
TAG=${1:-call10}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -n 4 $O/pytest_gpu.log
export ARTIS_HANG_S=60
show() {
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$1.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$1", round(d["ms_per_step"],1), "ms value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]), "roofline %.3f / %.3f" % (d["roofline"]["frac"], d["roofline"].get("frac_every_continuum", 0)), "Mcyc", c["warp_Mcycles_by_kind"][:6], "thin begin/kappa", c.get("thin_role_Mcycles_begin_kappa"), c["warp_Mcycles_refill_push_idle"], "busy", c["mean_busy_lanes_by_kind"][:6])
except Exception as e:
    print("$1 FAILED", e, open("$O/bench_$1.err").read()[-600:])
PY
}
for w in classic vpkt; do
  ( timeout 600 python bench.py --workload $w --steps 3 --warmup 2 --cpu-baseline off > $O/bench_$w.json 2> $O/bench_$w.err ); show $w
done
( ARTIS_BF_LIVE=0 timeout 600 python bench.py --steps 3 --warmup 2 --cpu-baseline off > $O/bench_classic_all.json 2> $O/bench_classic_all.err ); show classic_all
( timeout 600 python bench.py --steps 3 --warmup 2 --cpu-baseline off --scheduler-kernels 1 > $O/bench_classic_single.json 2> $O/bench_classic_single.err ); show classic_single
