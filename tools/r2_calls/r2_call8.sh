#!/bin/bash
# gpurun call: GPU tests with the live-continuum lists; every workload with and without them (ARTIS_BF_LIVE=0), same box
TAG=${1:-call8}
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
    print("$1", round(d["ms_per_step"],1), "ms value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]), "roofline %.3f" % d["roofline"]["frac"], "Mcyc", c["warp_Mcycles_by_kind"][:6], c["warp_Mcycles_refill_push_idle"], "busy", c["mean_busy_lanes_by_kind"][:6], "upload", c["s_per_timestep_parts_ms"])
except Exception as e:
    print("$1 FAILED", e, open("$O/bench_$1.err").read()[-600:])
PY
}
for w in classic vpkt nebular kilonova; do
  ( timeout 600 python bench.py --workload $w --steps 3 --warmup 2 --cpu-baseline off > $O/bench_$w.json 2> $O/bench_$w.err ); show $w
  ( ARTIS_BF_LIVE=0 timeout 600 python bench.py --workload $w --steps 3 --warmup 2 --cpu-baseline off > $O/bench_${w}_all.json 2> $O/bench_${w}_all.err ); show ${w}_all
done
( timeout 900 python bench.py --workload 3d --steps 1 --warmup 1 --cpu-baseline off > $O/bench_3d.json 2> $O/bench_3d.err ); show 3d
