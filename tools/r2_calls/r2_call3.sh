#!/bin/bash
TAG=${1:-call3}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -n 5 $O/pytest_gpu.log
export ARTIS_HANG_S=60
for m in $VARIANTS; do
  LIBENV="ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_vpkt_$m.so"; [ "$m" = "default" ] && LIBENV="ARTIS_X=0"
  ( env $LIBENV timeout 600 python bench.py --workload vpkt --steps 1 --warmup 1 --cpu-baseline off > $O/bench_$m.json 2> $O/bench_$m.err )
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_$m.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("$m", round(d["ms_per_step"],1), "ms", "Mcyc", c["warp_Mcycles_by_kind"][:6], "prof", c.get("vpkt_profile"), "req", c.get("vpkt_requests_per_step"), c.get("vpkt_list_restarts_per_step"))
except Exception as e:
    print("$m FAILED", e, open("$O/bench_$m.err").read()[-500:])
PY
done
