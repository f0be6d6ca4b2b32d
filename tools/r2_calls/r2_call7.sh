#!/bin/bash
TAG=${1:-call7}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q -k "vpkt or bench_workload" ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -n 5 $O/pytest_gpu.log
export ARTIS_HANG_S=60
for m in $VARIANTS; do
  LIBENV="ARTIS_GPU_LIB=$PWD/variants/libartis_gpu_vpkt_$m.so"; [ "$m" = "default" ] && LIBENV="ARTIS_X=0"
  ( env $LIBENV timeout 600 python bench.py --workload vpkt --steps 2 --warmup 1 --cpu-baseline off > $O/bench_vpkt_$m.json 2> $O/bench_vpkt_$m.err )
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_vpkt_$m.json").read().strip().splitlines()[-1])
    print("$m", round(d["ms_per_step"],1), "ms value %.4g" % d["value"], "launches", d["gpu_launches"])
except Exception as e:
    print("$m FAILED", e, open("$O/bench_vpkt_$m.err").read()[-600:])
PY
done
