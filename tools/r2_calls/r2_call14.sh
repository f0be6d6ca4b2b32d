#!/bin/bash
# config 4 with a smaller row arena, so that the live-continuum lists fit next to it
TAG=${1:-call14}
O=gpurun_out/$TAG
mkdir -p $O
export ARTIS_HANG_S=120
for mb in $ARENAS; do
  ( ARTIS_ROW_ARENA_MB=$mb timeout 900 python bench.py --workload 3d --steps 1 --warmup 1 --cpu-baseline off > $O/bench_3d_$mb.json 2> $O/bench_3d_$mb.err )
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_3d_$mb.json").read().strip().splitlines()[-1])
    c=d["config"]
    print("arena $mb MB:", round(d["ms_per_step"],1), "ms value %.4g" % d["value"], "Mcyc", c["warp_Mcycles_by_kind"][:6], "rows", c.get("rows_on_demand"), "table GB", round(c["table_bytes"]/1e9,1), "evaluated/listed", d["roofline"]["algorithmic_bytes_per_launch"], d["roofline"]["algorithmic_bytes_per_launch_every_continuum"])
except Exception as e:
    print("$mb FAILED", e, open("$O/bench_3d_$mb.err").read()[-600:])
PY
done
