#!/bin/bash
# the GPU test suite and smoke on the current tree (what the driver runs at round end)
TAG=${1:-check}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q -rs ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
( python bench.py --steps 3 --warmup 3 --cpu-baseline off ) > $O/bench.json 2> $O/bench.err
tail -n 5 $O/pytest_gpu.log; tail -n 2 $O/smoke.log; cut -c1-200 $O/bench.json
