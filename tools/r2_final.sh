#!/bin/bash
# the final build: GPU tests (all), smoke, the default bench line with its reference arm, config 5 and config 3 lines
TAG=${1:-final}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests -m gpu -x -q -rs ) > $O/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
( time python bench.py --impl reference ) > $O/bench_reference.json 2> $O/bench_reference.err
( time python bench.py ) > $O/bench.json 2> $O/bench.err
( time python bench.py --workload vpkt --cpu-baseline off ) > $O/bench_vpkt.json 2> $O/bench_vpkt.err
( time python bench.py --workload kilonova --cpu-baseline off ) > $O/bench_kilonova.json 2> $O/bench_kilonova.err
( time python bench.py --workload nebular --cpu-baseline off ) > $O/bench_nebular.json 2> $O/bench_nebular.err
tail -n 5 $O/pytest_gpu.log; tail -n 2 $O/smoke.log
for f in bench bench_reference bench_vpkt bench_kilonova bench_nebular; do echo "== $f"; cut -c1-300 $O/$f.json; tail -n 2 $O/$f.err | cut -c1-200; done
