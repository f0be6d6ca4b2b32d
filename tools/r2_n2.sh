#!/bin/bash
# gpurun --gpus 2: the library's NCCL exchange test, then the classic and config-4 lines at N = 2 (weak scaling)
TAG=${1:-n2}
O=gpurun_out/$TAG
mkdir -p $O
( time python -m pytest tests/test_nccl.py -m gpu -x -q -rs ) > $O/pytest_nccl.log 2>&1
echo "pytest rc=$?" >> $O/pytest_nccl.log
tail -n 4 $O/pytest_nccl.log
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 ) > $O/bench_n2.json 2> $O/bench_n2.err
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 ) > $O/bench_n2_reference.json 2> $O/bench_n2_reference.err
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload 3d --steps 2 --warmup 1 ) > $O/bench_3d_n2.json 2> $O/bench_3d_n2.err
for f in bench_n2 bench_n2_reference bench_3d_n2; do echo "== $f"; grep '"metric"' $O/$f.json | cut -c1-300; tail -n 3 $O/$f.err | cut -c1-200; done
