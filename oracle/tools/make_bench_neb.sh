#!/bin/sh
# TEST INFRASTRUCTURE. Recreates oracle/_ref/runs/bench_neb, the workload of `bench.py --workload nebular` (BASELINE.json
# config 2): the unmodified reference in nebular NLTE mode (oracle/modes/nebular.sh: artisoptions_nltenebular.h with the
# overrides of tests/setup_nebularonezone.sh, 50^3 grid, 4096-point Spencer-Fano grid; 100000 packets instead of 1e6 so
# that the snapshot stays small -- bench.py replicates them) on the one-zone test model with synthetic Fe/Co/Ni atomic
# data, run up to timestep 8 and left ready to restart there, plus the hook's snapshot of that update_packets call.
# Needs /root/reference.   sh oracle/tools/make_bench_neb.sh
set -e
cd "$(dirname "$0")/../.."
RUN=oracle/_ref/runs/bench_neb
ARGS="--nions 4 --nlevels 100 --dense-levels 12 --sparsity 0.51"  # 48577 lines / 1035 continua / 1200 levels
rm -f oracle/_ref/nebular/inc/artisoptions.h
ARTIS_MPKTS=100000 make -C oracle ref MODE=nebular -j8 > /dev/null
rm -rf $RUN
python oracle/tools/make_run.py nebular1zone $RUN --ftstep 9 $ARGS
(cd $RUN && ../../nebular/sn3d_ref > stdout.txt 2>&1)
python oracle/tools/make_run.py nebular1zone $RUN --itstep 8 --ftstep 9 --resume $ARGS
(cd $RUN && ARTIS_HOOK_SNAPSHOT_TS=8 ARTIS_HOOK_DIR=. ARTIS_HOOK_STOP_AFTER_TS=8 ARTIS_HOOK_SKIP_POST=1 \
   ../../nebular/sn3d_ref > stdout_snap.txt 2>&1)
ls -la $RUN/snap_ts8.bin $RUN/gridsave_ts8.tmp $RUN/packets_0000_ts8.tmp
