#!/bin/sh
# TEST INFRASTRUCTURE. Recreates oracle/_ref/runs/bench_base, the workload of bench.py: the unmodified reference
# (oracle/_ref/classic/sn3d_ref: artisoptions_classic.h, 100^3 grid, 15000 packets) run on the classic-mode test model
# with synthetic atomic data (46255 lines / 3204 bf continua / 2886 levels) through timestep 25, left ready to restart at
# timestep 25, plus the hook's snapshot of the inputs of that update_packets call (snap_ts25.bin).
# Needs /root/reference; takes a few minutes on one core.   sh oracle/tools/make_bench_run.sh
set -e
cd "$(dirname "$0")/../.."
RUN=oracle/_ref/runs/bench_base
ARGS="--nlevels 120 --dense-levels 20 --sparsity 0.25"
make -C oracle ref MODE=classic -j8 > /dev/null
rm -rf $RUN
python oracle/tools/make_run.py classic1d $RUN --ftstep 26 $ARGS  # runs timesteps 0-25, leaves the *_ts25.tmp checkpoint
(cd $RUN && ../../classic/sn3d_ref > stdout.txt 2>&1)
python oracle/tools/make_run.py classic1d $RUN --itstep 25 --ftstep 26 --resume $ARGS
(cd $RUN && ARTIS_HOOK_SNAPSHOT_TS=25 ARTIS_HOOK_DIR=. ARTIS_HOOK_STOP_AFTER_TS=25 ARTIS_HOOK_SKIP_POST=1 \
   ../../classic/sn3d_ref > stdout_snap.txt 2>&1)
ls -la $RUN/snap_ts25.bin $RUN/gridsave_ts25.tmp $RUN/packets_0000_ts25.tmp
