#!/bin/sh
# TEST INFRASTRUCTURE. Recreates oracle/_ref/runs/bench_vpkt, the workload of `bench.py --workload vpkt` (BASELINE.json
# config 5): the unmodified reference built with VPKT_ON (oracle/modes/classic_vpkt.sh: 100^3 grid, 15000
# packets, POL_ON, DIPOLE) on the classic-mode test model with the synthetic atomic data of make_bench_run.sh and a vpkt.txt
# of five observers (cos theta -0.8 ... 0.8, phi 0), one spectrum each, compiled windows, tau_max 10, no velocity maps, run
# through timestep 29 and left ready to restart at timestep 30 (11.9-12.5 d: inside the compiled 10-30 d window of vpkt.h,
# so every eligible event of the timestep sends its five virtual packets), plus the hook's snapshot of that call.
# Needs /root/reference.   sh oracle/tools/make_bench_vpkt.sh
set -e
cd "$(dirname "$0")/../.."
RUN=oracle/_ref/runs/bench_vpkt
ARGS="--nlevels 120 --dense-levels 20 --sparsity 0.25 --vpkt s5"
make -C oracle ref MODE=classic_vpkt -j8 > /dev/null
rm -rf $RUN
python oracle/tools/make_run.py classic1d $RUN --ftstep 31 $ARGS  # runs timesteps 0-30, leaves the *_ts30.tmp checkpoint
(cd $RUN && ../../classic_vpkt/sn3d_ref > stdout.txt 2>&1)
python oracle/tools/make_run.py classic1d $RUN --itstep 30 --ftstep 31 --resume $ARGS
(cd $RUN && ARTIS_HOOK_SNAPSHOT_TS=30 ARTIS_HOOK_DIR=. ARTIS_HOOK_STOP_AFTER_TS=30 ARTIS_HOOK_SKIP_POST=1 \
   ../../classic_vpkt/sn3d_ref > stdout_snap.txt 2>&1)
ls -la $RUN/snap_ts30.bin $RUN/gridsave_ts30.tmp $RUN/packets_0000_ts30.tmp
