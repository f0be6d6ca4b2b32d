#!/bin/sh
# TEST INFRASTRUCTURE. Recreates oracle/_ref/runs/bench_kn, the workload of `bench.py --workload kilonova` (BASELINE.json
# config 3): the unmodified reference in kilonova LTE mode (oracle/modes/kilonova.sh: artisoptions_kilonova_lte.h
# with the overrides of tests/setup_kilonova.sh, 50^3 grid, 40000 packets) on the 1-D kilonova test model
# (tests/kilonova_inputfiles) with a synthetic line list of more than 1e6 lines -- 12 ions (Fe I-V, Co II-IV, Ni II-V) x 500
# levels, 70 % of all level pairs connected -- the list that stresses closest_transition and the line sweep; run through
# timestep 6 (the grey phase ends after timestep 4) and left ready to restart at timestep 7, plus the hook's snapshot of
# that update_packets call.   Needs /root/reference.   sh oracle/tools/make_bench_kn.sh
set -e
cd "$(dirname "$0")/../.."
RUN=oracle/_ref/runs/bench_kn
ARGS="--nlevels 500 --dense-levels 20 --sparsity 0.7"
make -C oracle ref MODE=kilonova -j8 > /dev/null
rm -rf $RUN
python oracle/tools/make_run.py kilonova1d $RUN --ftstep 8 $ARGS
(cd $RUN && ../../kilonova/sn3d_ref > stdout.txt 2>&1)
python oracle/tools/make_run.py kilonova1d $RUN --itstep 7 --ftstep 8 --resume $ARGS
(cd $RUN && ARTIS_HOOK_SNAPSHOT_TS=7 ARTIS_HOOK_DIR=. ARTIS_HOOK_STOP_AFTER_TS=7 ARTIS_HOOK_SKIP_POST=1 \
   ../../kilonova/sn3d_ref > stdout_snap.txt 2>&1)
ls -la $RUN/snap_ts7.bin $RUN/gridsave_ts7.tmp $RUN/packets_0000_ts7.tmp
