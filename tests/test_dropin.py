"""The drop-in, run: the reference's own driver (sn3d.cc:514-673, every object compiled from the unmodified sources) with
its update_packets call site served by libartis_gpu.so through the C ABI (oracle/refglue/hook_gpu.cc -> sn3d_gpu), over
NTS consecutive timesteps with the host's update_grid between them -- the estimators the device returns drive the
temperatures of the next timestep, which drive the opacities the device transports through next.

Checked against an ensemble of runs of the same driver with its own host transport (sn3d_ref, 24 seeds;
tests/golden/dropin_classic.npz by oracle/tools/ref_dropin.py) on what the driver writes itself: gamma-ray deposition per
timestep (deposition.out), summed light curves (light_curve.out) and the plasma state of every cell after the last
timestep (T_R, T_e, W, T_J, n_e of estimators_0000.out). Two-sample z per observable.

sn3d_gpu and its template run directory are made by __graft_entry__.build() where /root/reference exists and travel to the
GPU box under oracle/_ref (git-ignored); without them the test is skipped with the reason."""
import os
import shutil
import tempfile

import numpy as np
import pytest

import dropinlib

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(REPO, "oracle", "_ref", "classic_small", "sn3d_gpu")
TEMPLATE = os.path.join(REPO, "oracle", "_ref", "runs", "dropin")
GOLDEN = os.path.join(REPO, "tests", "golden", "dropin_classic.npz")
NRUNS = 4


def test_dropin_golden_is_consistent():
    z = np.load(GOLDEN)
    rows, names = z["ref_rows"], [str(n) for n in z["names"]]
    assert rows.shape[0] >= 16 and rows.shape[1] == len(names) and int(z["nts"]) == dropinlib.NTS
    assert "gammadep_ts00" in names and any(n.startswith("TJ_cell") for n in names)
    assert np.all(np.isfinite(rows))


@pytest.mark.gpu
def test_reference_driver_runs_on_the_gpu_path_for_many_timesteps(gpu_lib):
    if not (os.path.exists(EXE) and os.path.isdir(TEMPLATE)):
        pytest.skip("oracle/_ref/classic_small/sn3d_gpu or its run template is missing (made by __graft_entry__.build() "
                    "where /root/reference exists)")
    z = np.load(GOLDEN)
    ref_rows, names = z["ref_rows"], [str(n) for n in z["names"]]
    rows = []
    with tempfile.TemporaryDirectory(prefix="dropin_gpu_") as tmp:
        dirs = []
        for k in range(NRUNS):
            d = os.path.join(tmp, f"g{k}")
            dropinlib.clone_run(TEMPLATE, d)
            dirs.append(d)
        # one driver at a time: the scheduler's two persistent kernels expect the whole GPU (two processes time-slicing it
        # would each see the other's kernels holding the SMs)
        for k0 in range(0, NRUNS, 1):
            procs = [(d, dropinlib.run(EXE, d, 550001 + 104729 * (k0 + i))) for i, d in enumerate(dirs[k0:k0 + 1])]
            for d, p in procs:
                rc = p.wait()
                assert rc == 0, open(os.path.join(d, "stdout.txt")).read()[-3000:]
                log = open(os.path.join(d, "hook_timing.txt")).read().split("\n")
                assert sum(1 for line in log if line.endswith("gpu")) == dropinlib.NTS  # every timestep went through the C ABI
                o = dropinlib.observables(d)
                rows.append([o[n] for n in names])
    res = dropinlib.ztest(ref_rows, rows, names)
    zs = np.array([r[3] for r in res])
    worst = sorted(res, key=lambda r: -abs(r[3]))[:5]
    print("drop-in:", len(names), "observables, max|z|", np.abs(zs).max(), "rms z", np.sqrt(np.mean(zs[np.isfinite(zs)] ** 2)), worst)
    assert np.all(np.isfinite(zs)), [r for r in res if not np.isfinite(r[3])]
    assert np.abs(zs).max() < 5.0, worst
    assert np.sqrt(np.mean(zs ** 2)) < 1.6
