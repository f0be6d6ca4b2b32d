"""GPU parity on the BENCH workloads themselves: the models bench.py is quoted on -- the classic
100^3 model at timestep 25, the same with virtual packets + polarisation at timestep 30 (config 5), the kilonova model with
1.1e6 lines at timestep 7 (config 3), the one-zone nebular NLTE model on the 50^3 grid at timestep 8 (config 2) -- z-tested
against ensembles of the unmodified reference run on exactly those inputs
(oracle/tools/ref_bench_rows.py; tests/golden/bench_rows_*.npz hold the reference's summary rows and a fingerprint of the
inputs). The inputs are the run directories that __graft_entry__.build() makes where /root/reference exists
(oracle/_ref/runs/bench_*; bit-reproducible; they travel to the GPU box with the working tree, as for bench.py). The summary
includes the Stokes Q / U statistics of the REAL packets (statlib.summary_pol).

Collected late (the file name) because it is the slowest GPU test; a missing or different run directory is a skip with the
reason, not a failure of the library.
"""
import hashlib
import os

import numpy as np
import pytest

import statlib
from artis_b200 import snapshot
from artis_b200.transport import Transport

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)


def fingerprint(snap):  # oracle/tools/ref_bench_rows.py fingerprint()
    h = hashlib.sha1(np.ascontiguousarray(snap["packets.pre"]).tobytes())
    st = snapshot.static_tables(snap)
    h.update(f"{int(st['nlines'])}|{int(st['nbfcontinua'])}|{int(st['nlevels_total'])}|{int(st['npts_model'])}".encode())
    return h.hexdigest()


@pytest.mark.parametrize("workload", ["classic", "vpkt", "kilonova", "nebular"])
def test_cuda_ensemble_on_bench_workload(gpu_lib, workload):
    gold = os.path.join(HERE, "golden", f"bench_rows_{workload}.npz")
    if not os.path.exists(gold):
        pytest.skip(f"{gold} not generated")
    z = np.load(gold, allow_pickle=False)
    nts = int(z["nts"])
    snap_path = os.path.join(REPO, "oracle", "_ref", "runs", str(z["run"]), f"snap_ts{nts}.bin")
    if not os.path.exists(snap_path):
        pytest.skip(f"{snap_path} is missing: the bench workloads are made by __graft_entry__.build() where /root/reference exists")
    s = snapshot.load(snap_path)
    if fingerprint(s) != str(z["fingerprint"]):
        pytest.skip("the run directory was regenerated with different inputs than the committed reference rows: rerun "
                    "oracle/tools/ref_bench_rows.py")
    st, cs, pre = snapshot.static_tables(s), snapshot.cell_state(s), s["packets.pre"].copy()
    names = [str(x) for x in z["ref_names"]]
    T = Transport(gpu_lib)
    rows = []
    for k in range(16):
        T.init(st, seed=424242 + 104729 * k)
        T.upload_cellstate(cs)
        T.zero_estimators()
        pk = pre.copy()
        T.update_packets(pk)
        est = T.download_estimators()
        d = statlib.summary(st, cs, est, pk, pre)
        d.update(statlib.summary_pol(pk, pre))
        rows.append([d[n] for n in names])
    T.finalize()
    res = statlib.ztest(z["ref_rows"], np.array(rows), names)
    zs = np.array([zz for _, r, d, zz in res if (r != 0 or d != 0)])
    pol = [x for x in res if x[0].startswith("pol_") and (x[1] != 0 or x[2] != 0)]
    print(workload, "quantities", zs.size, "max|z|", np.abs(zs).max(), "rms z", np.sqrt((zs ** 2).mean()),
          "| polarisation quantities", len(pol), "max|z|", max([abs(x[3]) for x in pol], default=0.0))
    assert np.all(np.isfinite(zs)), [x for x in res if not np.isfinite(x[3])]
    assert np.abs(zs).max() < 4.5, [x for x in res if abs(x[3]) >= 4.5]
    assert np.sqrt((zs ** 2).mean()) < 1.5
