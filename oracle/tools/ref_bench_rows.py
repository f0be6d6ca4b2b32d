#!/usr/bin/env python
"""Reference ensembles of the BENCH workloads themselves (TEST INFRASTRUCTURE; needs /root/reference, a built
oracle/_ref/<mode>/sn3d_ref and the workload's run directory, i.e. __graft_entry__.build()).

The golden fixtures under tests/golden pin the physics on shrunken builds (16^3-20^3 grids). This pins the models the bench
is quoted on -- the classic 100^3 model at timestep 25, the same with virtual packets at timestep 30, the kilonova model with
1.1e6 lines at timestep 7 -- in the same way: K runs of the unmodified reference over the workload timestep from the
workload's checkpoint with K seeds, summarised by tests/statlib.summary() + summary_pol(). Only the K summary rows and a
fingerprint of the inputs are committed (tests/golden/bench_rows_<workload>.npz, a few kB): the inputs are the run
directory that build() regenerates (oracle/_ref/runs/bench_*, bit-reproducible, travels to the GPU box), and
tests/test_zy_bench_workloads.py checks the fingerprint before comparing.

  python oracle/tools/ref_bench_rows.py classic|vpkt|kilonova [--seeds 24] [--procs 8]
"""
import argparse
import hashlib
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
from artis_b200 import snapshot  # noqa: E402
import statlib  # noqa: E402

WORKLOADS = {  # name -> (run directory, build mode, timestep)
    "classic": ("bench_base", "classic", 25),
    "vpkt": ("bench_vpkt", "classic_vpkt", 30),
    "kilonova": ("bench_kn", "kilonova", 7),
    "nebular": ("bench_neb", "nebular", 8),
}
SKIP = (".out", "output_", "stdout", "snap_", "hook_timing", "packets00_")


def fingerprint(snap):
    """identifies the inputs of the workload's update_packets call: the packets before the call and the sizes of the tables"""
    h = hashlib.sha1(np.ascontiguousarray(snap["packets.pre"]).tobytes())
    st = snapshot.static_tables(snap)
    h.update(f"{int(st['nlines'])}|{int(st['nbfcontinua'])}|{int(st['nlevels_total'])}|{int(st['npts_model'])}".encode())
    return h.hexdigest()


def set_seed(rundir, seed):
    path = os.path.join(rundir, "input.txt")
    lines = open(path).read().splitlines()
    for i, line in enumerate(lines):
        if line.strip() and not line.strip().startswith("#"):
            lines[i] = f"{seed}".ljust(25) + "# pre_zseed (set by ref_bench_rows.py)"
            break
    open(path, "w").write("\n".join(lines) + "\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workload", choices=sorted(WORKLOADS))
    ap.add_argument("--seeds", type=int, default=24)
    ap.add_argument("--procs", type=int, default=8)
    a = ap.parse_args()
    run, mode, nts = WORKLOADS[a.workload]
    src = os.path.join(REPO, "oracle", "_ref", "runs", run)
    exe = os.path.join(REPO, "oracle", "_ref", mode, "sn3d_ref")
    fp0 = fingerprint(snapshot.load(os.path.join(src, f"snap_ts{nts}.bin")))
    rows, names = [], None
    with tempfile.TemporaryDirectory(prefix="artis_rows_") as tmp:
        dirs = []
        for r in range(min(a.procs, a.seeds)):
            d = os.path.join(tmp, f"r{r}")
            os.makedirs(d)
            for f in os.listdir(src):
                if not f.startswith(SKIP[1:]) and not f.endswith(SKIP[0]):
                    shutil.copy(os.path.join(src, f), d)
            dirs.append(d)
        k = 0
        while k < a.seeds:
            batch = []
            for d in dirs:
                if k >= a.seeds:
                    break
                set_seed(d, 700001 + 7919 * k)
                env = dict(os.environ, ARTIS_HOOK_SNAPSHOT_TS=str(nts), ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(nts))
                batch.append((d, subprocess.Popen([exe], cwd=d, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)))
                k += 1
            for d, proc in batch:
                assert proc.wait() == 0, d
                s = snapshot.load(os.path.join(d, f"snap_ts{nts}.bin"))
                assert fingerprint(s) == fp0, "restart did not reproduce the workload's start state"
                st, cs, est = snapshot.static_tables(s), snapshot.cell_state(s), snapshot.estimators(s)
                dd = statlib.summary(st, cs, est, s["packets.post"], s["packets.pre"])
                dd.update(statlib.summary_pol(s["packets.post"], s["packets.pre"]))
                if names is None:
                    names = list(dd)
                rows.append([dd[nm] for nm in names])
                print(len(rows), "interactions", dd["interactions"], flush=True)
    out = os.path.join(REPO, "tests", "golden", f"bench_rows_{a.workload}.npz")
    np.savez_compressed(out, ref_rows=np.array(rows), ref_names=np.array(names), nts=nts, fingerprint=np.array(fp0),
                        run=np.array(run))
    print("wrote", out, os.path.getsize(out) / 1e3, "kB")


if __name__ == "__main__":
    main()
