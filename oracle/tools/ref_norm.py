#!/usr/bin/env python
"""Golden vectors for artis_gpu_normalise_estimators (TEST INFRASTRUCTURE; needs /root/reference, oracle/_ref/classic_small).

Runs the unmodified reference (classic mode, small grid, default synthetic atomic data of make_run.py) to timestep NTS,
restarts it there with ARTIS_HOOK_NORMALISE=1: after update_packets the hook (oracle/refglue/hook.cc) first writes the raw
estimators of the timestep as usual and then passes them through the reference's own radfield::normalise_J /
get_T_J_from_J / normalise_nuJ, the ffheating / colheating scaling of update_grid.cc:1041-1134 and normalise_grey
(grey_emissivities.cc:124-139). The fixture keeps the static tables, the cell state, the raw estimators and the normalised
per-cell values.

  python oracle/tools/ref_norm.py tests/golden/norm_classic.npz
"""
import os
import subprocess
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from artis_b200 import snapshot  # noqa: E402

NTS = 14


def main():
    out = sys.argv[1]
    run = os.path.join(REPO, "oracle", "_ref", "runs", "norm")
    exe = os.path.join(REPO, "oracle", "_ref", "classic_small", "sn3d_ref")
    subprocess.run(["make", "-C", os.path.join(REPO, "oracle"), "ref", "MODE=classic_small", "-j8"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["rm", "-rf", run])
    mk = [sys.executable, os.path.join(REPO, "oracle", "tools", "make_run.py"), "classic1d", run]
    subprocess.run(mk + ["--ftstep", str(NTS + 1)], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([exe], cwd=run, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    subprocess.run(mk + ["--itstep", str(NTS), "--ftstep", str(NTS + 1), "--resume"], check=True, stdout=subprocess.DEVNULL)
    env = dict(os.environ, ARTIS_HOOK_SNAPSHOT_TS=str(NTS), ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(NTS), ARTIS_HOOK_NORMALISE="1")
    subprocess.run([exe], cwd=run, env=env, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    s = snapshot.load(os.path.join(run, f"snap_ts{NTS}.bin"))
    keep = {k: np.array(v) for k, v in s.items() if k.startswith(("static.", "cell.", "estim.", "norm.", "meta."))}
    for k in ("J", "nuJ", "ff", "col", "rpkt_emiss", "TJ"):
        a = keep[f"norm.{k}"]
        print(k, a.size, "cells,", int((a != 0).sum()), "non-zero, max", a.max())
    print("gamma_dep_pathint, nprocs, MINTEMP, MAXTEMP", keep["norm.scalars"])
    np.savez_compressed(out, **keep)
    print("wrote", out, os.path.getsize(out) / 1e6, "MB")


if __name__ == "__main__":
    main()
