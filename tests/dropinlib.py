"""Observables of a multi-timestep run of the reference driver (sn3d_ref, or sn3d_gpu = the same driver with its
update_packets call served by libartis_gpu.so), read from the files the driver itself writes: deposition.out
(update_grid.cc / sn3d.cc write_deposition_file), light_curve.out (light_curve.cc:10-32), estimators_0000.out
(update_grid.cc write_to_estimators_file). Used by oracle/tools/ref_dropin.py (golden ensemble) and tests/test_dropin.py."""
import os
import re
import shutil
import subprocess

import numpy as np

NTS = 20            # timesteps 0 .. NTS-1 of the classic-mode test model (3.0 - 7.7 days)
KEEP = ("adata.txt", "transitiondata.txt", "phixsdata_v2.txt", "compositiondata.txt", "model.txt", "abundances.txt",
        "input.txt", "ratecoeff.dat", "bflist.dat")


def set_seed(rundir, seed):
    path = os.path.join(rundir, "input.txt")
    lines = open(path).read().splitlines()
    for i, line in enumerate(lines):
        if line.strip() and not line.strip().startswith("#"):
            lines[i] = f"{seed}".ljust(25) + "# pre_zseed"
            break
    open(path, "w").write("\n".join(lines) + "\n")


def clone_run(template, dst):
    os.makedirs(dst)
    for f in os.listdir(template):
        if f in KEEP or f.endswith("_lines.txt") or f.endswith("decays.txt") or f.startswith(("auger", "collion", "binding")):
            shutil.copy(os.path.join(template, f), dst)


def run(exe, rundir, seed, env=None):
    set_seed(rundir, seed)
    e = dict(os.environ)
    e.update(env or {})
    out = open(os.path.join(rundir, "stdout.txt"), "w")
    return subprocess.Popen([exe], cwd=rundir, env=e, stdout=out, stderr=subprocess.STDOUT)


def observables(rundir):
    """dict name -> float"""
    d = {}
    dep = np.loadtxt(os.path.join(rundir, "deposition.out"), comments="#", ndmin=2)
    assert dep.shape[0] >= NTS, dep.shape
    for ts in range(NTS):
        d[f"gammadep_ts{ts:02d}"] = dep[ts, 4]      # gamma_dep of the timestep: gamma transport + absorption
        d[f"gammadeppathint_ts{ts:02d}"] = dep[ts, 5]  # the same from the path-length estimator rpkt_emiss
    lc = np.loadtxt(os.path.join(rundir, "light_curve.out"), ndmin=2)
    d["lum_sum"] = lc[:NTS, 1].sum()
    d["lumcmf_sum"] = lc[NTS:2 * NTS, 1].sum() if lc.shape[0] >= 2 * NTS else lc[:NTS, 2].sum()
    # the plasma state that update_grid derives from the transport's estimators, last timestep
    pat = re.compile(rf"^timestep {NTS - 1} modelgridindex (\d+) titeration \d+ TR (\S+) Te (\S+) W (\S+) TJ (\S+) grey_depth (\S+) thick (\d+) nne (\S+)")
    ncell = 0
    for line in open(os.path.join(rundir, "estimators_0000.out")):
        m = pat.match(line)
        if m:
            mgi = int(m.group(1))
            d[f"TR_cell{mgi:02d}"] = float(m.group(2))
            d[f"Te_cell{mgi:02d}"] = float(m.group(3))
            d[f"W_cell{mgi:02d}"] = float(m.group(4))
            d[f"TJ_cell{mgi:02d}"] = float(m.group(5))
            d[f"thick_cell{mgi:02d}"] = float(m.group(7))
            d[f"nne_cell{mgi:02d}"] = float(m.group(8))
            ncell += 1
    assert ncell > 0, "no estimator lines of the last timestep"
    return d


def ztest(ref_rows, rows, names):
    """two-sample z per observable; entries without any scatter must agree to 1e-6 relative"""
    out = []
    ra, rb = np.asarray(ref_rows), np.asarray(rows)
    for j, n in enumerate(names):
        a, b = ra[:, j], rb[:, j]
        va, vb = a.var(ddof=1) / a.size, (b.var(ddof=1) / b.size if b.size > 1 else 0.0)
        if va + vb == 0:
            ok = np.isclose(a.mean(), b.mean(), rtol=1e-6, atol=0)
            out.append((n, a.mean(), b.mean(), 0.0 if ok else np.inf))
        else:
            out.append((n, a.mean(), b.mean(), (b.mean() - a.mean()) / np.sqrt(va + vb)))
    return out
