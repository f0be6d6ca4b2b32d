#!/usr/bin/env python
"""Golden ensemble for the drop-in test (TEST INFRASTRUCTURE; needs /root/reference).

Prepares oracle/_ref/runs/dropin (classic-mode test model, small grid, synthetic atomic data; the template run directory
that tests/test_dropin.py clones on the GPU box) and runs the UNMODIFIED reference K times over timesteps 0..NTS-1 with K
seeds; the observables of tests/dropinlib.py of every run go to tests/golden/dropin_classic.npz.

  python oracle/tools/ref_dropin.py [--seeds 24] [--template-only]
"""
import argparse
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(REPO, "tests"))
import dropinlib  # noqa: E402

TEMPLATE = os.path.join(REPO, "oracle", "_ref", "runs", "dropin")
EXE = os.path.join(REPO, "oracle", "_ref", "classic_small", "sn3d_ref")


def make_template():
    subprocess.run(["make", "-C", os.path.join(REPO, "oracle"), "ref", "MODE=classic_small", "-j8"], check=True, stdout=subprocess.DEVNULL)
    shutil.rmtree(TEMPLATE, ignore_errors=True)
    subprocess.run([sys.executable, os.path.join(REPO, "oracle", "tools", "make_run.py"), "classic1d", TEMPLATE, "--ftstep",
                    str(dropinlib.NTS)], check=True, stdout=subprocess.DEVNULL)
    # ratecoeff.dat / bflist.dat are written by the first run of a directory: make them part of the template
    p = dropinlib.run(EXE, TEMPLATE, 1281360349, env={"ARTIS_HOOK_STOP_AFTER_TS": "0"})
    p.wait()
    for f in os.listdir(TEMPLATE):
        if f.endswith(".out") or f.endswith(".tmp") or f.startswith(("output_", "stdout", "hook_")):
            os.remove(os.path.join(TEMPLATE, f))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=24)
    ap.add_argument("--template-only", action="store_true")
    a = ap.parse_args()
    make_template()
    if a.template_only:
        return
    rows, names = [], None
    with tempfile.TemporaryDirectory(prefix="dropin_ref_") as tmp:
        for k0 in range(0, a.seeds, 8):
            procs = []
            for k in range(k0, min(a.seeds, k0 + 8)):
                d = os.path.join(tmp, f"r{k}")
                dropinlib.clone_run(TEMPLATE, d)
                procs.append((d, dropinlib.run(EXE, d, 910001 + 7919 * k)))
            for d, p in procs:
                assert p.wait() == 0, open(os.path.join(d, "stdout.txt")).read()[-2000:]
                o = dropinlib.observables(d)
                if names is None:
                    names = sorted(o)
                rows.append([o[n] for n in names])
                print(os.path.basename(d), "lum_sum", o["lum_sum"], flush=True)
    out = os.path.join(REPO, "tests", "golden", "dropin_classic.npz")
    np.savez_compressed(out, ref_rows=np.array(rows), names=np.array(names), nts=dropinlib.NTS)
    print("wrote", out, len(names), "observables x", len(rows), "runs")


if __name__ == "__main__":
    main()
