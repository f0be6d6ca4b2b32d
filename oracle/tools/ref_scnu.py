#!/usr/bin/env python
"""Golden vectors for select_continuum_nu (TEST INFRASTRUCTURE; needs /root/reference and oracle/_ref/classic_small).

Runs the unmodified reference (classic mode, small grid, synthetic atomic data with the default recipe of make_run.py) for
a few timesteps, restarts it at the last one with the hook's probes switched on, and keeps what the probe
`probe.scnu_*` of oracle/refglue/hook.cc recorded: the reference's own select_continuum_nu (ratecoeff.cc:628-684, GSL qag
through the stand-in) called with an injected random number for a sample of continua, two electron temperatures and eight
draws -- together with the static tables and cell state of that run, so that the device probe can be initialised with
exactly the atomic data the reference had.

  python oracle/tools/ref_scnu.py tests/golden/scnu_classic.npz
"""
import os
import subprocess
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from artis_b200 import snapshot  # noqa: E402

NTS = 4


def main():
    out = sys.argv[1]
    run = os.path.join(REPO, "oracle", "_ref", "runs", "scnu")
    exe = os.path.join(REPO, "oracle", "_ref", "classic_small", "sn3d_ref")
    subprocess.run(["make", "-C", os.path.join(REPO, "oracle"), "ref", "MODE=classic_small", "-j8"], check=True,
                   stdout=subprocess.DEVNULL)
    subprocess.run(["rm", "-rf", run])
    mk = [sys.executable, os.path.join(REPO, "oracle", "tools", "make_run.py"), "classic1d", run]
    subprocess.run(mk + ["--ftstep", str(NTS + 1)], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([exe], cwd=run, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    subprocess.run(mk + ["--itstep", str(NTS), "--ftstep", str(NTS + 1), "--resume"], check=True, stdout=subprocess.DEVNULL)
    env = dict(os.environ, ARTIS_HOOK_SNAPSHOT_TS=str(NTS), ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(NTS),
               ARTIS_HOOK_PROBES="8", ARTIS_HOOK_SKIP_POST="1")
    subprocess.run([exe], cwd=run, env=env, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    s = snapshot.load(os.path.join(run, f"snap_ts{NTS}.bin"))
    keep = {k: np.array(v) for k, v in s.items() if k.startswith(("static.", "cell.", "meta.", "probe.scnu_"))}
    np.savez_compressed(out, nts=NTS, **keep)
    print("wrote", out, os.path.getsize(out) / 1e6, "MB;", keep["probe.scnu_nu"].size, "probes")


if __name__ == "__main__":
    main()
