#!/usr/bin/env python
"""Golden vectors for artis_gpu_bin_escaped (TEST INFRASTRUCTURE; needs a built oracle/_ref/<mode>/sn3d_ref).

Runs the shim-built, unmodified reference over one timestep from a checkpoint with ARTIS_HOOK_SPECTRA=1: the hook
(oracle/refglue/hook.cc write_spectra) passes the post-transport packets through the reference's own add_to_lc_res /
add_to_spec_res and stores the light curves, the spectrum and the Stokes spectra next to the packets. The fixture keeps
the packets (raw bytes) and those arrays; the static tables come from the ensemble fixture of the same run directory.

  python oracle/tools/ref_spectra.py <rundir prepared to restart at ts N> <N> <out.npz> [--mode classic_1d_small]
"""
import argparse
import os
import subprocess
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from artis_b200 import snapshot  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rundir")
    ap.add_argument("nts", type=int)
    ap.add_argument("out")
    ap.add_argument("--mode", default="classic_1d_small")
    a = ap.parse_args()
    exe = os.path.join(REPO, "oracle", "_ref", a.mode, "sn3d_ref")
    env = dict(os.environ, ARTIS_HOOK_SNAPSHOT_TS=str(a.nts), ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(a.nts),
               ARTIS_HOOK_SPECTRA="1")
    subprocess.run([exe], cwd=a.rundir, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
    s = snapshot.load(os.path.join(a.rundir, f"snap_ts{a.nts}.bin"))
    post = s["packets.post"]
    flux, stokes = np.asarray(s["spectra.flux"]), np.asarray(s["spectra.stokes"])
    print("packets", post.size, "escaped", int((post["type"] == 32).sum()), "nonzero flux bins", int((flux != 0).sum()),
          "nonzero lc bins", int((np.asarray(s["spectra.lc"]) != 0).sum()))
    np.savez_compressed(a.out, packets_post=np.frombuffer(post.tobytes(), dtype=np.uint8), params=s["spectra.params"],
                        lc=s["spectra.lc"], flux=flux, stokes=stokes, delta_freq=s["spectra.delta_freq"])
    print("wrote", a.out, os.path.getsize(a.out) / 1e6, "MB")


if __name__ == "__main__":
    main()
