#!/usr/bin/env python
"""Golden vectors for artis_gpu_bin_escaped_resolved (TEST INFRASTRUCTURE; needs /root/reference, oracle/_ref/classic_small).

Runs the unmodified reference (classic mode, small grid, default synthetic atomic data of make_run.py) to timestep NTS,
restarts it there with ARTIS_HOOK_SPECTRA_RES=1: the hook (oracle/refglue/hook.cc write_spectra_resolved) passes the
post-transport packets through the reference's own add_to_spec (do_emission_res) and, for each of the 100 escape-direction
bins, add_to_spec_res / add_to_lc_res. The fixture keeps the packets, the static tables of that run and the non-zero bins
of the five arrays (index + value: the dense arrays hold 13 million entries, a few thousand of them non-zero).

  python oracle/tools/ref_spectra_resolved.py tests/golden/spectra_resolved_classic.npz
"""
import os
import subprocess
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from artis_b200 import snapshot  # noqa: E402

NTS = 34


def main():
    out = sys.argv[1]
    run = os.path.join(REPO, "oracle", "_ref", "runs", "specres")
    exe = os.path.join(REPO, "oracle", "_ref", "classic_small", "sn3d_ref")
    subprocess.run(["make", "-C", os.path.join(REPO, "oracle"), "ref", "MODE=classic_small", "-j8"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["rm", "-rf", run])
    mk = [sys.executable, os.path.join(REPO, "oracle", "tools", "make_run.py"), "classic1d", run]
    subprocess.run(mk + ["--ftstep", str(NTS + 1)], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([exe], cwd=run, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    subprocess.run(mk + ["--itstep", str(NTS), "--ftstep", str(NTS + 1), "--resume"], check=True, stdout=subprocess.DEVNULL)
    env = dict(os.environ, ARTIS_HOOK_SNAPSHOT_TS=str(NTS), ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(NTS), ARTIS_HOOK_SPECTRA_RES="1")
    subprocess.run([exe], cwd=run, env=env, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    s = snapshot.load(os.path.join(run, f"snap_ts{NTS}.bin"))
    post = s["packets.post"]
    keep = {k: np.array(v) for k, v in s.items() if k.startswith(("static.", "meta."))}
    for name in ("emission", "trueemission", "absorption", "lc", "flux"):
        a = np.asarray(s[f"spectra.res_{name}"]).reshape(-1)
        nz = np.flatnonzero(a)
        keep[f"res_{name}_idx"] = nz.astype(np.int64)
        keep[f"res_{name}_val"] = a[nz]
        keep[f"res_{name}_size"] = np.array([a.size])
        print(name, a.size, "entries,", nz.size, "non-zero")
    print("packets", post.size, "escaped r-packets", int(((post["type"] == 32) & (post["escape_type"] == 11)).sum()))
    np.savez_compressed(out, packets_post=np.frombuffer(post.tobytes(), dtype=np.uint8), params=s["spectra.res_params"], **keep)
    print("wrote", out, os.path.getsize(out) / 1e6, "MB")


if __name__ == "__main__":
    main()
