#!/usr/bin/env python
"""Golden fixture generator (TEST INFRASTRUCTURE; needs /root/reference and a built oracle/_ref/<mode>/sn3d_ref).

Runs the shim-built, unmodified reference K times over the SAME timestep from the SAME checkpoint with K different
random seeds (the reference re-seeds its generator on restart, input.cc:1908) and stores, for a tests/golden fixture:
  - the inputs of that update_packets call (static tables, cell state, packets before) exactly as the hook
    (oracle/refglue/hook.cc) flattened them, and
  - the K summary vectors of tests/statlib.summary() computed from the reference's outputs.
The parity tests run the CUDA path on the stored inputs with their own seeds and z-test every summary entry.

  python oracle/tools/ref_ensemble.py <rundir prepared to restart at ts N> <N> <out.npz> [--seeds 24] [--mode classic_small]
"""
import argparse
import os
import subprocess
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
from artis_b200 import snapshot  # noqa: E402
import statlib  # noqa: E402


def set_seed(rundir, seed):
    path = os.path.join(rundir, "input.txt")
    lines = open(path).read().splitlines()
    for i, line in enumerate(lines):
        if line.strip() and not line.strip().startswith("#"):
            lines[i] = f"{seed}".ljust(25) + "# pre_zseed (set by ref_ensemble.py)"
            break
    open(path, "w").write("\n".join(lines) + "\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rundir")
    ap.add_argument("nts", type=int)
    ap.add_argument("out")
    ap.add_argument("--seeds", type=int, default=24)
    ap.add_argument("--mode", default="classic_small")
    ap.add_argument("--probes", type=int, default=1500)
    a = ap.parse_args()
    exe = os.path.join(REPO, "oracle", "_ref", a.mode, "sn3d_ref")
    rows, names, inputs = [], None, None
    for k in range(a.seeds):
        set_seed(a.rundir, 700001 + 7919 * k)
        env = dict(os.environ, ARTIS_HOOK_SNAPSHOT_TS=str(a.nts), ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(a.nts))
        if k == 0:
            env["ARTIS_HOOK_PROBES"] = str(a.probes)  # deterministic sub-kernel probes: inputs + the reference's outputs
        subprocess.run([exe], cwd=a.rundir, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
        s = snapshot.load(os.path.join(a.rundir, f"snap_ts{a.nts}.bin"))
        st, cs, est = snapshot.static_tables(s), snapshot.cell_state(s), snapshot.estimators(s)
        d = statlib.summary(st, cs, est, s["packets.post"], s["packets.pre"])
        if names is None:
            names = list(d)
            inputs = {k2: np.array(v) for k2, v in s.items()
                      if k2.startswith(("static.", "cell.", "meta.", "probe.")) or k2 == "packets.pre"}
        else:
            assert np.array_equal(inputs["packets.pre"], s["packets.pre"]), "restart did not reproduce the start state"
        rows.append([d[nm] for nm in names])
        print(k, "interactions", d["interactions"], flush=True)
    # packets.pre is a structured array: store its raw bytes
    for key in ("packets.pre", "probe.packets", "probe.gpackets"):
        inputs[key] = np.frombuffer(inputs[key].tobytes(), dtype=np.uint8)
    np.savez_compressed(a.out, ref_rows=np.array(rows), ref_names=np.array(names), nts=a.nts, **inputs)
    print("wrote", a.out, os.path.getsize(a.out) / 1e6, "MB")


if __name__ == "__main__":
    main()
