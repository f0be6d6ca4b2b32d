"""Loader for the tests/golden/*.npz fixtures written by oracle/tools/ref_ensemble.py, and the ensemble runner used by
the parity tests."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
from artis_b200 import abi  # noqa: E402
import statlib  # noqa: E402

GOLDEN = os.path.join(HERE, "golden")


def load_fixture(name):
    z = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    snap = {}
    for k in z.files:
        if k.startswith(("static.", "cell.", "meta.")):
            snap[k] = z[k]
    snap["packets.pre"] = np.frombuffer(z["packets.pre"].tobytes(), dtype=abi.PACKET_DTYPE).copy()
    probe = {k[6:]: z[k] for k in z.files if k.startswith("probe.")}
    for key in ("packets", "gpackets"):
        if key in probe:
            probe[key] = np.frombuffer(probe[key].tobytes(), dtype=abi.PACKET_DTYPE).copy()
    static = {f[2]: (snap["static." + f[2]][0].item() if f[0] == "S" else snap["static." + f[2]]) for f in abi.STATIC_FIELDS}
    cell = {f[2]: (snap["cell." + f[2]][0].item() if f[0] == "S" else snap["cell." + f[2]]) for f in abi.CELL_FIELDS}
    return {"static": static, "cell": cell, "pre": snap["packets.pre"], "ref_rows": z["ref_rows"],
            "names": [str(x) for x in z["ref_names"]], "nts": int(z["nts"]), "probe": probe}


def run_ensemble(T, fx, nseeds, seed0=424242):
    """T: artis_b200.transport.Transport. Returns rows in the order of fx['names']."""
    rows = []
    for k in range(nseeds):
        T.init(fx["static"], seed=seed0 + 104729 * k)
        T.upload_cellstate(fx["cell"])
        T.zero_estimators()
        pk = fx["pre"].copy()
        T.update_packets(pk)
        est = T.download_estimators()
        d = statlib.summary(fx["static"], fx["cell"], est, pk, fx["pre"])
        rows.append([d[n] for n in fx["names"]])
    return np.array(rows)
