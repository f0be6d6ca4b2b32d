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


def _section(snap, prefix, fields):
    """Fields that a fixture written before an ABI extension does not hold are zero / empty: every extension is a
    switch that is off (0 entries) in the older option sets."""
    d = {}
    for kind, _, fname, _ in fields:
        key = prefix + fname
        if key in snap:
            d[fname] = snap[key][0].item() if kind == "S" else snap[key]
        elif kind == "S":
            d[fname] = 0
    return d


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
    static = _section(snap, "static.", abi.STATIC_FIELDS)
    cell = _section(snap, "cell.", abi.CELL_FIELDS)
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
