"""Reader/writer for the flat snapshot container written by oracle/refglue/hook.cc (and by the synthetic
workload generators): magic "ARTSNAP1", then records {name[64], dtype[8], uint64 count, raw data}.

A snapshot holds the three inputs of one update_packets call (update_packets.cc:234): `static.*` (ArtisStaticTables),
`cell.*` (ArtisCellState), `packets.pre`, and, when produced by the reference, its outputs `estim.*`, `packets.post`.
"""
import numpy as np

from . import abi

_DT = {"i4": np.int32, "f4": np.float32, "f8": np.float64, "i8": np.int64, "pkt": abi.PACKET_DTYPE}


def load(path):
    raw = np.fromfile(path, dtype=np.uint8)
    assert raw[:8].tobytes() == b"ARTSNAP1", "not a snapshot file"
    pos = 8
    out = {}
    n = raw.size
    while pos < n:
        name = raw[pos:pos + 64].tobytes().split(b"\0")[0].decode()
        dt = raw[pos + 64:pos + 72].tobytes().split(b"\0")[0].decode()
        count = int(raw[pos + 72:pos + 80].view(np.uint64)[0])
        pos += 80
        dtype = np.dtype(_DT[dt])
        nbytes = count * dtype.itemsize
        out[name] = raw[pos:pos + nbytes].view(dtype)
        pos += nbytes
    return out


def save(path, records):
    with open(path, "wb") as f:
        f.write(b"ARTSNAP1")
        for name, arr in records.items():
            arr = np.ascontiguousarray(arr)
            code = "pkt" if arr.dtype == abi.PACKET_DTYPE else {np.dtype(np.int32): "i4", np.dtype(np.float32): "f4",
                                                                np.dtype(np.float64): "f8",
                                                                np.dtype(np.int64): "i8"}[arr.dtype]
            f.write(name.encode().ljust(64, b"\0"))
            f.write(code.encode().ljust(8, b"\0"))
            f.write(np.uint64(arr.size).tobytes())
            f.write(arr.tobytes())


def section(snap, prefix, fields):
    """dict of python scalars / numpy arrays for one struct ('static', 'cell', 'estim')."""
    d = {}
    for kind, ctype, fname, _ in fields:
        key = f"{prefix}.{fname}"
        if key not in snap:
            # written before an ABI extension: every extension is a switch that is off (0 / no entries) in older dumps
            if kind == "S":
                d[fname] = 0
            continue
        d[fname] = snap[key][0].item() if kind == "S" else snap[key]
    return d


def static_tables(snap):
    return section(snap, "static", abi.STATIC_FIELDS)


def cell_state(snap):
    return section(snap, "cell", abi.CELL_FIELDS)


def estimators(snap):
    return section(snap, "estim", abi.ESTIM_FIELDS)
