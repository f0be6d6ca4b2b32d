"""ctypes mirror of include/artis_gpu.h, generated at import time from include/artis_gpu_fields.def.

Host-side plumbing only: it builds the three table structs (ArtisStaticTables, ArtisCellState, ArtisEstimators)
out of numpy arrays and declares the argument types of the C-ABI entry points. The packet record mirrors
`struct packet` of the reference (packet.h:28-73; 304 bytes).
"""
import ctypes as C
import os
import re

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INCLUDE = os.path.join(REPO, "include")

_CT = {"int": (C.c_int, np.int32), "float": (C.c_float, np.float32), "double": (C.c_double, np.float64),
       "long long": (C.c_longlong, np.int64)}

# struct packet (packet.h:28-73) as a numpy structured dtype with explicit C padding
PACKET_DTYPE = np.dtype({
    "names": ["where", "type", "last_cross", "interactions", "nscatterings", "last_event", "pos", "dir", "e_cmf",
              "e_rf", "nu_cmf", "nu_rf", "next_trans", "emissiontype", "em_pos", "em_time", "prop_time",
              "absorptiontype", "trueemissiontype", "trueem_time", "absorptionfreq", "absorptiondir", "stokes",
              "pol_dir", "tdecay", "escape_type", "escape_time", "scat_count", "number",
              "originated_from_particlenotgamma", "pellet_decaytype", "pellet_nucindex", "trueemissionvelocity",
              "ma_element", "ma_ion", "ma_level", "ma_activatingline"],
    "formats": ["i4", "i4", "i4", "i4", "i4", "i4", ("f8", 3), ("f8", 3), "f8", "f8", "f8", "f8", "i4", "i4",
                ("f8", 3), "i4", "f8", "i4", "i4", "i4", "f8", ("f8", 3), ("f8", 3), ("f8", 3), "f8", "i4", "i4", "i4",
                "i4", "u1", "i4", "i4", "f4", "i4", "i4", "i4", "i4"],
    "offsets": [0, 4, 8, 12, 16, 20, 24, 48, 72, 80, 88, 96, 104, 108, 112, 136, 144, 152, 156, 160, 168, 176, 200,
                224, 248, 256, 260, 264, 268, 272, 276, 280, 284, 288, 292, 296, 300],
    "itemsize": 304,
})

ABI_VERSION = int(re.search(r"#define ARTIS_GPU_ABI_VERSION (\d+)", open(os.path.join(INCLUDE, "artis_gpu.h")).read()).group(1))

ERR_UNSUPPORTED = 6  # ARTIS_ERR_UNSUPPORTED
CTR_ROW_FLUSHES, CTR_ROW_UNITS = 91, 92  # ARTIS_CTR_ROW_* (include/artis_gpu.h)
TYPE_ESCAPE, TYPE_PELLET, TYPE_GAMMA, TYPE_RPKT, TYPE_KPKT, TYPE_MA = 32, 100, 10, 11, 12, 13
TYPE_NTLEPTON, TYPE_NT_PREDEPOSIT, TYPE_PRE_KPKT, TYPE_GAMMA_KPKT = 20, 21, 120, 121


def _parse_fields(macro_name):
    """Returns [(kind, ctype, name, count_expr)] for one ARTIS_*_FIELDS list of the .def file."""
    text = open(os.path.join(INCLUDE, "artis_gpu_fields.def")).read()
    start = text.index("#define " + macro_name)
    body = []
    for line in text[start:].splitlines()[1:]:
        stripped = re.sub(r"/\*.*?\*/", "", line).rstrip()
        cont = stripped.endswith("\\")
        body.append(stripped.rstrip("\\"))
        if not cont:
            break
    fields = []
    for line in body:
        m = re.match(r"\s*S\(\s*([\w ]+?)\s*,\s*(\w+)\s*\)", line)
        if m:
            fields.append(("S", m.group(1), m.group(2), None))
            continue
        m = re.match(r"\s*A\(\s*([\w ]+?)\s*,\s*(\w+)\s*,\s*(.+)\)\s*$", line)
        if m:
            fields.append(("A", m.group(1), m.group(2), m.group(3).strip()))
    return fields


STATIC_FIELDS = _parse_fields("ARTIS_STATIC_FIELDS")
CELL_FIELDS = _parse_fields("ARTIS_CELL_FIELDS")
ESTIM_FIELDS = _parse_fields("ARTIS_ESTIM_FIELDS")


def _make_struct(name, fields):
    members = []
    for kind, ctype, fname, _ in fields:
        ct = _CT[ctype][0]
        members.append((fname, ct if kind == "S" else C.POINTER(ct)))
    return type(name, (C.Structure,), {"_fields_": members})


ArtisStaticTables = _make_struct("ArtisStaticTables", STATIC_FIELDS)
ArtisCellState = _make_struct("ArtisCellState", CELL_FIELDS)
ArtisEstimators = _make_struct("ArtisEstimators", ESTIM_FIELDS)


class _Scalars:
    def __init__(self, d):
        self.__dict__.update(d)


def count_of(expr, scalars):
    """Evaluates a count expression of the .def file (`s->x` syntax, C ternary) over a dict of scalars."""
    e = expr.replace("s->", "s.")
    m = re.match(r"\((.+)\?(.+):(.+)\)$", e)
    s = _Scalars(scalars)
    if m:
        return int(eval(m.group(2), {"s": s}) if eval(m.group(1), {"s": s}) else eval(m.group(3), {"s": s}))
    return int(eval(e, {"s": s}))


def build_struct(cls, fields, data, writable=False):
    """data: dict name -> python scalar / numpy array. Returns (struct, keepalive list).
    Arrays are converted to the exact dtype and checked against the .def count expression."""
    st = cls()
    keep = []
    scalars = {f[2]: data[f[2]] for f in fields if f[0] == "S"}
    for kind, ctype, fname, expr in fields:
        ct, npdt = _CT[ctype]
        if kind == "S":
            v = data[fname]
            setattr(st, fname, float(v) if ctype in ("float", "double") else int(v))
        else:
            n = count_of(expr, scalars)
            arr = data.get(fname)
            if arr is None:
                arr = np.zeros(max(n, 1), dtype=npdt)
            arr = np.ascontiguousarray(arr, dtype=npdt)
            if arr.size < n:
                raise ValueError(f"{cls.__name__}.{fname}: have {arr.size} elements, need {n}")
            if writable and not arr.flags.writeable:
                arr = arr.copy()
            keep.append(arr)
            setattr(st, fname, arr.ctypes.data_as(C.POINTER(ct)))
            data[fname] = arr
    return st, keep


def empty_estimators(npts_model, nelements, max_nions, nlines, nradfieldbins=0, nbfrate=0, nvstokes=0, nvgrid=0):
    d = {"npts_model": npts_model, "nelements": nelements, "max_nions": max_nions, "nlines": nlines,
         "nradfieldbins": nradfieldbins, "nbfrate": nbfrate, "nvstokes": nvstokes, "nvgrid": nvgrid}
    for kind, ctype, fname, expr in ESTIM_FIELDS:
        if kind == "A":
            d[fname] = np.zeros(max(count_of(expr, d), 1), dtype=_CT[ctype][1])
    return d


def declare(lib):
    """Sets argtypes/restype for every entry point of include/artis_gpu.h."""
    P = C.POINTER
    pd, pi, vp = P(C.c_double), P(C.c_int), C.c_void_p
    sig = {
        "artis_gpu_init": [P(ArtisStaticTables), C.c_int, C.c_int, C.c_int, C.c_ulonglong],
        "artis_gpu_finalize": [],
        "artis_gpu_last_error": [C.c_char_p, C.c_size_t],
        "artis_gpu_upload_cellstate": [P(ArtisCellState)],
        "artis_gpu_zero_estimators": [],
        "artis_gpu_update_packets": [C.c_int, C.c_int, vp, C.c_int],
        "artis_gpu_packets_upload": [vp, C.c_int],
        "artis_gpu_transport_resident": [C.c_int],
        "artis_gpu_packets_download": [vp, C.c_int],
        "artis_gpu_last_timings": [pd, C.c_int],
        "artis_gpu_set_tuning": [C.c_int, C.c_int],
        "artis_gpu_scheduler_geometry": [pi, C.c_int],
        "artis_gpu_comm_unique_id": [C.c_char_p],
        "artis_gpu_comm_init": [C.c_char_p, C.c_int, C.c_int],
        "artis_gpu_reduce_estimators": [C.c_int],
        "artis_gpu_download_estimators": [P(ArtisEstimators)],
        "artis_gpu_estimator_block": [P(vp), P(C.c_size_t)],
        "artis_gpu_bin_escaped": [C.c_double, C.c_int, C.c_int, C.c_int, pd, pd, pd],
        "artis_gpu_packets_download_escaped": [vp, C.c_int, C.c_int, pi],
        "artis_gpu_bin_escaped_resolved": [C.c_double, C.c_int, C.c_int, pd, pd, pd, pd, pd, pd],
        "artis_gpu_normalise_estimators": [pd, C.c_int, C.c_double, pd, pd],
        "artis_gpu_cooling_rates": [pd, pd, C.c_int],
        "artis_gpu_upload_estimators": [P(ArtisEstimators)],
        "artis_gpu_packets_write_temp": [C.c_char_p],
        "artis_gpu_packets_read_temp": [C.c_char_p, C.c_int],
        "artis_gpu_packets_write_text": [C.c_char_p],
        "artis_gpu_pack_escaped_noalloc": [vp, C.c_int, vp, C.c_int, vp, vp],
        "artis_gpu_probe_boundary_cross": [vp, C.c_int, pd, pi, pi],
        "artis_gpu_probe_kappa_cont": [vp, C.c_int, pd],
        "artis_gpu_probe_closest_transition": [pd, pi, C.c_int, pi],
        "artis_gpu_probe_get_event": [vp, C.c_int, pd, pd, pd, pi, pi, pi],
        "artis_gpu_probe_macroatom_rates": [pi, pi, pi, pi, C.c_int, C.c_int, pd],
        "artis_gpu_probe_cooling_ion": [pi, pi, pi, C.c_int, pd],
        "artis_gpu_probe_gamma_kappa": [vp, C.c_int, pd],
        "artis_gpu_probe_levelpop": [pi, pi, pi, pi, C.c_int, pd],
        "artis_gpu_probe_select_continuum_nu": [pi, pi, pi, pi, P(C.c_float), pd, C.c_int, pd],
    }
    for name, args in sig.items():
        if name in CUDA_ONLY_SYMBOLS and not hasattr(lib, name):
            continue  # the oracle's host build of the ABI (test infrastructure) has no device-pointer entry points
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    return sig


CUDA_ONLY_SYMBOLS = ("artis_gpu_pack_escaped_noalloc",)

EXPORTED_SYMBOLS = [
    "artis_gpu_init", "artis_gpu_finalize", "artis_gpu_last_error", "artis_gpu_upload_cellstate",
    "artis_gpu_zero_estimators", "artis_gpu_update_packets", "artis_gpu_packets_upload",
    "artis_gpu_transport_resident", "artis_gpu_packets_download", "artis_gpu_last_timings", "artis_gpu_set_tuning",
    "artis_gpu_comm_unique_id", "artis_gpu_comm_init", "artis_gpu_reduce_estimators",
    "artis_gpu_download_estimators", "artis_gpu_estimator_block", "artis_gpu_probe_boundary_cross",
    "artis_gpu_probe_kappa_cont", "artis_gpu_probe_closest_transition", "artis_gpu_probe_get_event",
    "artis_gpu_probe_macroatom_rates", "artis_gpu_probe_cooling_ion", "artis_gpu_probe_gamma_kappa",
    "artis_gpu_probe_levelpop", "artis_gpu_launch_count", "artis_gpu_table_bytes", "artis_gpu_bin_escaped",
    "artis_gpu_packets_download_escaped", "artis_gpu_pack_escaped_noalloc", "artis_gpu_scheduler_geometry", "artis_gpu_probe_select_continuum_nu",
    "artis_gpu_packets_write_temp", "artis_gpu_packets_read_temp", "artis_gpu_packets_write_text", "artis_gpu_bin_escaped_resolved",
    "artis_gpu_normalise_estimators", "artis_gpu_upload_estimators", "artis_gpu_cooling_rates",
]
