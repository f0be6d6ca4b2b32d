"""Estimator normalisation on the device (artis_gpu_normalise_estimators) against the reference's own
radfield::normalise_J / normalise_nuJ / get_T_J_from_J (radfield.cc:1434-1475), the ffheating / colheating scaling of
update_grid (update_grid.cc:1041-1134) and normalise_grey (grey_emissivities.cc:124-139).

Golden vector: tests/golden/norm_classic.npz by oracle/tools/ref_norm.py (static tables, cell state and RAW estimators of
timestep 14 of a classic-mode reference run, and what the reference's functions made of them). The raw estimators go to the
device through artis_gpu_upload_estimators; per-cell values agree to rounding (TOL: a pow() and three multiplications), the
gamma_dep_pathint sum to TOL_SUM (the device adds the cells in no fixed order)."""
import os

import numpy as np
import pytest

import goldenlib
from artis_b200 import abi
from artis_b200.transport import Transport

TOL = 1e-13
TOL_SUM = 1e-12


def load():
    z = np.load(os.path.join(goldenlib.GOLDEN, "norm_classic.npz"))
    snap = {k: z[k] for k in z.files}
    static = goldenlib._section(snap, "static.", abi.STATIC_FIELDS)
    cell = goldenlib._section(snap, "cell.", abi.CELL_FIELDS)
    est = goldenlib._section(snap, "estim.", abi.ESTIM_FIELDS)
    norm = {k[5:]: z[k] for k in z.files if k.startswith("norm.")}
    return static, cell, est, norm


def check(lib):
    static, cell, est, norm = load()
    gamma_ref, nprocs, mintemp, maxtemp = norm["scalars"]
    assert mintemp == static["opt_mintemp"]
    T = Transport(lib)
    T.init(static)
    T.upload_cellstate(cell)
    T.zero_estimators()
    T.upload_estimators(est)
    back = T.download_estimators()           # the upload is the inverse of the download
    for k in ("J", "nuJ", "ffheatingestimator", "colheatingestimator", "rpkt_emiss", "ts_scalars"):
        n = min(back[k].size, np.asarray(est[k]).size)
        assert np.array_equal(back[k][:n], np.asarray(est[k])[:n]), k
    for rep in range(2):                      # the device estimators stay raw: a second call gives the same answer
        out, gamma = T.normalise_estimators(norm["vol_init"], int(nprocs), float(maxtemp))
        nm = int(static["npts_model"])
        assert out.shape == (nm, 6)
        for col, name in enumerate(("J", "nuJ", "ff", "col", "rpkt_emiss", "TJ")):
            w, g = norm[name][:nm], out[:, col]
            assert (w != 0).sum() > 5, name
            assert np.array_equal(g != 0, w != 0), f"{name}: different set of non-zero cells"
            nz = w != 0
            rel = np.abs(g[nz] - w[nz]) / np.abs(w[nz])
            assert rel.max() <= TOL, (name, rel.max())
        assert abs(gamma / gamma_ref - 1) <= TOL_SUM, (gamma, gamma_ref)
    # clamps of get_T_J_from_J: a cell with a huge J sits at MAXTEMP, one with a tiny J at MINTEMP, NaN falls back to T_R
    good = np.flatnonzero(norm["J"][:nm] != 0)[:3]
    e2 = {k: np.array(v, copy=True) if isinstance(v, np.ndarray) else v for k, v in est.items()}
    e2["J"][good[0]] = 1e300
    e2["J"][good[1]] = 1e-300
    e2["J"][good[2]] = np.nan
    T.upload_estimators(e2)
    out, _ = T.normalise_estimators(norm["vol_init"], int(nprocs), float(maxtemp))
    assert out[good[0], 5] == maxtemp and out[good[1], 5] == mintemp
    assert out[good[2], 5] == np.float32(cell["TR"][good[2]])
    with pytest.raises(Exception):
        T.normalise_estimators(norm["vol_init"], 0, float(maxtemp))
    T.finalize()


def test_restatement_normalise_matches_reference(oracle_lib):
    check(oracle_lib)


@pytest.mark.gpu
def test_cuda_normalise_matches_reference(gpu_lib):
    check(gpu_lib)
