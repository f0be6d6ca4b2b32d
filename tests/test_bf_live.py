"""Live continua (rpkt.cuh bf_is_live, CellTables.bf_live_*): the bound-free opacity sums run over the per-cell list of the
continua whose term can be non-zero instead of over every continuum of the list. The terms left out are exact zeros (the
stimulated-recombination factor is clamped at zero over the whole frequency range of the continuum's cross-section table, or
the level is empty), so kappa_cont may only change by the rounding of the additions: checked here against the sums over every
continuum (ARTIS_BF_LIVE=0) on real packet states, to 1e-13 relative where the reference probes allow 1e-10, and through
whole timesteps (same packet histories). The reference's own kappa_cont is compared with both by the probes of
test_oracle.py / test_gpu_parity.py."""
import os

import numpy as np
import pytest

import goldenlib
from artis_b200 import abi
from artis_b200.transport import Transport

FIXTURES = ["classic_nogrey_ts12.npz", "kilonova_lte_ts1.npz", "nebular_nlte_ts8.npz"]


def ready(lib, fx, live, seed=11):
    old = os.environ.get("ARTIS_BF_LIVE")
    os.environ["ARTIS_BF_LIVE"] = "1" if live else "0"
    try:
        T = Transport(lib)
        T.init(fx["static"], seed=seed)
    finally:
        if old is None:
            os.environ.pop("ARTIS_BF_LIVE", None)
        else:
            os.environ["ARTIS_BF_LIVE"] = old
    T.upload_cellstate(fx["cell"])
    T.zero_estimators()
    return T


def check(lib, name):
    fx = goldenlib.load_fixture(name)
    r = fx["pre"][fx["pre"]["type"] == abi.TYPE_RPKT].copy()
    out = []
    for live in (True, False):
        T = ready(lib, fx, live)
        kap = T.probe_kappa_cont(r)
        pk = fx["pre"].copy()
        T.update_packets(pk)
        c = T.download_estimators()["counters"]
        out.append((kap, pk, c))
        T.finalize()
    (ka, pa, ca), (kb, pb, cb) = out
    ok = np.isfinite(ka).all(axis=1) & np.isfinite(kb).all(axis=1)  # r-packets in grey / empty cells have no table row
    assert np.array_equal(np.isfinite(ka), np.isfinite(kb)) and ok.sum() > 50
    ka, kb = ka[ok], kb[ok]
    assert (kb[:, 3] > 0).any(), "the fixture does not exercise a non-zero bound-free opacity"
    scale = np.maximum(np.abs(kb[:, 0]), 1e-300)
    err = np.abs(ka - kb).max(axis=1) / scale
    print(name, "kappa_cont live vs every continuum: max rel", err.max(), "| terms evaluated", int(ca[38]), "of", int(ca[51]),
          "listed; every continuum:", int(cb[38]))
    assert err.max() < 1e-13
    assert np.array_equal(ka[:, 3] > 0, kb[:, 3] > 0)       # a sum is zero in one exactly when it is zero in the other
    assert ca[38] <= ca[51] and cb[38] == cb[51]
    same = (pa["type"] == pb["type"]) & (pa["where"] == pb["where"]) & (pa["interactions"] == pb["interactions"])
    assert same.mean() > 0.99, same.mean()


@pytest.mark.parametrize("name", FIXTURES)
def test_restatement_live_continua_leave_kappa_alone(oracle_lib, name):
    check(oracle_lib, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", FIXTURES)
def test_cuda_live_continua_leave_kappa_alone(gpu_lib, name):
    check(gpu_lib, name)
