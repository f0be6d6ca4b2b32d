"""calculate_cooling_rates on the device (artis_gpu_cooling_rates) against the reference's own values: the
`totalcooling` and `cooling_contrib_ion` of every golden fixture's cell state ARE what the reference's
calculate_cooling_rates (kpkt.cc:84-165, called at update_grid.cc:1205) left in grid::modelgrid for that timestep.
Cells the reference skips keep their stale values there, so the comparison is over the cells the transport can enter a
k-packet in (non-empty, not grey). Tolerance: the same terms in the same order -- rounding of libm calls only."""
import numpy as np
import pytest

import goldenlib
import statlib
from artis_b200.transport import Transport

TOL = {"oracle": 1e-13, "gpu": 1e-11}


def check(lib, name, mode):
    fx = goldenlib.load_fixture(name)
    st, cs = fx["static"], fx["cell"]
    T = Transport(lib)
    T.init(st)
    T.upload_cellstate(cs)
    total, contrib = T.cooling_rates()
    n, ni = int(st["npts_model"]), int(st["includedions"])
    good = statlib.good_cells(st, cs)
    want_c = np.asarray(cs["cooling_contrib_ion"]).reshape(n + 1, ni)[:n][good]
    got_c = contrib[:n][good]
    nz = want_c != 0
    assert nz.sum() > 10
    rel = np.abs(got_c[nz] / want_c[nz] - 1)
    want_t, got_t = np.asarray(cs["totalcooling"])[:n][good], total[:n][good]
    relt = np.abs(got_t / want_t - 1)
    print(name, mode, "cells", int(good.sum()), "max rel err per ion", rel.max(), "total", relt.max())
    assert rel.max() <= TOL[mode] and relt.max() <= TOL[mode]
    assert np.array_equal(got_c == 0, want_c == 0)
    # install: the k-packet path then runs from the device's own numbers -- same packets as from the host's
    pk0 = fx["pre"].copy()
    T.zero_estimators()
    T.update_packets(pk0)
    T.upload_cellstate(dict(cs, totalcooling=np.zeros_like(cs["totalcooling"]), cooling_contrib_ion=np.zeros_like(cs["cooling_contrib_ion"])))
    T.cooling_rates(install=True)
    pk1 = fx["pre"].copy()
    T.zero_estimators()
    T.update_packets(pk1)
    same = (pk0["type"] == pk1["type"]) & (pk0["where"] == pk1["where"]) & (pk0["interactions"] == pk1["interactions"])
    assert same.mean() > 0.97, same.mean()
    T.finalize()


@pytest.mark.parametrize("name", ["classic_nogrey_ts12.npz", "kilonova_lte_ts1.npz", "nebular_nlte_ts8.npz"])
def test_restatement_cooling_rates_match_reference(oracle_lib, name):
    check(oracle_lib, name, "oracle")


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["classic_nogrey_ts12.npz", "kilonova_lte_ts1.npz", "nebular_nlte_ts8.npz"])
def test_cuda_cooling_rates_match_reference(gpu_lib, name):
    check(gpu_lib, name, "gpu")
