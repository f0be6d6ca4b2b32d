"""VPKT_ON in the barrier-free scheduler: events record virtual-packet requests, k_vpkt_trace traces the list when the
scheduler has finished (vpkt.cuh). Checked here: a request list so small that it fills many times per timestep
(the scheduler drains, the list is traced and started afresh, the unfinished packets are launched again) gives every packet
its full timestep, conserves the counters and agrees with the one-pass run -- virtual-packet count, radiation field --
within Monte Carlo noise (a packet in flight at a restart redraws a pending macro-atom action, as with the row arena)."""
import os

import numpy as np
import pytest

import goldenlib
import statlib
from artis_b200 import abi
from artis_b200.transport import Transport

pytestmark = pytest.mark.gpu
CTR_VPKT_REQUESTS, CTR_VPKT_RESTARTS = 57, 58


def run(lib, fx, soft_mb=None, seed=31):
    old = os.environ.get("ARTIS_VPKT_REQ_SOFT_MB")
    if soft_mb is not None:
        os.environ["ARTIS_VPKT_REQ_SOFT_MB"] = str(soft_mb)
    try:
        T = Transport(lib)
        T.init(fx["static"], seed=seed)
    finally:
        if old is None:
            os.environ.pop("ARTIS_VPKT_REQ_SOFT_MB", None)
        else:
            os.environ["ARTIS_VPKT_REQ_SOFT_MB"] = old
    T.upload_cellstate(fx["cell"])
    T.zero_estimators()
    pk = fx["pre"].copy()
    T.update_packets(pk)
    est = T.download_estimators()
    T.finalize()
    return pk, est


def test_request_list_restarts_keep_the_timestep_whole(gpu_lib):
    fx = goldenlib.load_fixture("classic_vpkt_ts30.npz")
    p1, e1 = run(gpu_lib, fx)
    c1 = e1["counters"]
    assert c1[CTR_VPKT_REQUESTS] > 0 and c1[CTR_VPKT_RESTARTS] == 0
    nobs = int(fx["static"]["vpkt_nobs"])
    assert e1["ts_scalars"][9] <= c1[CTR_VPKT_REQUESTS] * nobs  # at most one virtual packet per request and observer
    p2, e2 = run(gpu_lib, fx, soft_mb=0.02)  # 160 requests
    c2 = e2["counters"]
    print("requests", int(c1[CTR_VPKT_REQUESTS]), "/", int(c2[CTR_VPKT_REQUESTS]), "list restarts", int(c2[CTR_VPKT_RESTARTS]))
    assert c2[CTR_VPKT_RESTARTS] >= 2
    st = fx["static"]
    t2 = st["ts_start"][fx["nts"]] + st["ts_width"][fx["nts"]]
    assert np.array_equal(p2["number"], fx["pre"]["number"])
    assert ((p2["type"] == abi.TYPE_ESCAPE) | (p2["prop_time"] == t2)).all()
    assert c2[45] == p2["interactions"].sum()
    good = statlib.good_cells(st, fx["cell"])
    n = st["npts_model"]
    assert np.isclose(e2["J"][:n][good].sum(), e1["J"][:n][good].sum(), rtol=3e-2)
    assert np.isclose(e2["ts_scalars"][9], e1["ts_scalars"][9], rtol=0.05)      # virtual packets created
    assert np.isclose(c2[CTR_VPKT_REQUESTS], c1[CTR_VPKT_REQUESTS], rtol=0.05)
