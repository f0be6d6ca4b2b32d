"""artis_gpu_update_packets streams an ensemble that does not fit the device through it in batches (BASELINE.json config 5:
1e9 packets; engine.inc update_batch_packets). Packets are independent within a timestep -- every packet owns its random
stream -- so a batched call must return the packets of the one-piece call bit for bit and estimators that differ only by the
order of the additions. ARTIS_BATCH_PACKETS forces small batches here."""
import os

import numpy as np
import pytest

import goldenlib
import statlib
from artis_b200.transport import Transport

FIXTURE = "classic_nogrey_ts12.npz"


def run(lib, fx, batch):
    old = os.environ.get("ARTIS_BATCH_PACKETS")
    if batch:
        os.environ["ARTIS_BATCH_PACKETS"] = str(batch)
    else:
        os.environ.pop("ARTIS_BATCH_PACKETS", None)
    try:
        T = Transport(lib)
        T.init(fx["static"], seed=4242)
        T.upload_cellstate(fx["cell"])
        T.zero_estimators()
        pk = fx["pre"].copy()
        T.update_packets(pk)
        est = T.download_estimators()
        T.finalize()
    finally:
        if old is None:
            os.environ.pop("ARTIS_BATCH_PACKETS", None)
        else:
            os.environ["ARTIS_BATCH_PACKETS"] = old
    return pk, est


def check(lib, exact_histories):
    fx = goldenlib.load_fixture(FIXTURE)
    p1, e1 = run(lib, fx, 0)
    p2, e2 = run(lib, fx, 1500)  # 4000 packets: batches of 1500, 1500, 1000
    if exact_histories:
        for f in p1.dtype.names:
            assert np.array_equal(p1[f], p2[f], equal_nan=True), f
    else:  # the device's estimator-dependent branches are none: histories are identical there too
        same = (p1["type"] == p2["type"]) & (p1["where"] == p2["where"]) & (p1["interactions"] == p2["interactions"])
        assert same.all()
        assert np.array_equal(p1["nu_cmf"], p2["nu_cmf"])
    good = statlib.good_cells(fx["static"], fx["cell"])
    n = fx["static"]["npts_model"]
    for k in ("J", "nuJ", "ffheatingestimator"):
        assert np.allclose(e1[k][:n][good], e2[k][:n][good], rtol=1e-10), k
    assert np.array_equal(e1["counters"][:34], e2["counters"][:34])
    assert e2["counters"][46] >= 3 * e1["counters"][46]  # ARTIS_CTR_PASSES: one (or more) per batch


def test_restatement_batched_update_packets(oracle_lib):
    check(oracle_lib, True)


@pytest.mark.gpu
def test_cuda_batched_update_packets(gpu_lib):
    check(gpu_lib, False)
