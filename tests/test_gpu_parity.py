"""GPU tests (-m gpu): the parity tests proper. Everything here goes through the C ABI of
artis_b200/libartis_gpu.so (CUDA, sm_100a); the oracle's restatement and the golden vectors are only the checkers."""
import numpy as np
import pytest

import goldenlib
import probelib
import statlib
from artis_b200 import abi
from artis_b200.transport import Transport

pytestmark = pytest.mark.gpu


def _ready(lib, fx, seed=1281360349):
    T = Transport(lib)
    T.init(fx["static"], seed=seed)
    T.upload_cellstate(fx["cell"])
    T.zero_estimators()
    return T


def test_cuda_probes_match_reference(gpu_lib, fixture_name):
    """A4-A8, A11, A12, A14, A18 with injected random numbers against the reference's own outputs"""
    fx = goldenlib.load_fixture(fixture_name)
    T = _ready(gpu_lib, fx)
    out = probelib.check_probes(T, fx, "gpu")
    print(fixture_name, {k: float(v) for k, v in out.items()})
    T.finalize()


def test_cuda_ensemble_matches_reference_ensemble(gpu_lib, fixture_name):
    """spectrum bins, J / nuJ / heating estimators, deposition scalars, event counters, packet census: two-sample
    z-test of 24 reference runs against 24 CUDA runs of the same timestep (tests/statlib.py)"""
    fx = goldenlib.load_fixture(fixture_name)
    T = Transport(gpu_lib)
    rows = goldenlib.run_ensemble(T, fx, 24)
    res = statlib.ztest(fx["ref_rows"], rows, fx["names"])
    zs = np.array([z for _, r, d, z in res if (r != 0 or d != 0)])
    print(fixture_name, "max|z|", np.abs(zs).max(), "rms z", np.sqrt((zs ** 2).mean()))
    assert np.all(np.isfinite(zs))
    assert np.abs(zs).max() < 4.5, [x for x in res if abs(x[3]) >= 4.5]
    assert np.sqrt((zs ** 2).mean()) < 1.5
    T.finalize()


def test_cuda_matches_restatement_same_streams(gpu_lib, oracle_lib, fixture_name):
    """same seeded inputs, same Philox streams: the CUDA path and the CPU restatement follow the same histories except
    where a comparison is decided in the last ulps of a libm call"""
    fx = goldenlib.load_fixture(fixture_name)
    outs = []
    for lib in (gpu_lib, oracle_lib):
        T = _ready(lib, fx, seed=777)
        pk = fx["pre"].copy()
        T.update_packets(pk)
        outs.append((pk, T.download_estimators()))
        T.finalize()
    (a, ea), (b, eb) = outs
    same = (a["type"] == b["type"]) & (a["where"] == b["where"]) & (a["interactions"] == b["interactions"])
    print(fixture_name, "identical histories:", same.mean())
    assert same.mean() > 0.97
    close = same & (a["type"] == abi.TYPE_RPKT)
    assert np.allclose(a["nu_cmf"][close], b["nu_cmf"][close], rtol=1e-6)
    good = statlib.good_cells(fx["static"], fx["cell"])
    n = fx["static"]["npts_model"]
    assert np.isclose(ea["J"][:n][good].sum(), eb["J"][:n][good].sum(), rtol=2e-2)


def test_cuda_postconditions_and_conservation(gpu_lib, fixture_name):
    fx = goldenlib.load_fixture(fixture_name)
    T = _ready(gpu_lib, fx, seed=99)
    pk = fx["pre"].copy()
    T.update_packets(pk)
    st = fx["static"]
    t2 = st["ts_start"][fx["nts"]] + st["ts_width"][fx["nts"]]
    assert np.array_equal(pk["number"], fx["pre"]["number"])
    assert ((pk["type"] == abi.TYPE_ESCAPE) | (pk["prop_time"] == t2)).all()
    est = T.download_estimators()
    newly = (pk["type"] == abi.TYPE_ESCAPE) & (fx["pre"]["type"] != abi.TYPE_ESCAPE)
    assert est["counters"][33] == newly.sum()
    r_esc = newly & (pk["escape_type"] == abi.TYPE_RPKT)
    assert np.isclose(est["ts_scalars"][7], pk["e_cmf"][r_esc].sum(), rtol=1e-12)
    c = est["counters"]
    assert c[0] + c[1] + c[2] + c[3] + c[4] + c[5] + c[6] == c[7] + c[8] + c[9] + c[10] + (pk["type"] == abi.TYPE_MA).sum()
    assert est["counters"][45] == pk["interactions"].sum()  # device tally == sum of pkt.interactions (stats.cc:172)
    r = pk["type"] == abi.TYPE_RPKT
    assert np.allclose((pk["dir"][r] ** 2).sum(1), 1.0, atol=1e-10)
    assert T.launch_count() > 0
    T.finalize()


def test_cuda_edge_cases_and_scale(gpu_lib):
    fx = goldenlib.load_fixture("classic_nogrey_ts12.npz")
    T = _ready(gpu_lib, fx)
    T.update_packets(np.zeros(0, dtype=abi.PACKET_DTYPE))  # empty input
    esc = fx["pre"][fx["pre"]["type"] == abi.TYPE_ESCAPE][:50].copy()
    before = esc.copy()
    T.update_packets(esc)  # nothing to do
    assert np.array_equal(esc["pos"], before["pos"]) and np.array_equal(esc["type"], before["type"])
    # ragged size (not a multiple of the block) and a large replicated population: linearity of the estimators in
    # the number of statistically identical packets (size-independent property at full scale)
    reps = 25
    big = np.tile(fx["pre"], reps)
    big["number"] = np.arange(big.size, dtype=np.int32)
    big = big[: big.size - 37].copy()
    T.zero_estimators()
    T.update_packets(big)
    e_big = T.download_estimators()
    T.zero_estimators()
    small = fx["pre"].copy()
    T.update_packets(small)
    e_small = T.download_estimators()
    good = statlib.good_cells(fx["static"], fx["cell"])
    n = fx["static"]["npts_model"]
    ratio = e_big["J"][:n][good].sum() / e_small["J"][:n][good].sum()
    assert abs(ratio / (big.size / small.size) - 1) < 0.01
    st = fx["static"]
    t2 = st["ts_start"][fx["nts"]] + st["ts_width"][fx["nts"]]
    assert ((big["type"] == abi.TYPE_ESCAPE) | (big["prop_time"] == t2)).all()
    T.finalize()



def test_scheduler_variants_give_identical_packets(gpu_lib, fixture_name):
    """scheduling must not change physics: one persistent kernel with every role (a macro-atom iteration in one piece) and
    the two-kernel scheduler (jumps drawn by the jump kernel on 32 bytes of the packet, every other action handed to the
    event role of the other kernel through pend[]) produce the same packets bit for bit -- every packet owns its Philox
    stream and its draw counter travels with it"""
    fx = goldenlib.load_fixture(fixture_name)
    outs = []
    for mode in (-3, -4):
        T = _ready(gpu_lib, fx, seed=99)
        T.set_tuning(block_threads=mode)  # one kernel (-3), two kernels (-4)
        geo = T.scheduler_geometry()
        assert geo[0] == (1 if mode == -3 else 2), geo
        pk = fx["pre"].copy()
        T.update_packets(pk)
        outs.append(pk)
        T.set_tuning(block_threads=-4)  # back to the default for whatever runs next in this process
        T.finalize()
    for name in outs[0].dtype.names:
        assert np.array_equal(outs[0][name], outs[1][name], equal_nan=True), ("two-kernel", name)


def test_classic_build_and_runtime_option_build_agree(gpu_lib, monkeypatch):
    """libartis_gpu_classic.so (the switches of artisoptions_classic.h compiled in; what Transport picks for that option
    set) against libartis_gpu.so (every switch a run-time field) on a classic fixture: the probes hold for both, and the
    same seeded packets live the same lives (two builds of the sources: a few packets in a hundred see a last-bit
    difference of a contracted multiply-add and part ways somewhere in a thousand events)"""
    from artis_b200 import transport
    fx = goldenlib.load_fixture("classic_nogrey_ts12.npz")
    outs = []
    for generic in (False, True):
        if generic:
            monkeypatch.setenv("ARTIS_GPU_GENERIC", "1")
        T = _ready(gpu_lib, fx, seed=4242)
        want = transport.DEFAULT_LIB if generic else transport.CLASSIC_LIB
        assert T.libpath == want, (T.libpath, want)
        probelib.check_probes(T, fx, "gpu")
        pk = fx["pre"].copy()
        T.update_packets(pk)
        outs.append((pk, T.download_estimators()))
        T.finalize()
    (a, ea), (b, eb) = outs
    same = (a["type"] == b["type"]) & (a["where"] == b["where"]) & (a["interactions"] == b["interactions"])
    print("same history", same.mean())
    assert same.mean() > 0.97
    assert np.allclose(a["nu_cmf"][same], b["nu_cmf"][same], rtol=1e-6)
    # and the classic build refuses another option set instead of mis-running it
    neb = goldenlib.load_fixture("nebular_nlte_ts8.npz")
    Tc = Transport(transport.CLASSIC_LIB)
    st = dict(neb["static"])
    st["abi_version"] = abi.ABI_VERSION
    s, keep = abi.build_struct(abi.ArtisStaticTables, abi.STATIC_FIELDS, st)
    import ctypes as C
    assert Tc.lib.artis_gpu_init(C.byref(s), 0, 0, 1, C.c_ulonglong(1)) == abi.ERR_UNSUPPORTED
