"""Light curves and spectrum of the resident packets (artis_gpu_bin_escaped) against the reference's own
add_to_lc_res / add_to_spec_res output.

Golden vector: tests/golden/spectra_classic_vpkt_ts30.npz, written by oracle/tools/ref_spectra.py -- the post-transport
packets of one reference timestep and the arrays the reference's functions made of them. The sums are FP64 atomics on
the device, so a bin may differ from the reference's serial sum by rounding: TOL below. Which bins are non-zero is exact.
"""
import os

import numpy as np
import pytest

import goldenlib
from artis_b200 import abi
from artis_b200.transport import Transport

TOL = 1e-13  # relative, per bin: at most a few hundred additions of positive terms per bin in these cases


def load_spectra_fixture():
    z = np.load(os.path.join(goldenlib.GOLDEN, "spectra_classic_vpkt_ts30.npz"))
    fx = goldenlib.load_fixture("classic_vpkt_ts30.npz")  # static tables of the same run directory
    static = dict(fx["static"])
    static["opt_vpkt_on"] = 0  # binning needs no virtual packets: stay on the default library
    post = np.frombuffer(z["packets_post"].tobytes(), dtype=abi.PACKET_DTYPE).copy()
    ntstep = int(static["ntstep"])
    tmax, nprocs, nnubins = float(z["params"][0]), int(z["params"][1]), int(z["params"][2])
    return {"static": static, "post": post, "tmax": tmax, "nprocs": nprocs, "nnubins": nnubins,
            "lc": z["lc"].reshape(4, ntstep), "flux": z["flux"].reshape(ntstep, nnubins),
            "stokes": z["stokes"].reshape(3, ntstep, nnubins)}


def assert_bins(got, want, what):
    assert np.array_equal(got != 0, want != 0), f"{what}: different set of non-zero bins"
    nz = want != 0
    if nz.any():
        rel = np.abs(got[nz] - want[nz]) / np.abs(want[nz])
        assert rel.max() <= TOL, (what, rel.max())


def check_against_golden(lib):
    g = load_spectra_fixture()
    T = Transport(lib)
    T.init(g["static"])
    T.packets_upload(g["post"])
    lc, flux, stokes = T.bin_escaped(g["tmax"], g["nprocs"], g["nnubins"], stokes=True)
    assert (g["flux"] != 0).sum() > 300 and (g["lc"][2] != 0).sum() > 0  # the fixture exercises r-packets and gammas
    assert_bins(lc, g["lc"], "light curves")
    assert_bins(flux, g["flux"], "spectrum")
    assert_bins(stokes, g["stokes"], "stokes spectra")
    # without the Stokes planes the other outputs are unchanged; a second call starts from zero again
    lc2, flux2, none = T.bin_escaped(g["tmax"], g["nprocs"], g["nnubins"])
    assert none is None
    assert_bins(lc2, g["lc"], "light curves (2nd call)")
    assert_bins(flux2, g["flux"], "spectrum (2nd call)")
    # nprocs is a plain normalisation (spectrum.cc:360)
    lc3, flux3, _ = T.bin_escaped(g["tmax"], 4, g["nnubins"])
    assert np.allclose(lc3 * 4, g["lc"] * g["nprocs"], rtol=1e-12, atol=0)
    assert np.allclose(flux3 * 4, g["flux"] * g["nprocs"], rtol=1e-12, atol=0)
    # no escaped packet, and no packet at all
    none_escaped = g["post"][g["post"]["type"] != abi.TYPE_ESCAPE][:64].copy()
    T.packets_upload(none_escaped)
    lc4, flux4, _ = T.bin_escaped(g["tmax"], 1, g["nnubins"])
    assert not lc4.any() and not flux4.any()
    T.packets_upload(none_escaped[:0])
    lc5, flux5, _ = T.bin_escaped(g["tmax"], 1, g["nnubins"])
    assert not lc5.any() and not flux5.any()
    with pytest.raises(Exception):
        T.bin_escaped(g["tmax"], 0, g["nnubins"])  # nprocs = 0 would divide by zero
    T.finalize()
    return g


def synthetic_escaped(g, n, seed):
    """n packets, three quarters of them escaped, with arrival times and frequencies spread over (and beyond) the time
    and frequency grids so that every range check of add_to_spec / add_to_lc_res is hit."""
    rng = np.random.default_rng(seed)
    st = g["static"]
    rows = g["post"].view(np.uint8).reshape(-1, 304)[np.arange(n) % g["post"].size]  # (np.resize would repack the dtype)
    p = np.ascontiguousarray(rows).view(abi.PACKET_DTYPE).reshape(-1)
    p["type"] = np.where(rng.random(n) < 0.75, abi.TYPE_ESCAPE, 11)
    p["escape_type"] = np.where(rng.random(n) < 0.8, 11, 10)
    tmin, tmax = float(st["tmin"]), g["tmax"]
    p["escape_time"] = np.exp(rng.uniform(np.log(tmin * 0.8), np.log(tmax * 1.2), n)).astype(np.int32)
    mu = rng.uniform(-1, 1, n)
    phi = rng.uniform(0, 2 * np.pi, n)
    d = np.stack([np.sqrt(1 - mu * mu) * np.cos(phi), np.sqrt(1 - mu * mu) * np.sin(phi), mu], axis=1)
    p["dir"] = d
    p["pos"] = rng.uniform(-1, 1, (n, 3)) * float(st["rmax"]) * (p["escape_time"] / tmin)[:, None]
    p["nu_rf"] = np.exp(rng.uniform(np.log(float(st["nu_min_r"]) * 0.5), np.log(float(st["nu_max_r"]) * 2), n))
    p["e_rf"] = rng.uniform(0.5, 1.5, n) * 1e40
    p["e_cmf"] = p["e_rf"] * rng.uniform(0.9, 1.1, n)
    p["stokes"] = rng.uniform(-1, 1, (n, 3))
    return p


def check_escaped_compaction(lib, n):
    """array-order compaction of the escaped packets; only_new delivers every escaped packet exactly once"""
    g = load_spectra_fixture()
    pk = g["post"].copy() if n is None else synthetic_escaped(g, n, 11)
    pk["number"] = np.arange(pk.size)
    T = Transport(lib)
    T.init(g["static"])
    T.packets_upload(pk)
    # whole 304-byte records, padding included (a boolean index on the structured array would drop the padding bytes)
    want = pk.view(np.uint8).reshape(-1, 304)[pk["type"] == abi.TYPE_ESCAPE]
    raw = lambda a: a.view(np.uint8).reshape(-1, 304)  # noqa: E731
    got = T.download_escaped()
    assert got.size == want.shape[0] and np.array_equal(raw(got), want)
    assert np.array_equal(got["number"], np.flatnonzero(pk["type"] == abi.TYPE_ESCAPE))  # array order
    first = T.download_escaped(only_new=True)
    assert np.array_equal(raw(first), want)
    assert T.download_escaped(only_new=True).size == 0            # all delivered
    assert np.array_equal(raw(T.download_escaped()), want)        # the plain call is not affected by the marks
    with pytest.raises(Exception):
        T.download_escaped(capacity=want.shape[0] - 1)            # too small a buffer is an error, not a truncation
    # a new ensemble forgets the marks; an ensemble without escapes gives nothing
    T.packets_upload(pk)
    assert T.download_escaped(only_new=True).size == want.shape[0]
    T.packets_upload(pk[pk["type"] != abi.TYPE_ESCAPE].copy())
    assert T.download_escaped().size == 0
    T.finalize()


def test_restatement_escaped_compaction(oracle_lib):
    check_escaped_compaction(oracle_lib, None)
    check_escaped_compaction(oracle_lib, 5000)


@pytest.mark.gpu
def test_cuda_escaped_compaction(gpu_lib):
    check_escaped_compaction(gpu_lib, None)
    for n in (1, 255, 256, 257, 1_000_003):  # chunk edges and more chunks than one scan sweep
        check_escaped_compaction(gpu_lib, n)


def test_restatement_spectra_match_reference(oracle_lib):
    check_against_golden(oracle_lib)


def energy_checksum(g, pk, lc, nprocs):
    """every escaped r-packet that arrives inside (tmin, tmax) lands in exactly one light-curve bin"""
    esc = (pk["type"] == abi.TYPE_ESCAPE) & (pk["escape_type"] == 11)
    t_arr = pk["escape_time"] - (pk["pos"] * pk["dir"]).sum(1) / 2.99792458e10
    inr = esc & (t_arr > float(g["static"]["tmin"])) & (t_arr < g["tmax"])
    width = np.asarray(g["static"]["ts_width"])
    assert np.isclose((lc[0] * width).sum() * nprocs, pk["e_rf"][inr].sum(), rtol=1e-10)


def test_restatement_spectra_energy_checksum(oracle_lib):
    g = load_spectra_fixture()
    pk = synthetic_escaped(g, 200_000, 5)  # arrival times before tmin and after tmax included
    T = Transport(oracle_lib)
    T.init(g["static"])
    T.packets_upload(pk)
    lc, flux, _ = T.bin_escaped(g["tmax"], 3, g["nnubins"])
    energy_checksum(g, pk, lc, 3)
    assert (flux != 0).mean() > 0.5
    T.finalize()


@pytest.mark.gpu
def test_cuda_spectra_match_reference(gpu_lib):
    check_against_golden(gpu_lib)


@pytest.mark.gpu
def test_cuda_spectra_match_restatement_at_scale(gpu_lib, oracle_lib):
    g = load_spectra_fixture()
    pk = synthetic_escaped(g, 2_000_000, 7)
    out = []
    for lib in (gpu_lib, oracle_lib):
        T = Transport(lib)
        T.init(g["static"])
        T.packets_upload(pk)
        out.append(T.bin_escaped(g["tmax"], 3, g["nnubins"], stokes=True))
        T.finalize()
    (lc_g, fl_g, st_g), (lc_o, fl_o, st_o) = out
    assert (fl_o != 0).mean() > 0.5  # most bins hit
    assert np.array_equal(fl_g != 0, fl_o != 0)
    # light-curve bins hold ~1e4 terms each, summed in a different order on the device
    assert np.allclose(lc_g, lc_o, rtol=1e-11, atol=0)
    assert np.allclose(fl_g, fl_o, rtol=1e-12, atol=0)
    # Stokes terms have both signs: compare against the size of the terms, not of the (cancelling) sum
    scale = np.abs(fl_o)[None]
    assert (np.abs(st_g - st_o) <= 1e-12 * scale).all()
    energy_checksum(g, pk, lc_g, 3)
