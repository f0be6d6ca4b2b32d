"""Emission / absorption-resolved and angle-resolved spectra and light curves of the resident packets
(artis_gpu_bin_escaped_resolved) against the reference's own add_to_spec (do_emission_res, spectrum.cc:339-452),
add_to_spec_res / add_to_lc_res with current_abin = 0..99 (spectrum.cc:602-620, light_curve.cc:55-62) and
get_escapedirectionbin (vectors.h:158-192).

Golden vector: tests/golden/spectra_resolved_classic.npz by oracle/tools/ref_spectra_resolved.py (the post-transport packets
of timestep 34 of a classic-mode reference run, 687 escaped r-packets, and the non-zero bins of the five arrays the
reference's functions made of them). Which bins are non-zero is exact; a bin's value agrees to rounding (TOL)."""
import os

import numpy as np
import pytest

import goldenlib
from artis_b200 import abi
from artis_b200.transport import Transport

TOL = 1e-13


def load():
    z = np.load(os.path.join(goldenlib.GOLDEN, "spectra_resolved_classic.npz"))
    snap = {k: z[k] for k in z.files}
    static = goldenlib._section(snap, "static.", abi.STATIC_FIELDS)
    post = np.frombuffer(z["packets_post"].tobytes(), dtype=abi.PACKET_DTYPE).copy()
    want = {}
    for name in ("emission", "trueemission", "absorption", "lc", "flux"):
        a = np.zeros(int(z[f"res_{name}_size"][0]))
        a[z[f"res_{name}_idx"]] = z[f"res_{name}_val"]
        want[name] = a
    return static, post, z["params"], want


def check(lib):
    static, post, params, want = load()
    tmax, nprocs, nnubins, syn_dir = float(params[0]), int(params[1]), int(params[2]), params[3:6]
    T = Transport(lib)
    T.init(static)
    T.packets_upload(post)
    em, tem, ab, lc, fl = T.bin_escaped_resolved(tmax, syn_dir, nprocs, nnubins)
    got = {"emission": em, "trueemission": tem, "absorption": ab, "lc": lc, "flux": fl}
    for name, w in want.items():
        g = got[name].reshape(-1)
        assert g.size == w.size, name
        assert (w != 0).sum() > 500, name
        assert np.array_equal(g != 0, w != 0), f"{name}: different set of non-zero bins"
        nz = w != 0
        rel = np.abs(g[nz] - w[nz]) / np.abs(w[nz])
        assert rel.max() <= TOL, (name, rel.max())
    # consistency with the unresolved call: the process columns add up to the spectrum, the angle bins average to it
    lc0, flux0, _ = T.bin_escaped(tmax, nprocs, nnubins)
    assert np.allclose(em.sum(-1), flux0, rtol=1e-12, atol=0)
    assert np.allclose(tem.sum(-1), flux0, rtol=1e-12, atol=0)
    assert np.allclose(fl.sum(0) / 100.0, flux0, rtol=1e-12, atol=0)
    assert np.allclose(lc.sum(0) / 100.0, lc0[0], rtol=1e-12, atol=0)
    T.finalize()


def test_restatement_resolved_spectra_match_reference(oracle_lib):
    check(oracle_lib)


@pytest.mark.gpu
def test_cuda_resolved_spectra_match_reference(gpu_lib):
    check(gpu_lib)
