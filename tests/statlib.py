"""Summary statistics of one update_packets call, shared by the golden-fixture generator
(oracle/tools/ref_ensemble.py, which runs the reference) and the parity tests (which run the CUDA path).

A Monte Carlo transport step cannot be compared packet by packet (the reference consumes one GSL ran3
stream in packet-visit order, the device draws from per-packet Philox streams), so parity of the full step is
statistical: an ensemble of K reference runs of the SAME timestep from the SAME start state with K seeds is compared
with an ensemble of device runs. For every scalar below the test is a two-sample z-test with both ensemble variances.
"""
import numpy as np

TYPES = (10, 11, 12, 13, 20, 21, 32, 100, 120)
NSPEC = 12


def good_cells(static, cell):
    """non-empty, non-grey model cells: the only ones whose estimators mean something in the reference (in grey cells
    the reference adds whatever opacity scratch its host thread last computed for some other cell)."""
    n = int(static["npts_model"])
    return (np.asarray(static["nonemptymgi_of_mgi"][:n]) >= 0) & (np.asarray(cell["thick"][:n]) != 1)


def nonempty_cells(static):
    n = int(static["npts_model"])
    return np.asarray(static["nonemptymgi_of_mgi"][:n]) >= 0


def summary(static, cell, est, post, pre):
    """dict name -> float for one transported timestep."""
    n = int(static["npts_model"])
    ng = int(static["nelements"]) * int(static["max_nions"])
    good = good_cells(static, cell)
    ne = nonempty_cells(static)
    d = {}
    # radiation-field estimators: totals over non-grey cells and three radial groups of them
    idx = np.nonzero(good)[0]
    groups = np.array_split(idx, 3) if idx.size >= 3 else [idx]
    for key, short in (("J", "J"), ("nuJ", "nuJ"), ("ffheatingestimator", "ff"), ("colheatingestimator", "col")):
        a = np.asarray(est[key][:n])
        d[short + "_tot"] = a[good].sum()
        for g, ii in enumerate(groups):
            d[f"{short}_g{g}"] = a[ii].sum()
    a = np.asarray(est["rpkt_emiss"][:n])
    d["rpkt_emiss_tot"] = a[ne].sum()
    for key, short in (("gammaestimator", "gamma"), ("bfheatingestimator", "bfheat")):
        a = np.asarray(est[key][: n * ng]).reshape(n, ng)
        d[short + "_tot"] = a[good].sum()
    # multi-bin radiation field and detailed bf-rate estimators (nebular NLTE option set; radfield.cc:764-874): totals and
    # four contiguous groups of bins / continua, summed over the non-empty cells
    nb = int(est.get("nradfieldbins", 0) or 0)
    if nb > 0:
        for key, short in (("radfieldbin_J_raw", "binJ"), ("radfieldbin_nuJ_raw", "binnuJ"),
                           ("radfieldbin_contribcount", "bincount")):
            a = np.asarray(est[key][:nb], dtype=float).reshape(-1, int(static["opt_radfieldbincount"])).sum(0)
            d[short + "_tot"] = a.sum()
            for g, part in enumerate(np.array_split(a, 4)):
                d[f"{short}_q{g}"] = part.sum()
    nr = int(est.get("nbfrate", 0) or 0)
    if nr > 0:
        a = np.asarray(est["bfrate_raw"][:nr]).reshape(-1, int(static["nbfcontinua"])).sum(0)
        d["bfrate_tot"] = a.sum()
        for g, part in enumerate(np.array_split(a, 4)):
            d[f"bfrate_q{g}"] = part.sum()
    ts = np.asarray(est["ts_scalars"])
    for i, nm in enumerate(("gamma_dep", "positron_dep", "electron_dep", "electron_emission", "alpha_dep",
                            "alpha_emission", "gamma_emission", "cmf_lum", "nt_energy_deposited")):
        d["ts_" + nm] = ts[i]
    # virtual packets (VPKT_ON; vpkt.cc:388-406,581-627): counts, Stokes I per (observer, spectrum), I / Q / U of the
    # full spectrum in three frequency bands per observer, and the velocity maps' totals
    nvs = int(est.get("nvstokes", 0) or 0)
    if nvs > 0:
        for i, nm in enumerate(("nvpkt", "nvpkt_esc1", "nvpkt_esc2", "nvpkt_esc3")):
            d["ts_" + nm] = ts[9 + i]
        nobs, nspec = int(static["vpkt_nobs"]), int(static["vpkt_nspectra"])
        v = np.asarray(est["vstokes"][:nvs]).reshape(3, int(static["vpkt_vmtbins"]), nobs * nspec,
                                                       int(static["vpkt_vmnubins"]))
        for ic in range(nobs * nspec):
            d[f"vI_o{ic // nspec}_s{ic % nspec}"] = v[0, :, ic, :].sum()
        for ob in range(nobs):
            for s3, nm in enumerate("IQU"):
                for g, part in enumerate(np.array_split(v[s3, :, ob * nspec, :].sum(0), 3)):
                    d[f"v{nm}_o{ob}_band{g}"] = part.sum()
        nvg = int(est.get("nvgrid", 0) or 0)
        if nvg > 0:
            g4 = np.asarray(est["vgrid"][:nvg]).reshape(3, -1, int(static["vpkt_nrange_grid"]), nobs)
            for s3, nm in enumerate("IQU"):
                for r in range(int(static["vpkt_nrange_grid"])):
                    d[f"vgrid{nm}_r{r}"] = g4[s3, :, r, :].sum()
    ctr = np.asarray(est["counters"])
    for i in range(31):  # stats::eventcounters except UPDATECELL / COOLINGRATECALC / NESC (scheduler-specific)
        d[f"ctr{i:02d}"] = float(ctr[i])
    d["acounter_sum"] = float(np.asarray(est["acounter"]).sum())
    d["ecounter_sum"] = float(np.asarray(est["ecounter"]).sum())
    # packets
    d["interactions"] = float(post["interactions"].sum())
    for t in TYPES:
        d[f"ntype{t}"] = float((post["type"] == t).sum())
    newesc = (post["type"] == 32) & _was_active(post, pre)
    d["n_newly_escaped"] = float(newesc.sum())
    d["e_rf_newly_escaped_r"] = post["e_rf"][newesc & (post["escape_type"] == 11)].sum()
    d["e_rf_newly_escaped_gamma"] = post["e_rf"][newesc & (post["escape_type"] == 10)].sum()
    r = post["type"] == 11
    d["rpkt_e_cmf"] = post["e_cmf"][r].sum()
    d["rpkt_mean_lognu"] = np.log(post["nu_cmf"][r]).mean() if r.any() else 0.0
    d["rpkt_mean_radius"] = np.sqrt((post["pos"][r] ** 2).sum(1)).mean() if r.any() else 0.0
    d["rpkt_mean_mu"] = ((post["pos"][r] * post["dir"][r]).sum(1) / np.sqrt((post["pos"][r] ** 2).sum(1))).mean() if r.any() else 0.0
    d["rpkt_mean_nscat"] = post["nscatterings"][r].mean() if r.any() else 0.0
    # coarse emergent spectrum of the packets that escaped in this step (spectrum.cc:339 bins nu_rf logarithmically)
    er = newesc & (post["escape_type"] == 11)
    edges = np.logspace(np.log10(float(static["nu_min_r"])), np.log10(float(static["nu_max_r"])), NSPEC + 1)
    h, _ = np.histogram(post["nu_rf"][er], bins=edges, weights=post["e_rf"][er])
    for i in range(NSPEC):
        d[f"spec{i:02d}"] = h[i]
    return d


def summary_pol(post, pre):
    """Polarisation state of the REAL packets after the step (POL_ON builds; polarization.cc:6-140 escat_rpkt, the frame
    transforms of vpkt.cc:898-1069): second moments and mean magnitudes of the Stokes Q / U that the electron scatterings
    of this timestep leave in the resident r-packets and in those that escaped, their energy-weighted sums (zero within
    noise for a spherical model, so the z-test checks the sign conventions too), the polarisation reference axis, and the
    scattering counts that go with them. Kept apart from summary() so that the fixtures written before it keep their
    column lists; used by the bench-workload ensembles (oracle/tools/ref_bench_rows.py)."""
    d = {}
    newesc = (post["type"] == 32) & _was_active(post, pre) & (post["escape_type"] == 11)
    for tag, m in (("res", post["type"] == 11), ("esc", newesc)):
        q, u, i0 = post["stokes"][m, 1], post["stokes"][m, 2], post["stokes"][m, 0]
        e = post["e_rf"][m]
        n = max(int(m.sum()), 1)
        d[f"pol_{tag}_I_mean"] = i0.sum() / n
        d[f"pol_{tag}_Q2_mean"] = (q * q).sum() / n
        d[f"pol_{tag}_U2_mean"] = (u * u).sum() / n
        d[f"pol_{tag}_absQ_mean"] = np.abs(q).sum() / n
        d[f"pol_{tag}_absU_mean"] = np.abs(u).sum() / n
        d[f"pol_{tag}_frac_polarised"] = float(((q != 0) | (u != 0)).sum()) / n
        d[f"pol_{tag}_eQ_sum"] = (e * q).sum()
        d[f"pol_{tag}_eU_sum"] = (e * u).sum()
        d[f"pol_{tag}_poldir_z_abs_mean"] = np.abs(post["pol_dir"][m, 2]).sum() / n
        d[f"pol_{tag}_nscat_mean"] = post["nscatterings"][m].sum() / n
    return d


def _was_active(post, pre):
    """mask over `post` of packets that were not yet escaped in `pre` (matched by packet.number: the reference
    re-orders the array, update_packets.cc:256)."""
    esc_before = np.zeros(int(max(pre["number"].max(), post["number"].max())) + 1, dtype=bool)
    esc_before[pre["number"][pre["type"] == 32]] = True
    return ~esc_before[post["number"]]


def ztest(ref_rows, dev_rows, names):
    """two-sample z for every quantity; returns list of (name, ref_mean, dev_mean, z)."""
    out = []
    ra, da = np.asarray(ref_rows, float), np.asarray(dev_rows, float)
    for j, nm in enumerate(names):
        r, d = ra[:, j], da[:, j]
        se = np.sqrt(r.var(ddof=1) / r.size + d.var(ddof=1) / d.size)
        # rare-event quantities (e.g. a spectrum bin hit by a packet in a few runs out of hundreds): a small ensemble
        # that happens to hold no event has zero sample variance and the unpooled error is meaningless; under the
        # null hypothesis both ensembles share one distribution, so use its pooled variance
        if min((r != 0).mean(), (d != 0).mean()) < 0.2:
            both = np.concatenate([r, d])
            se = np.sqrt(both.var(ddof=1) * (1.0 / r.size + 1.0 / d.size))
        diff = d.mean() - r.mean()
        scale = max(abs(r.mean()), abs(d.mean()))
        if scale == 0 or abs(diff) <= 1e-9 * scale:  # identical up to summation order (deterministic quantities)
            z = 0.0
        elif se == 0:
            z = np.inf
        else:
            z = diff / se
        out.append((nm, r.mean(), d.mean(), z))
    return out
