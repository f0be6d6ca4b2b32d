#!/usr/bin/env python
"""Synthetic atomic data in the reference's own file formats (TEST INFRASTRUCTURE).

The reference's test configs download `atomicdata_feconi.tar.xz` from the network (tests/setup_classicmode.sh:7-9);
it is not available offline, so the oracle runs use this seeded generator instead.

Writes into <outdir>:
  compositiondata.txt   input.cc:759-794   nelements, T_preset, homogeneous, then per element
                                           `Z nions lowermost_ionstage uppermost_ionstage nlevelsmax abundance mass_amu`
  adata.txt             input.cc:849,349   per ion `Z ionstage nlevels ionpot_eV` then
                                           `levelindex(1-based) energy_eV statweight ntransitions`
  transitiondata.txt    input.cc:893,429   per ion `Z ionstage ntransitions` then `lower upper A coll_str forbidden`
  phixsdata_v2.txt      input.cc:249-270   NPHIXSPOINTS, NPHIXSNUINCREMENT, then per table
                                           `Z upperionstage upperlevel lowerionstage lowerlevel threshold_eV`,
                                           optional target list, NPHIXSPOINTS cross sections in Mb
"""
import argparse
import os

import numpy as np

# first five ionisation potentials [eV]
IONPOT = {
    14: [8.15, 16.35, 33.49, 45.14, 166.77],
    16: [10.36, 23.34, 34.79, 47.22, 72.59],
    20: [6.11, 11.87, 50.91, 67.27, 84.50],
    26: [7.90, 16.19, 30.65, 54.8, 75.0],
    27: [7.88, 17.08, 33.50, 51.3, 79.5],
    28: [7.64, 18.17, 35.19, 54.9, 76.06],
}
MASS = {14: 28.09, 16: 32.06, 20: 40.08, 26: 55.85, 27: 58.93, 28: 58.69}


def generate(outdir, elements=(26, 27, 28), nions=5, nlevels=100, dense_levels=20, sparsity=0.1, seed=20240607,
             nphixspoints=100, nphixsnuincrement=0.1, multitarget_every=7, ion_ranges=None, abundance_only=(),
             single_level_top_ion=True):
    """ion_ranges: optional {Z: (lowermost_ionstage, uppermost_ionstage)} (default 1..nions for every element);
    abundance_only: [(Z, mass_amu)] elements without ions (compositiondata rows with nions = 0, as in
    tests/kilonova_inputfiles/compositiondata.txt); single_level_top_ion as in artisoptions_*.h."""
    rng = np.random.default_rng(seed)
    os.makedirs(outdir, exist_ok=True)
    rows = {}
    adata, trans, phixs = [], [], [f"{nphixspoints}", f"{nphixsnuincrement}"]
    nlines_total = 0
    for Z, mass in abundance_only:
        rows[Z] = f"{Z} 0 0 0 1 0.0 {mass}"
    nions_default = nions
    for Z in elements:
        lowest, highest = (ion_ranges or {}).get(Z, (1, nions_default))
        nions = highest - lowest + 1
        rows[Z] = f"{Z} {nions} {lowest} {highest} {nlevels} 0.0 {MASS[Z]}"
        ion_levels = []
        for ion in range(nions):
            ionpot = IONPOT[Z][lowest - 1 + ion]
            nlev = nlevels
            # level energies: ground at 0, a few low-lying metastable levels, the rest spread below 0.95*ionpot
            nlow = min(6, nlev - 1)
            e_low = np.sort(rng.uniform(0.02, 0.25, size=nlow)) * min(ionpot, 12.0) * 0.3
            e_high = np.sort(rng.uniform(0.12, 0.95, size=nlev - 1 - nlow)) * ionpot
            energies = np.concatenate([[0.0], np.sort(np.concatenate([e_low, e_high]))])
            g = rng.choice([1, 3, 5, 7, 9], size=nlev).astype(float)
            g[0] = rng.choice([5, 7, 9])
            ion_levels.append((energies, g))
            # transitions: all pairs among the lowest `dense_levels`, `sparsity` of the rest
            tl = []
            for up in range(1, nlev):
                for lo in range(up):
                    dense = up < dense_levels
                    if dense or rng.random() < sparsity:
                        forbidden = 1 if (rng.random() < 0.3 or up <= nlow) else 0
                        # A-values follow from an oscillator strength and the transition energy, so that close-lying
                        # levels get small A (A ~ nu^2 f): A_ul = 6.6702e15 g_l/g_u f / lambda[Angstrom]^2.
                        # Forbidden lines: f ~ 1e-9..1e-6.
                        lam = 12398.42 / max(energies[up] - energies[lo], 1e-6)
                        f = 10 ** (rng.uniform(-9.0, -6.0) if forbidden else rng.uniform(-3.5, -0.3))
                        A = 6.6702e15 * g[lo] / g[up] * f / lam ** 2
                        tl.append((lo + 1, up + 1, A, -1.0, forbidden))
            ntr = np.zeros(nlev, dtype=int)
            for lo, up, *_ in tl:
                ntr[up - 1] += 1
            adata.append(f"{Z} {lowest + ion} {nlev} {ionpot:.6f}")
            for i in range(nlev):
                adata.append(f"{i + 1} {energies[i]:.8f} {g[i]:.1f} {ntr[i]}")
            adata.append("")
            # transitiondata sorted by lower then upper (input.cc:429-470 expects grouped by lower level)
            tl.sort(key=lambda x: (x[0], x[1]))
            trans.append(f"{Z} {lowest + ion} {len(tl)}")
            for lo, up, A, cs, fb in tl:
                trans.append(f"{lo} {up} {A:.6e} {cs:.1f} {fb}")
            trans.append("")
            if ion < nions - 1:
                nlines_total += len(tl)
        # photoionisation tables for every level of every ion below the top one
        for ion in range(nions - 1):
            energies, g = ion_levels[ion]
            up_energies, _ = ion_levels[ion + 1]
            ionpot = IONPOT[Z][lowest - 1 + ion]
            top_is_next = single_level_top_ion and (ion + 1 == nions - 1)  # then only level 0 exists there
            for lev in range(len(energies)):
                sigma0 = 10 ** rng.uniform(-1.0, 1.0)
                x = 1.0 + nphixsnuincrement * np.arange(nphixspoints)
                sig = sigma0 * x ** -3.0
                multi = (not top_is_next) and (lev % multitarget_every == 0)
                if multi:
                    thr = ionpot + up_energies[0] - energies[lev]
                    phixs.append(f"{Z} {lowest + ion + 1} -1 {lowest + ion} {lev + 1} {thr:.8f}")
                    phixs.append("2")
                    phixs.append("1 0.7")
                    phixs.append("2 0.3")
                else:
                    thr = ionpot - energies[lev]
                    phixs.append(f"{Z} {lowest + ion + 1} 1 {lowest + ion} {lev + 1} {thr:.8f}")
                phixs.extend(f"{s:.6e}" for s in sig)
    comp = [f"{len(rows)}", "0", "0"] + [rows[Z] for Z in sorted(rows)]
    open(os.path.join(outdir, "compositiondata.txt"), "w").write("\n".join(comp) + "\n")
    open(os.path.join(outdir, "adata.txt"), "w").write("\n".join(adata) + "\n")
    open(os.path.join(outdir, "transitiondata.txt"), "w").write("\n".join(trans) + "\n")
    open(os.path.join(outdir, "phixsdata_v2.txt"), "w").write("\n".join(phixs) + "\n")
    return nlines_total


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("outdir")
    ap.add_argument("--elements", default="26,27,28")
    ap.add_argument("--nions", type=int, default=5)
    ap.add_argument("--nlevels", type=int, default=100)
    ap.add_argument("--dense-levels", type=int, default=20)
    ap.add_argument("--sparsity", type=float, default=0.1)
    ap.add_argument("--seed", type=int, default=20240607)
    a = ap.parse_args()
    n = generate(a.outdir, tuple(int(z) for z in a.elements.split(",")), a.nions, a.nlevels, a.dense_levels,
                 a.sparsity, a.seed)
    print("lines (below top ions):", n)
