#!/usr/bin/env python
"""BASELINE.json config 4 from the config-1 workload: a 3-D model whose model cells ARE the propagation
cells, sampled from the 1-D profile.

The bench snapshot (oracle/_ref/runs/bench_base: the reference's classic-mode run on a 100^3 Cartesian grid) maps its
1e6 propagation cells onto 78 radial shells (grid.cc:988-1004 for a 1-D model.txt). promote() gives every non-empty
propagation cell its own model cell, with the state (T_e, T_R, W, n_e, composition, populations' inputs, cooling sums,
photoionisation renormalisation) of the shell its centre lies in: the rho(r), X(r) sampling of S4 applied to the
timestep's state instead of to model.txt (running the reference itself on a 5e5-cell model to timestep 25 is hours of
host update_grid). Every per-model-cell array of the ABI then has 5e5 rows -- the per-cell tables, the estimators (a
272 MB block to all-reduce) and the cell state -- which is what config 4 is about.

groups=G instead makes G model cells per shell (cell index modulo G): a model small enough for the CPU tests in
which packets still change model cell where the 1-D run does not.

rotate_copies() spreads the tiled copies of the snapshot's packets over the sphere (the model is spherically symmetric:
a rotated packet is statistically the same packet), so that the copies of one packet do not share a cell.

Test and bench infrastructure; no reference code."""
import numpy as np

PER_CELL = ("Te", "TR", "TJ", "W", "nne", "nnetot", "rho", "kappagrey", "ffegrp", "thick", "totalcooling")
PER_CELL_X = (("elem_abundance", "nelements"), ("groundlevelpop", "includedions"), ("partfunct", "includedions"),
              ("cooling_contrib_ion", "includedions"))


def _i(v):
    return int(np.asarray(v).reshape(-1)[0])


def promote(static, cell, groups=None):
    """-> (static3d, cell3d, shell_of_newmgi). Only for Cartesian grids with LTE populations and no per-cell arrays
    beyond the ones listed above (the classic option set)."""
    st, cs = dict(static), dict(cell)
    assert _i(st["opt_grid_type"]) == 0, "promote() needs the Cartesian propagation grid"
    for k in ("total_nlte_levels", "nradfieldbins", "nprevbfrate", "nt_nions"):
        assert _i(cs[k]) == 0, f"promote(): per-cell array {k} is not handled"
    npts = _i(st["npts_model"])
    old_mgi = np.asarray(st["cell_mgi"]).astype(np.int64)
    nonempty_old = np.asarray(st["nonemptymgi_of_mgi"])[:npts] >= 0
    alive = (old_mgi < npts) & nonempty_old[np.minimum(old_mgi, npts - 1)]
    cells = np.flatnonzero(alive)
    if groups is None:
        new_of_cell = np.arange(cells.size, dtype=np.int64)
        shell_of_new = old_mgi[cells]
    else:
        key = old_mgi[cells] * groups + (cells % groups)
        uniq, new_of_cell = np.unique(key, return_inverse=True)
        shell_of_new = uniq // groups
    n_new = int(shell_of_new.size)
    new_mgi = np.full(old_mgi.size, n_new, dtype=np.int32)   # the EMPTY cell
    new_mgi[cells] = new_of_cell
    st["cell_mgi"] = new_mgi
    st["npts_model"] = n_new
    st["nonempty_npts_model"] = n_new
    st["nonemptymgi_of_mgi"] = np.concatenate([np.arange(n_new, dtype=np.int32), np.array([-1], np.int32)])
    src = np.concatenate([shell_of_new, [npts]])             # rows of the old arrays, the EMPTY cell's row last
    cs["npts_model"] = n_new
    for k in PER_CELL:
        cs[k] = np.ascontiguousarray(np.asarray(cs[k])[src])
    for k, dim in PER_CELL_X:
        d = _i(st[dim])
        cs[k] = np.ascontiguousarray(np.asarray(cs[k]).reshape(npts + 1, d)[src].reshape(-1))
    ncorr = _i(cs["ncorrphotoionrenorm"])
    if ncorr > 0:
        per = ncorr // npts
        cs["corrphotoionrenorm"] = np.ascontiguousarray(np.asarray(cs["corrphotoionrenorm"]).reshape(npts, per)[shell_of_new].reshape(-1))
        cs["ncorrphotoionrenorm"] = n_new * per
    return st, cs, shell_of_new


def _rotations(n, rng):
    """n uniformly distributed rotation matrices (unit quaternions)"""
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1)[:, None]
    w, x, y, z = q.T
    R = np.empty((n, 3, 3))
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - z * w); R[:, 0, 2] = 2 * (x * z + y * w)
    R[:, 1, 0] = 2 * (x * y + z * w); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - x * w)
    R[:, 2, 0] = 2 * (x * z - y * w); R[:, 2, 1] = 2 * (y * z + x * w); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def rotate_copies(pk, npre, static, seed, chunk=1 << 20):
    """pk: packets tiled from a snapshot of npre packets (bench.py replicate). Every packet beyond the first npre is
    rotated about the origin by its own random rotation and put into the propagation cell it then lies in; packets that
    would land in an empty cell while they are not free-streaming (r-packet, gamma, escaped) keep their original place."""
    from artis_b200 import abi
    rng = np.random.default_rng(seed)
    nx, ny, nz = (_i(static[k]) for k in ("ncoordgrid_x", "ncoordgrid_y", "ncoordgrid_z"))
    cmax = float(np.asarray(static["coordmax_x"]).reshape(-1)[0])
    tmin = float(np.asarray(static["tmin"]).reshape(-1)[0])
    wid = 2 * cmax / nx
    mgi = np.asarray(static["cell_mgi"])
    empty = _i(static["npts_model"])
    moved = 0
    for a in range(npre, pk.size, chunk):
        v = pk[a:a + chunk]
        R = _rotations(v.size, rng)
        pos = np.einsum("nij,nj->ni", R, v["pos"])
        idx = np.floor((pos * (tmin / v["prop_time"])[:, None] + cmax) / wid).astype(np.int64)
        inside = ((idx >= 0) & (idx < np.array([nx, ny, nz]))).all(1)
        idx = np.clip(idx, 0, np.array([nx, ny, nz]) - 1)
        where = idx[:, 0] + nx * (idx[:, 1] + ny * idx[:, 2])
        free = (v["type"] == abi.TYPE_RPKT) | (v["type"] == abi.TYPE_GAMMA) | (v["type"] == abi.TYPE_ESCAPE)
        ok = inside & (free | (mgi[where] != empty))
        for f in ("pos", "dir", "em_pos", "absorptiondir", "pol_dir"):
            rot = np.einsum("nij,nj->ni", R, v[f])
            v[f][ok] = rot[ok]
        v["where"][ok] = where[ok].astype(np.int32)
        v["last_cross"][ok] = 107  # ARTIS_CROSS_NONE: the rotated packet does not sit on the boundary it crossed last
        moved += int(ok.sum())
    return moved
