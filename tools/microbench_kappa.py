#!/usr/bin/env python
"""Micro-benchmark of the continuum-opacity evaluation in isolation: calculate_kappa_rpkt_cont (one cache miss per thread)
for the r-packets of the bench workload that sit in non-grey cells, through artis_gpu_probe_kappa_cont at full occupancy.
Prints evaluations/s and bf terms/s; compare with the in-kernel rate (bench.py counters) to see what the scheduler costs.

usage (GPU box): [ARTIS_GPU_LIB=...] python tools/microbench_kappa.py [n_evals]
"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from artis_b200 import abi
from artis_b200.transport import Transport

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
st, cs, pre, name = bench.load_workload()
mgi = np.asarray(st["cell_mgi"])[pre["where"]]
thick = np.asarray(cs["thick"])
thin = (pre["type"] == abi.TYPE_RPKT) & (mgi < int(st["npts_model"])) & (thick[np.minimum(mgi, thick.size - 1)] != 1)
src = pre[thin]
pk = np.tile(src, -(-n // src.size))[:n].copy()
rng = np.random.default_rng(1)
rng.shuffle(pk)  # like the rings: neighbouring lanes hold unrelated packets
T = Transport()
T.init(st)
T.upload_cellstate(cs)
for rep in range(3):
    k = T.probe_kappa_cont(pk)
    ms = T.last_timings()[1]
nu_edge = np.asarray(st["cont_nu_edge"])
terms = np.searchsorted(nu_edge, pk["nu_cmf"], side="right").astype(np.float64)
print(f"{n} evaluations, {src.size} distinct packets, mean terms {terms.mean():.0f} max {terms.max():.0f}: "
      f"{ms:.2f} ms -> {n / ms / 1e3:.2f} M evals/s, {terms.sum() / ms / 1e6:.2f} G terms/s; kappa checksum {k[:, 0].sum():.6e}")
