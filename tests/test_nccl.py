"""The library's own exchange step on two GPUs: artis_gpu_comm_init + artis_gpu_reduce_estimators (the replacement of
mpi_reduce_estimators, sn3d.cc:316-384) and the all-reduced light curves / spectrum of artis_gpu_bin_escaped
(spectrum.cc:688-699). Each rank transports half of the fixture's packets; the reduced estimator block must equal the
sum of the two ranks' own blocks, and the reduced spectrum the spectrum of the whole ensemble. Skipped on a box with
one GPU (the gloo test in test_sharding.py covers the host-side logic there)."""
import os
import sys

import numpy as np
import pytest

import goldenlib
import shardlib as sharding

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
FIXTURE = "classic_nogrey_ts12.npz"
FP64_FIELDS = ("J", "nuJ", "ffheatingestimator", "colheatingestimator", "rpkt_emiss", "gammaestimator", "bfheatingestimator",
               "ts_scalars")


def _worker(rank, world, port, lib, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, REPO)
    sys.path.insert(0, HERE)
    import torch.distributed as dist
    from artis_b200.transport import Transport
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only carries the NCCL id, like MPI_Bcast would
    fx = goldenlib.load_fixture(FIXTURE)
    T = Transport(lib)
    T.init(fx["static"], device=rank, rank=rank, nranks=world, seed=99)
    ids = [T.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    T.comm_init(ids[0], rank, world)
    T.upload_cellstate(fx["cell"])
    T.zero_estimators()
    lo, hi = sharding.shard_bounds(fx["pre"].size, rank, world)
    pk = fx["pre"][lo:hi].copy()
    T.packets_upload(pk)
    T.transport_resident()
    own = T.download_estimators()
    tmax = float(fx["static"]["ts_start"][-1] + fx["static"]["ts_width"][-1])
    lc_own, flux_own, _ = T.bin_escaped(tmax, world, 1000, allreduce=False)
    T.reduce_estimators()
    red = T.download_estimators()
    lc_red, flux_red, _ = T.bin_escaped(tmax, world, 1000, allreduce=True)
    np.savez(os.path.join(outdir, f"r{rank}.npz"), lc_own=lc_own, flux_own=flux_own, lc_red=lc_red, flux_red=flux_red,
             counters_own=own["counters"], counters_red=red["counters"],
             **{f"own_{k}": own[k] for k in FP64_FIELDS}, **{f"red_{k}": red[k] for k in FP64_FIELDS})
    dist.barrier()
    dist.destroy_process_group()
    T.finalize()


@pytest.mark.gpu
def test_two_gpus_reduce_estimators_and_spectra_over_nccl(gpu_lib, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, gpu_lib, str(tmp_path)), nprocs=2, join=True)
    r = [np.load(os.path.join(str(tmp_path), f"r{k}.npz")) for k in range(2)]
    for k in FP64_FIELDS:
        want = r[0][f"own_{k}"] + r[1][f"own_{k}"]
        assert want.any() or k == "colheatingestimator", k
        for rank in range(2):  # every rank holds the sum afterwards (an all-reduce, not a reduce to rank 0)
            assert np.allclose(r[rank][f"red_{k}"], want, rtol=1e-14, atol=0), k
    n_event = 34  # stats::eventcounters: whole numbers, exact
    assert np.array_equal(r[0]["counters_red"][:n_event], r[0]["counters_own"][:n_event] + r[1]["counters_own"][:n_event])
    assert np.array_equal(r[0]["counters_red"][:n_event], r[1]["counters_red"][:n_event])
    for key in ("lc", "flux"):
        want = r[0][f"{key}_own"] + r[1][f"{key}_own"]
        assert want.any()
        for rank in range(2):
            assert np.allclose(r[rank][f"{key}_red"], want, rtol=1e-14, atol=0), key
