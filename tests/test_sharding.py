"""N>1 path on CPU: two gloo ranks each transport half of the packets with the oracle's restatement standing in for the
device (the host-side sharding logic is backend independent), sum their estimators like mpi_reduce_estimators, and
must reproduce the unsharded run: packet for packet, and estimator sums up to the order of floating-point additions."""
import os
import sys

import numpy as np
import pytest

import goldenlib
import shardlib as sharding

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
FIXTURE = "classic_nogrey_ts12.npz"
SEED = 4242


def test_shard_bounds_cover_without_overlap():
    for n in (0, 1, 7, 15000, 10 ** 6 + 3):
        for world in (1, 2, 3, 8):
            cuts = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(10, 2, 2)


def _worker(rank, world, port, lib, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, REPO)
    sys.path.insert(0, HERE)
    import torch.distributed as dist
    from artis_b200.transport import Transport
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fx = goldenlib.load_fixture(FIXTURE)
    T = Transport(lib)
    # the rank enters the Philox counter; the comparison run below is a single shard, so both ranks use its key
    T.init(fx["static"], rank=0, nranks=1, seed=SEED)
    T.upload_cellstate(fx["cell"])
    T.zero_estimators()
    lo, hi = sharding.shard_bounds(fx["pre"].size, rank, world)
    pk = fx["pre"][lo:hi].copy()
    T.update_packets(pk)
    est = sharding.allreduce_estimators(T.download_estimators(), dist)
    np.save(os.path.join(outdir, f"pk{rank}.npy"), pk)
    if rank == 0:
        np.savez(os.path.join(outdir, "est.npz"), **{k: est[k] for k in sharding.REDUCED_FIELDS if k in est})
    dist.barrier()
    dist.destroy_process_group()
    T.finalize()


def test_two_gloo_ranks_reproduce_the_unsharded_step(oracle_lib, tmp_path):
    import torch.multiprocessing as mp
    from artis_b200.transport import Transport
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, oracle_lib, str(tmp_path)), nprocs=2, join=True)
    fx = goldenlib.load_fixture(FIXTURE)
    T = Transport(oracle_lib)
    T.init(fx["static"], rank=0, nranks=1, seed=SEED)
    T.upload_cellstate(fx["cell"])
    T.zero_estimators()
    whole = fx["pre"].copy()
    T.update_packets(whole)
    est1 = T.download_estimators()
    T.finalize()
    parts = np.concatenate([np.load(tmp_path / "pk0.npy"), np.load(tmp_path / "pk1.npy")])
    assert parts.size == whole.size
    for name in whole.dtype.names:  # independent packets: sharding does not change any history (padding aside)
        assert np.array_equal(parts[name], whole[name], equal_nan=True), name
    est2 = np.load(tmp_path / "est.npz")
    for k in est2.files:
        a, b = np.asarray(est1[k], dtype=np.float64), np.asarray(est2[k], dtype=np.float64)
        if k in ("counters",):
            # device work counters are sums too; pass/launch bookkeeping (slot 46 and above 50) is per process
            a, b = a[:46], b[:46]
        assert np.allclose(a, b, rtol=1e-12, atol=0), k
