"""TEST HELPER (not part of the product package). Packet sharding across ranks: every rank transports its own packets with the grid, atomic data and
cell tables replicated; the only exchange of a timestep is the sum of the estimator arrays, the reference's
mpi_reduce_estimators (sn3d.cc:316-384, called at sn3d.cc:582). On GPUs that sum is one NCCL all-reduce of the
contiguous estimator block (Transport.estimator_block, used by bench.py); this module holds the backend-independent
host logic, which tests/test_sharding.py runs over gloo with world_size 2.
"""
import numpy as np

# the members of ArtisEstimators that mpi_reduce_estimators sums over ranks
REDUCED_FIELDS = ("J", "nuJ", "ffheatingestimator", "colheatingestimator", "rpkt_emiss", "gammaestimator",
                  "bfheatingestimator", "ts_scalars", "counters", "ecounter", "acounter")


def shard_bounds(npackets, rank, world):
    """contiguous balanced split [lo, hi) of a packet array; the first npackets % world ranks get one more"""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(int(npackets), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_estimators(est, dist):
    """sum every reduced estimator array over the ranks of torch.distributed's default group, in place"""
    import torch
    for name in REDUCED_FIELDS:
        if name not in est:
            continue
        a = np.ascontiguousarray(est[name])
        t = torch.from_numpy(a)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        est[name] = t.numpy()
    return est
