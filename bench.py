#!/usr/bin/env python
"""bench.py -- throughput of the ARTIS update_packets hot path on B200, next to the reference on the host cores.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--packets P] [--impl reference]

A "step" is one pass of the hot path (one update_packets call, update_packets.cc:234) over one batch of packets of
the workload timestep. The workload is a snapshot of the inputs of that call, taken from a run of the unmodified
reference (oracle/refglue/hook.cc) on the repo's classic-mode test model with synthetic atomic data, and replicated
to P statistically identical packets per GPU (weak scaling: packets are sharded, grid and atomic data replicated,
one estimator all-reduce per step -- SURVEY 8e).

  value      interactions/s, whole job, packets resident in HBM when the timed region starts (device time, CUDA events
             inside the library on the stream the kernels run on, max over ranks; + the NCCL all-reduce for N > 1)
  e2e        the same through artis_gpu_update_packets with HOST packet buffers: H2D of the 304-byte packets, all
             passes, D2H of the packets and of the estimator block inside the timed region
  roofline   algorithmic bytes (SURVEY 8d formula over the library's device work counters) / transport-kernel time
  cpu_baseline / --impl reference : the reference's own update_packets on the host cores (oracle/_ref)
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

METRIC = "rpkt_interactions_per_s"
UNIT = "interactions/s"
BENCH_RUN = os.path.join(REPO, "oracle", "_ref", "runs", "bench_base")
BENCH_TS = 25
REF_EXE = os.path.join(REPO, "oracle", "_ref", "classic", "sn3d_ref")
WORKLOAD_NAME = ("classicmode W7-like 1D model (tests/classicmode_inputfiles, artisoptions_classic.h, 100^3 grid), "
                 "synthetic atomic data {nlines} lines / {nbf} bf continua / {nlev} levels, timestep {ts}")
FIXTURE = None  # --workload fixture: a committed golden fixture instead of the named workload (smoke runs only)
MODEL3D = False  # --workload 3d: BASELINE.json config 4, the classic workload promoted to one model cell per propagation cell


def select_workload(name):
    """--workload nebular: BASELINE.json config 2, the one-zone nebular NLTE model (tests/nebularonezone_inputfiles,
    artisoptions_nltenebular.h with the overrides of tests/setup_nebularonezone.sh, 50^3 grid) at timestep 8; the
    default stays the classic-mode model, whose timestep is the interaction-rich one the metric is about.
    --workload fixture: the 16^3-grid parity fixture tests/golden/classic_nogrey_ts12.npz; never the default, and its
    line says what it is (it has no reference arm)."""
    global BENCH_RUN, BENCH_TS, REF_EXE, WORKLOAD_NAME, FIXTURE, MODEL3D
    if name == "3d":
        # BASELINE.json config 4: tools/model3d.py promote() of the classic workload -- the state of the
        # classic run's timestep sampled onto a 3-D model with one model cell per non-empty propagation cell (5.2e5 of
        # them), 1.25e7 packets per GPU spread over the sphere. The reference arm of this workload is the classic 1-D-model
        # run it is promoted from (same physics per packet; the reference on the promoted model is hours of host update_grid).
        MODEL3D = True
        WORKLOAD_NAME = ("config 4: 3-D model, one model cell per non-empty cell of the 100^3 grid, promoted (tools/model3d.py) "
                         "from the " + WORKLOAD_NAME)
        os.environ.setdefault("ARTIS_ROW_ARENA_MB", "65536")
    elif name == "nebular":
        BENCH_RUN = os.path.join(REPO, "oracle", "_ref", "runs", "bench_neb")
        BENCH_TS = 8
        REF_EXE = os.path.join(REPO, "oracle", "_ref", "nebular", "sn3d_ref")
        WORKLOAD_NAME = ("nebular one-zone NLTE model (tests/nebularonezone_inputfiles, artisoptions_nltenebular.h with the "
                         "overrides of tests/setup_nebularonezone.sh, 50^3 grid), synthetic atomic data {nlines} lines / "
                         "{nbf} bf continua / {nlev} levels, timestep {ts}")
    elif name == "vpkt":
        # BASELINE.json config 5: the classic model with VPKT_ON + POL_ON + DIPOLE, five observers, at
        # timestep 30 (inside the compiled 10-30 d window of vpkt.h: every eligible event sends its five virtual packets)
        BENCH_RUN = os.path.join(REPO, "oracle", "_ref", "runs", "bench_vpkt")
        BENCH_TS = 30
        REF_EXE = os.path.join(REPO, "oracle", "_ref", "classic_vpkt", "sn3d_ref")
        WORKLOAD_NAME = ("config 5: " + WORKLOAD_NAME + ", VPKT_ON + POL_ON + DIPOLE, vpkt.txt with 5 observers x 1 spectrum "
                         "(SURVEY S5)")
    elif name == "kilonova":
        # BASELINE.json config 3: the kilonova LTE build mode on the 1-D kilonova test model with a line list
        # of more than 1e6 lines (the closest_transition / line-sweep stress), timestep 7
        BENCH_RUN = os.path.join(REPO, "oracle", "_ref", "runs", "bench_kn")
        BENCH_TS = 7
        REF_EXE = os.path.join(REPO, "oracle", "_ref", "kilonova", "sn3d_ref")
        WORKLOAD_NAME = ("config 3: kilonova LTE 1-D model (tests/kilonova_inputfiles, artisoptions_kilonova_lte.h with the "
                         "overrides of tests/setup_kilonova.sh, 50^3 grid), synthetic atomic data {nlines} lines / {nbf} bf "
                         "continua / {nlev} levels, timestep {ts}")
    elif name == "fixture":
        FIXTURE = "classic_nogrey_ts12.npz"


# ---------------------------------------------------------------------------------------------------------------------
def load_workload():
    """(static, cell, packets_pre, name). The bench snapshot lives under oracle/_ref/runs (git-ignored; made by
    __graft_entry__.build() where /root/reference exists and carried to the GPU box with the working tree). A missing
    snapshot is an error: the bench never measures another workload under the named one's flag."""
    from artis_b200 import snapshot
    if FIXTURE is not None:
        sys.path.insert(0, os.path.join(REPO, "tests"))
        import goldenlib
        fx = goldenlib.load_fixture(FIXTURE)
        return fx["static"], fx["cell"], fx["pre"], f"parity fixture tests/golden/{FIXTURE} (16^3 grid; --workload fixture)"
    snap = os.path.join(BENCH_RUN, f"snap_ts{BENCH_TS}.bin")
    if not os.path.exists(snap):
        raise SystemExit(f"bench.py: the workload snapshot {snap} is missing. It is a run of the unmodified reference made by "
                         "__graft_entry__.build() (oracle/tools/make_bench_run.sh / make_bench_neb.sh) where /root/reference "
                         "exists; run build() there before the bench. (`--workload fixture` runs the small parity fixture "
                         "instead and says so.)")
    s = snapshot.load(snap)
    st, cs = snapshot.static_tables(s), snapshot.cell_state(s)
    name = WORKLOAD_NAME.format(nlines=st["nlines"], nbf=st["nbfcontinua"], nlev=st["nlevels_total"], ts=BENCH_TS)
    if os.environ.get("ARTIS_BENCH_VPKT_OFF") and st.get("opt_vpkt_on"):  # A/B only: what the virtual packets of config 5 cost
        st["opt_vpkt_on"] = 0
        name += " [A/B: virtual packets switched OFF]"
    if MODEL3D:
        sys.path.insert(0, os.path.join(REPO, "tools"))
        import model3d
        st, cs, _ = model3d.promote(st, cs)
        name += f"; {int(st['npts_model'])} model cells"
    return st, cs, s["packets.pre"].copy(), name


def workload_key(name, packets):
    """identifies (workload, packets per GPU): the key under which an ncu DRAM-traffic capture is filed"""
    import hashlib
    return hashlib.sha1(f"{name}|{int(packets)}".encode()).hexdigest()[:16]


def replicate(pre, P, rank, step, static=None):
    """P statistically identical packets: the snapshot's packets tiled, each copy with its own packet.number (its
    Philox stream) and with energies scaled so that the total energy is that of the snapshot (packet.cc:106-108)."""
    reps = -(-P // pre.size)
    out = np.tile(pre, reps)[:P].copy()
    scale = pre.size / float(P)
    for f in ("e_cmf", "e_rf"):
        out[f] *= scale
    base = (step * 8 + rank) * P
    out["number"] = (np.arange(P, dtype=np.int64) + base).astype(np.int32)
    if MODEL3D and static is not None:  # the copies of a packet go to different places of the (spherically symmetric) model
        import model3d
        model3d.rotate_copies(out, pre.size, static, seed=20240607 + step * 8 + rank)
    return out


TRAFFIC_FILE = os.path.join("profiles", "r2_traffic.json")


def measured_traffic(name, packets):
    """dram__bytes_read.sum + dram__bytes_write.sum of one transport launch from the committed ncu capture of THIS
    workload at THIS packet count (matched on workload_key); None otherwise -- the number is never borrowed from a
    capture of something else"""
    try:
        for t in json.load(open(os.path.join(REPO, TRAFFIC_FILE))):
            if t.get("workload_key") == workload_key(name, packets):
                return float(t["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


def algorithmic_bytes_every_continuum(ctr):
    """the same formula with n_bf = the continua the REFERENCE's loop visits for the same opacity evaluations (device counter
    ARTIS_CTR_BF_LISTED): the library's own sums run over the live continua of the cell only, so its n_bf -- and
    with it the bytes its algorithm needs -- is several times smaller on the classic model"""
    return algorithmic_bytes(ctr) + 80 * (ctr[51] - ctr[38]) if ctr[51] > 0 else algorithmic_bytes(ctr)


def algorithmic_bytes(ctr):
    """B = 208 n_rpkt_step + 8 n_scan + 64 n_tau + 80 n_bf + 16 n_est + (72 n_ma_jump + 8 n_ma_probe)
    + (72 n_kpkt + 8 n_cool_probe) + 240 n_gamma_step, all exact device counters of the timed steps."""
    return (208 * ctr[35] + 8 * ctr[36] + 64 * ctr[37] + 80 * ctr[38] + 16 * ctr[39] + 240 * ctr[40] + 72 * ctr[41] +
            8 * ctr[42] + 72 * ctr[43] + 8 * ctr[44])


class ClockSampler(threading.Thread):
    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.rows, self.stop_flag = device, [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.device)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
def run_reference(nproc, steps, warmup):
    """Times the UNMODIFIED reference (oracle/_ref/classic/sn3d_ref: the reference's own sources built against the
    GSL shim) on the host cores: `nproc` independent single-thread processes (= MPI ranks of the reference: packets
    sharded, everything else replicated; MPI itself is not in the image) each transporting the workload timestep from
    the same checkpoint with its own seed. Returns (interactions, seconds, sample description)."""
    if FIXTURE is not None or not (os.path.exists(REF_EXE) and
                                   os.path.exists(os.path.join(BENCH_RUN, f"gridsave_ts{BENCH_TS}.tmp"))):
        return None
    keep = ("adata.txt", "transitiondata.txt", "phixsdata_v2.txt", "compositiondata.txt", "model.txt", "abundances.txt",
            "input.txt", "ratecoeff.dat", "bflist.dat", "recombrates.txt", "vpkt.txt", "gridcontributions.txt",
            "vspecpol_0_0_even.tmp", "vspecpol_0_0_odd.tmp", f"gridsave_ts{BENCH_TS}.tmp",
            f"packets_0000_ts{BENCH_TS}.tmp")
    tot_inter, tot_sec = 0.0, 0.0
    npk = 0
    with tempfile.TemporaryDirectory(prefix="artis_ref_") as tmp:
        dirs = []
        for r in range(nproc):
            d = os.path.join(tmp, f"rank{r}")
            os.makedirs(d)
            for f in os.listdir(BENCH_RUN):
                if f in keep or f.endswith("_lines.txt") or f.endswith("decays.txt") or f.startswith(("auger", "collion",
                                                                                                       "binding")):
                    shutil.copy(os.path.join(BENCH_RUN, f), d)
            dirs.append(d)
        for it in range(warmup + steps):
            procs = []
            for r, d in enumerate(dirs):
                lines = open(os.path.join(d, "input.txt")).read().splitlines()
                for i, line in enumerate(lines):
                    if line.strip() and not line.strip().startswith("#"):
                        lines[i] = f"{900001 + 7919 * (it * nproc + r)}".ljust(25) + "# pre_zseed"
                        break
                open(os.path.join(d, "input.txt"), "w").write("\n".join(lines) + "\n")
                if os.path.exists(os.path.join(d, "hook_timing.txt")):
                    os.remove(os.path.join(d, "hook_timing.txt"))
                env = dict(os.environ, ARTIS_HOOK_DIR=".", ARTIS_HOOK_STOP_AFTER_TS=str(BENCH_TS))
                procs.append(subprocess.Popen([REF_EXE], cwd=d, env=env, stdout=subprocess.DEVNULL,
                                              stderr=subprocess.DEVNULL))
            for p in procs:
                p.wait()
            inter, sec = 0.0, 0.0
            for d in dirs:
                row = open(os.path.join(d, "hook_timing.txt")).read().split()
                npk = int(row[1])
                inter += float(row[3])
                sec = max(sec, float(row[2]))  # the ranks run concurrently: the step takes as long as the slowest
            if it >= warmup:
                tot_inter += inter
                tot_sec += sec
    sample = (f"{nproc} concurrent single-thread sn3d_ref processes x {npk} packets, timestep {BENCH_TS} of the workload "
              f"run, update_packets call only (std::chrono in oracle/refglue/hook.cc), {steps} step(s) after {warmup} warm-up")
    return tot_inter, tot_sec, sample


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--packets", type=int, default=0, help="packets per GPU (default 1e6; 1.25e7 for --workload 3d)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-baseline", default="auto", choices=["auto", "off"])
    ap.add_argument("--max-events", type=int, default=0, help="events per packet and pass (0: library default)")
    ap.add_argument("--scheduler-kernels", type=int, default=0, choices=(0, 1, 2),
                    help="1: one persistent kernel with every role, 2: macro-atom jump kernel + every-other-role kernel (0: library default)")
    ap.add_argument("--workload", default="classic", choices=["classic", "nebular", "fixture", "3d", "vpkt", "kilonova"])
    a = ap.parse_args()
    select_workload(a.workload)
    if a.packets <= 0:
        a.packets = 12500000 if a.workload == "3d" else 1000000
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if a.impl == "reference":
        if rank != 0:
            return 0
        st, cs, pre, wname = load_workload()
        cores = host_cores()
        res = run_reference(cores, max(1, a.steps), min(a.warmup, 1))
        kind = "reference"
        if res is None:  # oracle/_ref did not travel: the CPU restatement of the path, one process per core
            res = run_port(st, cs, pre, cores, max(1, a.steps), min(a.warmup, 1))
            kind = "port"
        inter, sec, sample = res
        if MODEL3D:
            sample += "; the 1-D-model run the 3-D workload is promoted from"
        v = inter / sec
        line = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": 1e3 * sec / max(1, a.steps), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
                "config": {"workload": wname, "packets_per_rank": int(pre.size), "ranks": cores},
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist
    from artis_b200.transport import Transport
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    st, cs, pre, wname = load_workload()
    P = a.packets
    T = Transport()
    T.init(st, device=local_rank, rank=rank, nranks=world)
    T.upload_cellstate(cs)
    torch.cuda.synchronize()
    t0 = time.perf_counter()      # the other per-timestep call of a driver: cell state in, the per-cell tables rebuilt
    T.upload_cellstate(cs)
    torch.cuda.synchronize()
    cell_upload_wall_ms = (time.perf_counter() - t0) * 1e3
    if a.max_events:
        T.set_tuning(max_events_per_pass=a.max_events)
    if os.environ.get("ARTIS_HANG_S"):  # measurement runs of experimental builds: give up a stuck timestep quickly
        T.set_tuning(block_threads=-1000 - int(os.environ["ARTIS_HANG_S"]))
    if a.scheduler_kernels or os.environ.get("ARTIS_BENCH_SCHED"):
        T.set_tuning(block_threads=-3 if int(a.scheduler_kernels or os.environ["ARTIS_BENCH_SCHED"]) == 1 else -4)
    precompute_ms = T.last_timings()[3]
    ptr, ndoubles = T.estimator_block()
    if world > 1:
        # the exchange step goes through the library's own NCCL communicator (artis_gpu_comm_init /
        # artis_gpu_reduce_estimators, the replacement of mpi_reduce_estimators, sn3d.cc:316-384); torch.distributed only
        # hands the 128-byte NCCL id round and takes the max over ranks of the timings
        ids = [T.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        T.comm_init(ids[0], rank, world)
    # pinned host staging for the end-to-end leg
    nbytes = P * 304
    pinned = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    from artis_b200 import abi
    host_pk = pinned.numpy().view(abi.PACKET_DTYPE)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(step, timed):
        """device-resident step: returns (device ms of transport [+ all-reduce], interactions of this rank)"""
        pk = replicate(pre, P, rank, step, st)
        T.packets_upload(pk)  # untimed: inputs are resident in HBM when the timed region starts
        T.zero_estimators()
        barrier()
        T.transport_resident()
        ms = T.last_timings()[0]
        ar = 0.0
        if world > 1:
            T.reduce_estimators()          # one NCCL group on the transport stream, timed with CUDA events in the library
            ar = T.last_timings()[5]
        return ms + ar, ar

    for w in range(a.warmup):
        one_step(1000 + w, False)
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = T.launch_count()
    tot_ms = 0.0
    ctr_sum = np.zeros(96, dtype=np.int64)
    kernel_ms = 0.0
    allreduce_ms = 0.0
    for s in range(a.steps):
        ms, ar = one_step(s, True)
        allreduce_ms += ar
        kernel_ms += T.last_timings()[2]
        c = T.download_estimators()["counters"]
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        tot_ms += ms
        ctr_sum += c
    launches = T.launch_count() - launches0
    # light curves + spectrum of the packets left resident by the last step (artis_gpu_bin_escaped; outside the timed region)
    tmax_run = float(st["ts_start"][-1] + st["ts_width"][-1])
    for _ in range(3):
        lc_out, _, _ = T.bin_escaped(tmax_run, world, 1000, allreduce=False)
    bin_ms = T.last_timings()[4]
    # with N > 1 the library's all-reduce has summed the work counters over the shards as well (they ride in the same
    # NCCL group as the estimators): every rank holds the job's totals
    inter_all = float(ctr_sum[45])
    alg_bytes = float(algorithmic_bytes(ctr_sum)) / world  # of this rank's kernel (equal shards)
    alg_bytes_ref = float(algorithmic_bytes_every_continuum(ctr_sum)) / world
    value = inter_all / (tot_ms * 1e-3)

    # end-to-end through the C ABI with host buffers
    e2e_sec, e2e_inter = 0.0, 0.0
    for s in range(max(1, min(a.steps, 3)) + 1):
        host_pk[:] = replicate(pre, P, rank, 2000 + s, st)
        T.zero_estimators()
        barrier()
        t0 = time.perf_counter()
        T.update_packets(host_pk)                      # H2D packets, all passes, D2H packets
        if world > 1:
            T.reduce_estimators()
        T.download_estimators()                        # D2H of the step's result (the estimator blocks)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        if s > 0:  # first one is warm-up of the pinned path
            e2e_sec += dt
            e2e_inter += float(host_pk["interactions"].sum())
    if world > 1:
        t = torch.tensor([e2e_inter], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        e2e_inter = float(t.item())
    sampler.stop_flag = True
    sampler.join(timeout=2)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6537.6))
        achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": tot_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": wname, "packets_per_gpu": P, "parallelism": f"packets sharded x{world}, grid and "
                           "atomic data replicated, one estimator all-reduce per step",
                           "l2": "per-step working set (304 B x packets + per-cell tables) exceeds the 126 MB L2",
                           "s_per_timestep": (tot_ms / a.steps + cell_upload_wall_ms) * 1e-3,
                           "s_per_timestep_parts_ms": {"transport_and_reduce": tot_ms / a.steps, "upload_cellstate_wall": cell_upload_wall_ms,
                                                       "of_which_table_kernels": precompute_ms},
                           "cell_precompute_ms": precompute_ms,
                           "table_bytes": T.table_bytes(),
                           "rows_on_demand": ({"arena_restarts_per_step": float(ctr_sum[91]) / a.steps / world,
                                               "row_bytes_built_per_step": 128.0 * float(ctr_sum[92]) / a.steps / world}
                                              if ctr_sum[92] > 0 else None),
                           "estimator_block_bytes": int(ndoubles) * 8,
                           "passes_per_step": float(ctr_sum[46]) / a.steps / world, "handoffs_per_step": float(ctr_sum[50]) / a.steps / world,
                           "drain_ms_to_10pct_1pct_done": [round(float(c[61 + i]) * 1e-6, 1) for i in range(3)],
                           "warp_iters_by_kind": [float(x) / a.steps / world for x in ctr_sum[64:72]],
                           "mean_busy_lanes_by_kind": [round(float(ctr_sum[72 + i]) / max(float(ctr_sum[64 + i]), 1.0), 2) for i in range(8)],
                           "warp_Mcycles_by_kind": [round(float(x) / a.steps / world * 1e-6, 1) for x in ctr_sum[80:88]],
                           "rpkt_steps_and_kappa_evals_per_step": [float(ctr_sum[35]) / a.steps / world, float(ctr_sum[48]) / a.steps / world],
                           # -DARTIS_KAPPA_PROFILE builds only: prologue, barrier wait, term arithmetic, copy issue, owner set-up, reductions
                           "kappa_profile_Mcycles": ([round(float(ctr_sum[i]) / a.steps / world * 1e-6, 1) for i in (54, 55, 56, 58, 59, 60)]
                                                     if ctr_sum[54] > 0 and not st.get("opt_vpkt_on") else None),
                           "thin_role_Mcycles_begin_kappa": [round(float(x) / a.steps / world * 1e-6, 1) for x in ctr_sum[52:54]],
                           "warp_Mcycles_refill_push_idle": [round(float(x) / a.steps / world * 1e-6, 1) for x in ctr_sum[88:91]],
                           "scheduler_geometry_kernels_blocks_single_big_jump": T.scheduler_geometry(),
                           # -DARTIS_VPKT_PROFILE builds only (vpkt.cuh): Mcycles per step in boundary_cross + continuum sum, the two line
                           # sweeps, set-up + move; then rays and cells per step
                           "vpkt_requests_per_step": float(ctr_sum[57]) / a.steps / world if st.get("opt_vpkt_on") else None,
                           "vpkt_list_restarts_per_step": float(ctr_sum[58]) / a.steps / world if st.get("opt_vpkt_on") else None,
                           "vpkt_profile": ([round(float(x) / a.steps / world * 1e-6, 1) for x in ctr_sum[54:57]] +
                                            [float(x) / a.steps / world for x in ctr_sum[59:61]]) if ctr_sum[59] > 0 else None,
                           "allreduce_ms_per_step": allreduce_ms / a.steps,
                           "estimator_block_bytes": int(ndoubles * 8),
                           "events_per_s": float(ctr_sum[49]) / (tot_ms * 1e-3),
                           "bin_escaped_ms": bin_ms, "bin_escaped_record_GBps": P * 304 / (bin_ms * 1e-3) / 1e9 if bin_ms > 0 else None},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": measured_traffic(wname, P),
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6537.6",
                             "kernel": "k_transport_persistent", "launches_per_step": 1,
                             "algorithmic_bytes_per_launch": alg_bytes / a.steps,
                             # n_bf of the formula counted as the reference counts it (every continuum in range, exact zeros
                             # included) instead of the live continua the kernel evaluates: comparable with earlier rounds
                             "algorithmic_bytes_per_launch_every_continuum": alg_bytes_ref / a.steps,
                             "frac_every_continuum": (alg_bytes_ref / (kernel_ms * 1e-3) / 1e9 / peak) if kernel_ms > 0 else 0.0,
                             "traffic_source": TRAFFIC_FILE + " (ncu dram bytes per launch, filed under the workload key; "
                                               "null when no capture of this workload and packet count is committed)",
                             "workload_key": workload_key(wname, P)},
                "e2e": {"value": e2e_inter / e2e_sec, "unit": UNIT, "h2d_bytes_per_step": nbytes,
                        "d2h_bytes_per_step": nbytes + ndoubles * 8},
                "gpu_launches": int(launches), "clocks": sampler.summary()}
        if world == 1 and a.cpu_baseline != "off":
            # the reference's own update_packets on every host core of this box: one warm-up step, one timed step
            cores = host_cores()
            res = run_reference(cores, 1, 1)
            kind = "reference"
            if res is None and MODEL3D:
                raise SystemExit("bench.py --workload 3d: oracle/_ref (the reference build and its run) is missing")
            if res is None:
                res = run_port(st, cs, pre, cores, 1, 0)
                kind = "port"
            inter, sec, sample = res
            if MODEL3D:
                sample += "; the 1-D-model run the 3-D workload is promoted from"
            line["cpu_baseline"] = {"value": inter / sec, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def _port_worker(args):
    st, cs, pre, seed = args
    from artis_b200.transport import Transport
    O = Transport(os.path.join(REPO, "oracle", "libartis_oracle.so"))
    O.init(st, seed=seed)
    O.upload_cellstate(cs)
    O.zero_estimators()
    pk = pre.copy()
    t0 = time.perf_counter()
    O.update_packets(pk)
    return float(pk["interactions"].sum()), time.perf_counter() - t0


def run_port(st, cs, pre, nproc, steps, warmup):
    """Stand-in for run_reference when oracle/_ref did not travel: the oracle's CPU restatement of the path
    (oracle/libartis_oracle.so), `nproc` single-thread processes each transporting the workload's packets with its own
    seed. Returns (interactions, seconds, sample description)."""
    import multiprocessing as mp
    lib = os.path.join(REPO, "oracle", "libartis_oracle.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-C", os.path.join(REPO, "oracle"), "restatement"], check=True, stdout=subprocess.DEVNULL)
    tot_inter, tot_sec = 0.0, 0.0
    with mp.get_context("spawn").Pool(nproc) as pool:
        for it in range(warmup + steps):
            rows = pool.map(_port_worker, [(st, cs, pre, 900001 + 7919 * (it * nproc + r)) for r in range(nproc)])
            if it >= warmup:
                tot_inter += sum(r[0] for r in rows)
                tot_sec += max(r[1] for r in rows)
    return tot_inter, tot_sec, (f"oracle restatement (CPU port), {nproc} concurrent single-thread processes x {pre.size} "
                                f"packets, {steps} step(s) after {warmup} warm-up")


if __name__ == "__main__":
    sys.exit(main())
