"""Thin ctypes binding of the C ABI (include/artis_gpu.h) -- plumbing for tests/ and bench.py.

The product is the shared library `artis_b200/libartis_gpu.so` (built from artis_b200/csrc/transport.cu by
`__graft_entry__.build()`); the reference-side caller is the C++ glue of INTEGRATION.md. This module only moves
numpy arrays across the same boundary and mirrors the reference's call order around `update_packets`
(sn3d.cc:543-582): upload_cellstate -> zero_estimators -> update_packets -> reduce/download estimators.
There is no fallback: if the library is missing or no CUDA device is present, construction / init raises.
"""
import ctypes as C
import os

import numpy as np

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.environ.get("ARTIS_GPU_LIB", os.path.join(HERE, "libartis_gpu.so"))  # override: kernel tuning experiments
# VPKT_ON is a compile-time switch of the reference (artisoptions.h) and of this library: the virtual-packet ray trace
# costs the other build modes 5-19 % when it is merely present in the transport kernels (profiles/r1_notes.md), so it is
# compiled into a second library from the same sources. A VPKT_ON driver links this one instead.
VPKT_LIB = os.path.join(HERE, "libartis_gpu_vpkt.so")
ROWS_LIB = os.path.join(HERE, "libartis_gpu_rows.so")  # macro-atom / cooling rows on demand: models of ~1e5 cells and more
# the switches of artisoptions_classic.h compiled in, as the reference compiles them in (7 % faster than testing them at
# run time: the kernels are bound by instruction supply, profiles/r2_notes.md); chosen when the static tables carry exactly
# that option set. ARTIS_GPU_GENERIC=1 keeps the run-time-option library for every mode.
CLASSIC_LIB = os.path.join(HERE, "libartis_gpu_classic.so")


def is_classic_option_set(static):
    g = lambda k: int(np.asarray(static.get(k, 0)).reshape(-1)[0])  # noqa: E731
    return (g("opt_grid_type") == 0 and not g("opt_nt_on") and not g("opt_nlte_pops_on") and not g("opt_no_lut_photoion")
            and not g("opt_no_lut_bfheating") and not g("opt_relativistic_doppler") and g("opt_instant_particle_deposition")
            and not g("opt_detailed_bf_estimators") and not g("opt_multibin_radfield") and not g("opt_vpkt_on"))


class ArtisGpuError(RuntimeError):
    pass


class Transport:
    def __init__(self, libpath=None):
        path = libpath or DEFAULT_LIB
        if not os.path.exists(path):
            raise ArtisGpuError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(there is no CPU fallback)")
        self._load(path)
        self._keep = []
        self.static = None
        self.npkts = 0

    # ---- helpers
    def _load(self, path):
        self.libpath = path
        self.lib = C.CDLL(path)
        abi.declare(self.lib)
        self.lib.artis_gpu_launch_count.restype = C.c_longlong
        self.lib.artis_gpu_table_bytes.restype = C.c_size_t

    def _check(self, rc, what):
        if rc != 0:
            buf = C.create_string_buffer(1024)
            self.lib.artis_gpu_last_error(buf, 1024)
            raise ArtisGpuError(f"{what} failed with code {rc}: {buf.value.decode()}")

    # ---- life cycle
    def init(self, static, device=0, rank=0, nranks=1, seed=1281360349):
        static = dict(static)
        static["abi_version"] = abi.ABI_VERSION
        base = os.path.join(HERE, "libartis_gpu.so")
        want = VPKT_LIB if static.get("opt_vpkt_on") else base
        if want == base and is_classic_option_set(static) and os.path.exists(CLASSIC_LIB) and not os.environ.get("ARTIS_GPU_GENERIC"):
            want = CLASSIC_LIB
        if os.path.realpath(self.libpath) in (os.path.realpath(base), os.path.realpath(VPKT_LIB), os.path.realpath(ROWS_LIB),
                                              os.path.realpath(CLASSIC_LIB)) and \
                os.path.realpath(self.libpath) != os.path.realpath(want):
            if not os.path.exists(want):
                raise ArtisGpuError(f"{want} not found: build it with __graft_entry__.build() (there is no CPU fallback)")
            self.lib.artis_gpu_finalize()
            self._load(want)  # the library that matches the build mode, as a driver would link it
        st, keep = abi.build_struct(abi.ArtisStaticTables, abi.STATIC_FIELDS, static)
        self._keep.append(keep)
        self.static = static
        rc = self.lib.artis_gpu_init(C.byref(st), device, rank, nranks, C.c_ulonglong(seed))
        if rc == abi.ERR_UNSUPPORTED and os.path.realpath(self.libpath) in (os.path.realpath(base), os.path.realpath(CLASSIC_LIB)) \
                and os.path.exists(ROWS_LIB):
            buf = C.create_string_buffer(1024)
            self.lib.artis_gpu_last_error(buf, 1024)
            if b"libartis_gpu_rows.so" in buf.value:  # the model is too large for dense rows (or ARTIS_TABLES=sparse)
                self.lib.artis_gpu_finalize()
                self._load(ROWS_LIB)
                rc = self.lib.artis_gpu_init(C.byref(st), device, rank, nranks, C.c_ulonglong(seed))
        self._check(rc, "artis_gpu_init")

    def finalize(self):
        self.lib.artis_gpu_finalize()

    def upload_cellstate(self, cell):
        cs, keep = abi.build_struct(abi.ArtisCellState, abi.CELL_FIELDS, dict(cell))
        self._check(self.lib.artis_gpu_upload_cellstate(C.byref(cs)), "artis_gpu_upload_cellstate")
        self.nts = int(cell["nts"])

    def zero_estimators(self):
        self._check(self.lib.artis_gpu_zero_estimators(), "artis_gpu_zero_estimators")

    # ---- the hot path
    def update_packets(self, packets, nts=None, rank=0):
        """Drop-in for update_packets(my_rank, nts, packets): host array in, host array out (modified in place)."""
        assert packets.dtype == abi.PACKET_DTYPE and packets.flags.c_contiguous
        nts = self.nts if nts is None else nts
        self._check(self.lib.artis_gpu_update_packets(rank, nts, packets.ctypes.data_as(C.c_void_p), packets.size),
                    "artis_gpu_update_packets")
        return packets

    def packets_upload(self, packets):
        assert packets.dtype == abi.PACKET_DTYPE and packets.flags.c_contiguous
        self._check(self.lib.artis_gpu_packets_upload(packets.ctypes.data_as(C.c_void_p), packets.size),
                    "artis_gpu_packets_upload")
        self.npkts = packets.size

    def transport_resident(self, nts=None):
        self._check(self.lib.artis_gpu_transport_resident(self.nts if nts is None else nts),
                    "artis_gpu_transport_resident")

    def packets_download(self, out=None):
        if out is None:
            out = np.zeros(self.npkts, dtype=abi.PACKET_DTYPE)
        self._check(self.lib.artis_gpu_packets_download(out.ctypes.data_as(C.c_void_p), out.size),
                    "artis_gpu_packets_download")
        return out

    # ---- packet files from the resident ensemble (sn3d.cc:387-398, packet.cc:152-209)
    def packets_write_temp(self, path):
        self._check(self.lib.artis_gpu_packets_write_temp(os.fsencode(path)), "artis_gpu_packets_write_temp")

    def packets_read_temp(self, path, npkts):
        self._check(self.lib.artis_gpu_packets_read_temp(os.fsencode(path), int(npkts)), "artis_gpu_packets_read_temp")
        self.npkts = int(npkts)

    def packets_write_text(self, path):
        self._check(self.lib.artis_gpu_packets_write_text(os.fsencode(path)), "artis_gpu_packets_write_text")

    def last_timings(self):
        ms = (C.c_double * 6)()
        n = self.lib.artis_gpu_last_timings(ms, 6)
        return [ms[i] for i in range(n)]

    def set_tuning(self, max_events_per_pass=0, block_threads=0):
        self.lib.artis_gpu_set_tuning(max_events_per_pass, block_threads)

    def scheduler_geometry(self):
        """(kernels per timestep, blocks per SM of: the single kernel, the every-other-role kernel, the jump kernel)"""
        g = (C.c_int * 4)()
        self.lib.artis_gpu_scheduler_geometry(g, 4)
        return [g[i] for i in range(4)]

    def launch_count(self):
        return int(self.lib.artis_gpu_launch_count())

    def table_bytes(self):
        return int(self.lib.artis_gpu_table_bytes())

    # ---- estimators
    def comm_unique_id(self):
        """rank 0: the 128-byte NCCL id that every shard passes to comm_init (hand it round like MPI_Bcast would)"""
        buf = C.create_string_buffer(128)
        self._check(self.lib.artis_gpu_comm_unique_id(buf), "artis_gpu_comm_unique_id")
        return buf.raw

    def comm_init(self, unique_id, rank, nranks):
        self._check(self.lib.artis_gpu_comm_init(C.create_string_buffer(unique_id, 128), rank, nranks), "artis_gpu_comm_init")

    def reduce_estimators(self, nts=None):
        self._check(self.lib.artis_gpu_reduce_estimators(self.nts if nts is None else nts), "artis_gpu_reduce_estimators")

    def download_estimators(self):
        s = self.static
        nonempty = int(s["nonempty_npts_model"])
        nbins = nonempty * int(s["opt_radfieldbincount"]) if s.get("opt_multibin_radfield") else 0
        nbfrate = nonempty * int(s["nbfcontinua"]) if s.get("opt_detailed_bf_estimators") else 0
        nvstokes = nvgrid = 0
        if s.get("opt_vpkt_on"):
            nvstokes = 3 * int(s["vpkt_vmtbins"]) * int(s["vpkt_nspectra"]) * int(s["vpkt_nobs"]) * int(s["vpkt_vmnubins"])
            if int(s["vpkt_vgrid_flag"]) == 1:
                nvgrid = 3 * int(s["vpkt_ny_vgrid"]) * int(s["vpkt_nz_vgrid"]) * int(s["vpkt_nrange_grid"]) * int(s["vpkt_nobs"])
        d = abi.empty_estimators(s["npts_model"], s["nelements"], s["max_nions"], s["nlines"], nbins, nbfrate, nvstokes,
                                 nvgrid)
        est, keep = abi.build_struct(abi.ArtisEstimators, abi.ESTIM_FIELDS, d, writable=True)
        self._check(self.lib.artis_gpu_download_estimators(C.byref(est)), "artis_gpu_download_estimators")
        return d

    def upload_estimators(self, est):
        """dict of numpy arrays (the names of ArtisEstimators) -> the device-resident estimators"""
        s = self.static
        d = dict(est)
        for k, v in (("npts_model", s["npts_model"]), ("nelements", s["nelements"]), ("max_nions", s["max_nions"]),
                     ("nlines", s["nlines"]), ("nradfieldbins", 0), ("nbfrate", 0), ("nvstokes", 0), ("nvgrid", 0)):
            d.setdefault(k, v)
        st, keep = abi.build_struct(abi.ArtisEstimators, abi.ESTIM_FIELDS, d)
        self._check(self.lib.artis_gpu_upload_estimators(C.byref(st)), "artis_gpu_upload_estimators")

    def normalise_estimators(self, vol_init_modelcell, nprocs, maxtemp):
        """-> (cells[npts_model, 6] = J, nuJ, ff, col, rpkt_emiss normalised and T_J; gamma_dep_pathint)"""
        nm = int(self.static["npts_model"])
        vol = np.ascontiguousarray(vol_init_modelcell, np.float64)
        assert vol.size >= nm
        out = np.zeros((nm, 6))
        g = np.zeros(1)
        self._check(self.lib.artis_gpu_normalise_estimators(_pd(vol), int(nprocs), float(maxtemp), _pd(out), _pd(g)),
                    "artis_gpu_normalise_estimators")
        return out, float(g[0])

    def cooling_rates(self, install=False):
        """-> (totalcooling[npts_model + 1], cooling_contrib_ion[npts_model + 1, includedions]) of the uploaded cell state"""
        nm, ni = int(self.static["npts_model"]) + 1, int(self.static["includedions"])
        total, contrib = np.zeros(nm), np.zeros((nm, ni))
        self._check(self.lib.artis_gpu_cooling_rates(_pd(total), _pd(contrib), 1 if install else 0), "artis_gpu_cooling_rates")
        return total, contrib

    def estimator_block(self):
        ptr = C.c_void_p()
        n = C.c_size_t()
        self._check(self.lib.artis_gpu_estimator_block(C.byref(ptr), C.byref(n)), "artis_gpu_estimator_block")
        return ptr.value, n.value

    # ---- light curves + spectrum of the resident packets (write_partial_lightcurve_spectra, spectrum.cc:641)
    def bin_escaped(self, tmax, nprocs=1, nnubins=1000, allreduce=False, stokes=False):
        """-> lc[4, ntstep] (r lum, r lumcmf, gamma lum, gamma lumcmf), flux[ntstep, nnubins], stokes[3, ntstep, nnubins] | None"""
        ntstep = int(self.static["ntstep"])
        lc, flux = np.zeros((4, ntstep)), np.zeros((ntstep, nnubins))
        sto = np.zeros((3, ntstep, nnubins)) if stokes else None
        self._check(self.lib.artis_gpu_bin_escaped(float(tmax), int(nprocs), int(nnubins), int(bool(allreduce)), _pd(lc), _pd(flux),
                                                   _pd(sto) if stokes else None), "artis_gpu_bin_escaped")
        return lc, flux, sto

    def bin_escaped_resolved(self, tmax, syn_dir, nprocs=1, nnubins=1000):
        """-> emission, trueemission [ntstep, nnubins, 2 * nel * maxion + 1], absorption [ntstep, nnubins, nel * maxion],
        lc_res [100, ntstep], flux_res [100, ntstep, nnubins]"""
        s = self.static
        ntstep, nel, mi = int(s["ntstep"]), int(s["nelements"]), int(s["max_nions"])
        em = np.zeros((ntstep, nnubins, 2 * nel * mi + 1))
        tem = np.zeros_like(em)
        ab = np.zeros((ntstep, nnubins, nel * mi))
        lc = np.zeros((100, ntstep))
        fl = np.zeros((100, ntstep, nnubins))
        sd = np.ascontiguousarray(syn_dir, np.float64)
        self._check(self.lib.artis_gpu_bin_escaped_resolved(float(tmax), int(nprocs), int(nnubins), _pd(sd), _pd(em), _pd(tem), _pd(ab),
                                                            _pd(lc), _pd(fl)), "artis_gpu_bin_escaped_resolved")
        return em, tem, ab, lc, fl

    def download_escaped(self, only_new=False, capacity=None):
        """the escaped packets of the resident ensemble in array order (only_new: not delivered by an earlier only_new call)"""
        cap = self.npkts if capacity is None else capacity
        out = np.zeros(cap, dtype=abi.PACKET_DTYPE)
        n = C.c_int(0)
        self._check(self.lib.artis_gpu_packets_download_escaped(out.ctypes.data_as(C.c_void_p), cap, int(bool(only_new)),
                                                                C.byref(n)), "artis_gpu_packets_download_escaped")
        return out[:n.value]

    def pack_escaped_noalloc(self, dev_packets_ptr, npkts, dev_out_ptr, capacity, dev_n_out_ptr, stream=None):
        """artis_gpu_pack_escaped_noalloc: device pointers in, device pointers out, nothing allocated, no synchronisation"""
        self._check(self.lib.artis_gpu_pack_escaped_noalloc(C.c_void_p(dev_packets_ptr), int(npkts), C.c_void_p(dev_out_ptr),
                                                            int(capacity), C.c_void_p(dev_n_out_ptr), C.c_void_p(stream or 0)),
                    "artis_gpu_pack_escaped_noalloc")

    # ---- deterministic probes (K10)
    def probe_boundary_cross(self, pkts):
        n = pkts.size
        sd, sn, lc = np.zeros(n), np.zeros(n, np.int32), np.zeros(n, np.int32)
        self._check(self.lib.artis_gpu_probe_boundary_cross(pkts.ctypes.data_as(C.c_void_p), n, _pd(sd), _pi(sn), _pi(lc)),
                    "probe_boundary_cross")
        return sd, sn, lc

    def probe_kappa_cont(self, pkts):
        n = pkts.size
        k = np.zeros((n, 5))
        self._check(self.lib.artis_gpu_probe_kappa_cont(pkts.ctypes.data_as(C.c_void_p), n, _pd(k)), "probe_kappa_cont")
        return k

    def probe_closest_transition(self, nu_cmf, next_trans):
        nu = np.ascontiguousarray(nu_cmf, np.float64)
        nt = np.ascontiguousarray(next_trans, np.int32)
        out = np.zeros(nu.size, np.int32)
        self._check(self.lib.artis_gpu_probe_closest_transition(_pd(nu), _pi(nt), nu.size, _pi(out)), "probe_closest")
        return out

    def probe_get_event(self, pkts, tau_rnd, abort_dist):
        n = pkts.size
        t = np.ascontiguousarray(tau_rnd, np.float64)
        a = np.ascontiguousarray(abort_dist, np.float64)
        ed, et, nt, ma = np.zeros(n), np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros((n, 4), np.int32)
        self._check(self.lib.artis_gpu_probe_get_event(pkts.ctypes.data_as(C.c_void_p), n, _pd(t), _pd(a), _pd(ed), _pi(et),
                                                       _pi(nt), _pi(ma)), "probe_get_event")
        return ed, et, nt, ma

    def probe_macroatom_rates(self, mgi, element, ion, level):
        a = [np.ascontiguousarray(x, np.int32) for x in (mgi, element, ion, level)]
        n = a[0].size
        out = np.zeros((n, 9))
        self._check(self.lib.artis_gpu_probe_macroatom_rates(_pi(a[0]), _pi(a[1]), _pi(a[2]), _pi(a[3]), n, self.nts,
                                                             _pd(out)), "probe_macroatom_rates")
        return out

    def probe_cooling_ion(self, mgi, element, ion):
        a = [np.ascontiguousarray(x, np.int32) for x in (mgi, element, ion)]
        n = a[0].size
        out = np.zeros(n)
        self._check(self.lib.artis_gpu_probe_cooling_ion(_pi(a[0]), _pi(a[1]), _pi(a[2]), n, _pd(out)), "probe_cooling_ion")
        return out

    def probe_gamma_kappa(self, pkts):
        n = pkts.size
        out = np.zeros((n, 3))
        self._check(self.lib.artis_gpu_probe_gamma_kappa(pkts.ctypes.data_as(C.c_void_p), n, _pd(out)), "probe_gamma_kappa")
        return out

    def probe_select_continuum_nu(self, element, lowerion, lower, upper, T_e, zrand):
        a = [np.ascontiguousarray(x, np.int32) for x in (element, lowerion, lower, upper)]
        t = np.ascontiguousarray(T_e, np.float32)
        z = np.ascontiguousarray(zrand, np.float64)
        out = np.zeros(z.size)
        self._check(self.lib.artis_gpu_probe_select_continuum_nu(_pi(a[0]), _pi(a[1]), _pi(a[2]), _pi(a[3]),
                                                                 t.ctypes.data_as(C.POINTER(C.c_float)), _pd(z), z.size, _pd(out)),
                    "probe_select_continuum_nu")
        return out

    def probe_levelpop(self, mgi, element, ion, level):
        a = [np.ascontiguousarray(x, np.int32) for x in (mgi, element, ion, level)]
        n = a[0].size
        out = np.zeros(n)
        self._check(self.lib.artis_gpu_probe_levelpop(_pi(a[0]), _pi(a[1]), _pi(a[2]), _pi(a[3]), n, _pd(out)),
                    "probe_levelpop")
        return out


def _pd(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _pi(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))
