/* artis_gpu.h -- C ABI of the B200-native ARTIS packet-transport path.
 *
 * This is the whole drop-in boundary: the entry points below replace exactly
 *     update_packets(my_rank, nts, packets)          (update_packets.h; call site sn3d.cc:574)
 *     mpi_reduce_estimators(my_rank, nts)            (sn3d.cc:316-384; call site sn3d.cc:582)
 *     radfield::reduce_estimators()                  (radfield.cc:1502-1563)
 * of the reference and nothing else. Plain pointers and sizes only; no C++ or torch types.
 * The reference keeps every input/output of update_packets in process globals, so the host glue
 * (oracle/refglue/flatten.h is a working example) flattens those globals into the three
 * table structs declared here and passes them across.
 *
 * All functions return 0 on success and a non-zero code otherwise; artis_gpu_last_error() gives the text.
 * The reference aborts on failure (sn3d.h:50-62): the glue is expected to printout()+abort() on non-zero.
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with ARTIS_ERR_NODEVICE.
 */
#ifndef ARTIS_GPU_H
#define ARTIS_GPU_H

#include <stddef.h>
#include <stdint.h>

#include "artis_gpu_fields.def"

#ifdef __cplusplus
extern "C" {
#endif

#define ARTIS_GPU_ABI_VERSION 6

enum artis_gpu_error {
  ARTIS_OK = 0,
  ARTIS_ERR_NODEVICE = 1,   /* no CUDA device / driver */
  ARTIS_ERR_CUDA = 2,       /* a CUDA runtime call failed */
  ARTIS_ERR_ABI = 3,        /* abi_version mismatch or inconsistent table sizes */
  ARTIS_ERR_STATE = 4,      /* call order violated (e.g. transport before upload_cellstate) */
  ARTIS_ERR_TRANSPORT = 5,  /* a packet hit one of the reference's fatal branches (abort() in the reference) */
  ARTIS_ERR_UNSUPPORTED = 6,/* option combination not implemented yet */
  ARTIS_ERR_NCCL = 7
};

/* ---- A1: struct packet (packet.h:28-73), byte-identical 304-byte AoS layout -------------------------- */
/* packet_type (packet.h:6-17) */
enum {
  ARTIS_TYPE_ESCAPE = 32,
  ARTIS_TYPE_RADIOACTIVE_PELLET = 100,
  ARTIS_TYPE_GAMMA = 10,
  ARTIS_TYPE_RPKT = 11,
  ARTIS_TYPE_KPKT = 12,
  ARTIS_TYPE_MA = 13,
  ARTIS_TYPE_NTLEPTON = 20,
  ARTIS_TYPE_NONTHERMAL_PREDEPOSIT = 21,
  ARTIS_TYPE_PRE_KPKT = 120,
  ARTIS_TYPE_GAMMA_KPKT = 121
};
/* cell_boundary (boundary.h:4-12) */
enum { ARTIS_NEG_X = 101, ARTIS_POS_X = 102, ARTIS_NEG_Y = 103, ARTIS_POS_Y = 104, ARTIS_NEG_Z = 105,
       ARTIS_POS_Z = 106, ARTIS_CROSS_NONE = 107 };
/* decay::decaytypes used on the path (decay.h) */
enum { ARTIS_DECAYTYPE_ALPHA = 0, ARTIS_DECAYTYPE_ELECTRONCAPTURE = 1, ARTIS_DECAYTYPE_BETAPLUS = 2,
       ARTIS_DECAYTYPE_BETAMINUS = 3, ARTIS_DECAYTYPE_NONE = 4 };

typedef struct ArtisPacket {
  int32_t where;
  int32_t type;
  int32_t last_cross;
  int32_t interactions;
  int32_t nscatterings;
  int32_t last_event;
  double pos[3];
  double dir[3];
  double e_cmf;
  double e_rf;
  double nu_cmf;
  double nu_rf;
  int32_t next_trans;
  int32_t emissiontype;
  double em_pos[3];
  int32_t em_time;
  int32_t _pad0;
  double prop_time;
  int32_t absorptiontype;
  int32_t trueemissiontype;
  int32_t trueem_time;
  int32_t _pad1;
  double absorptionfreq;
  double absorptiondir[3];
  double stokes[3];
  double pol_dir[3];
  double tdecay;
  int32_t escape_type;
  int32_t escape_time;
  int32_t scat_count;
  int32_t number;
  uint8_t originated_from_particlenotgamma;
  uint8_t _pad2[3];
  int32_t pellet_decaytype;
  int32_t pellet_nucindex;
  float trueemissionvelocity;
  int32_t ma_element;       /* mastate.element */
  int32_t ma_ion;           /* mastate.ion */
  int32_t ma_level;         /* mastate.level */
  int32_t ma_activatingline;/* mastate.activatingline */
} ArtisPacket;

/* ---- the three table structs, generated from artis_gpu_fields.def ------------------------------------ */
#define ARTIS_DECL_S(type, name) type name;
#define ARTIS_DECL_A(type, name, count) const type *name;
#define ARTIS_DECL_AOUT(type, name, count) type *name;
typedef struct ArtisStaticTables { ARTIS_STATIC_FIELDS(ARTIS_DECL_S, ARTIS_DECL_A) } ArtisStaticTables;
typedef struct ArtisCellState { ARTIS_CELL_FIELDS(ARTIS_DECL_S, ARTIS_DECL_A) } ArtisCellState;
typedef struct ArtisEstimators { ARTIS_ESTIM_FIELDS(ARTIS_DECL_S, ARTIS_DECL_AOUT) } ArtisEstimators;

/* indices into ArtisEstimators.ts_scalars: members of globals::time_step[nts] written by the path
 * (globals.h:20-40; update_packets.cc:53-60,88-118,152-163; nonthermal.cc do_ntlepton) */
enum {
  ARTIS_TS_GAMMA_DEP = 0, ARTIS_TS_POSITRON_DEP = 1, ARTIS_TS_ELECTRON_DEP = 2, ARTIS_TS_ELECTRON_EMISSION = 3,
  ARTIS_TS_ALPHA_DEP = 4, ARTIS_TS_ALPHA_EMISSION = 5, ARTIS_TS_GAMMA_EMISSION = 6, ARTIS_TS_CMF_LUM = 7,
  ARTIS_TS_NT_ENERGY_DEPOSITED = 8, /* nonthermal::nt_energy_deposited */
  /* virtual-packet counts of the timestep (vpkt.cc:69-74), whole numbers kept in the FP64 block so that they ride the
   * same all-reduce: nvpkt, nvpkt_esc1 (electron scattering), nvpkt_esc2 (k-packet), nvpkt_esc3 (macro-atom) */
  ARTIS_TS_NVPKT = 9, ARTIS_TS_NVPKT_ESC1 = 10, ARTIS_TS_NVPKT_ESC2 = 11, ARTIS_TS_NVPKT_ESC3 = 12,
  ARTIS_TS_COUNT = 13
};
/* indices into ArtisEstimators.counters: [0,34) = stats::eventcounters (stats.h:48-84); then device counters */
enum {
  ARTIS_CTR_NESC = 33,            /* COUNTER_NESC / globals::nesc */
  ARTIS_CTR_PELLET_DECAYS = 34,   /* time_step[nts].pellet_decays */
  ARTIS_CTR_RPKT_STEPS = 35,      /* iterations of do_rpkt_step (rpkt.cc:623): the roofline unit */
  ARTIS_CTR_LINES_SCANNED = 36,   /* n_scan: line frequencies examined in get_event */
  ARTIS_CTR_LINES_TAU = 37,       /* n_tau : Sobolev optical depths evaluated */
  ARTIS_CTR_BF_EVALUATED = 38,    /* n_bf  : bf continua evaluated in calculate_kappa_bf */
  ARTIS_CTR_ESTIM_ATOMICS = 39,   /* n_est : estimator atomics issued (after warp aggregation) */
  ARTIS_CTR_GAMMA_STEPS = 40,     /* do_gamma calls */
  ARTIS_CTR_MA_JUMPS = 41,        /* macro-atom loop iterations */
  ARTIS_CTR_MA_TRANS_SCANNED = 42,/* transitions scanned while building/sampling macro-atom rates */
  ARTIS_CTR_KPKT_EVENTS = 43,     /* do_kpkt / do_kpkt_bb calls */
  ARTIS_CTR_COOL_SCANNED = 44,    /* cooling terms evaluated */
  ARTIS_CTR_INTERACTIONS = 45,    /* sum of pkt.interactions increments (stats.cc:172-178) */
  ARTIS_CTR_PASSES = 46,          /* bucket+transport passes in the last update_packets */
  ARTIS_CTR_SOFT_REPAIRS = 47,    /* clamps of SURVEY Q6 applied */
  ARTIS_CTR_KAPPA_EVALS = 48,     /* calculate_kappa_rpkt_cont recomputes (cache misses) */
  ARTIS_CTR_PACKET_UPDATES = 49,  /* do_packet calls */
  ARTIS_CTR_HANDOFFS = 50,        /* packets handed from one scheduler ring to another in mid-timestep */
  /* [51,61): unused since ABI 6 (were the per-kind scheduler statistics of five kinds), except: */
  ARTIS_CTR_BF_LISTED = 51,       /* bf continua the REFERENCE's loop visits for the same opacity evaluations (every continuum with
                                   * nu_edge <= nu <= nu_edge * last_phixs_nuovernuedge); BF_EVALUATED counts the live ones */
  ARTIS_CTR_THIN_BEGIN_CYCLES = 52, /* of WARP_CYCLES0 + 1 (r-packet steps in non-grey cells): SM cycles up to the opacity evaluation */
  ARTIS_CTR_THIN_KAPPA_CYCLES = 53, /* ... and inside it (coop_kappa_warm); the rest is get_event, the event and the estimators */
  ARTIS_CTR_VPKT_REQUESTS = 57,   /* VPKT_ON: interactions whose virtual packets were traced from the request list in the last transport */
  ARTIS_CTR_VPKT_RESTARTS = 58,   /* ... times the request list filled and the scheduler was launched again */
  ARTIS_CTR_DRAIN_NS_10PCT = 61,  /* ns from the transport launch until 10% / 1% / none of the packets are still active */
  ARTIS_CTR_DRAIN_NS_1PCT = 62,
  ARTIS_CTR_DRAIN_NS_DONE = 63,
  /* scheduler statistics per event kind k = 0..7 (kernels.cuh: 0 macro-atom events / k-packets, 1 r-packet in a non-grey
   * cell, 2 gamma, 3 pellets, 4 r-packet in a grey or empty cell, 5 macro-atom jumps within an ion) */
  ARTIS_CTR_WARP_ITERS0 = 64,     /* [64,72): scheduler iterations a warp ran in kind k */
  ARTIS_CTR_BUSY_LANES0 = 72,     /* [72,80): lanes holding a packet, summed over those iterations */
  ARTIS_CTR_WARP_CYCLES0 = 80,    /* [80,88): SM cycles (clock64) the warps spent in the event code of kind k */
  ARTIS_CTR_REFILL_CYCLES = 88,   /* ... claiming ring entries and loading packets */
  ARTIS_CTR_PUSH_CYCLES = 89,     /* ... storing packets and pushing them on the ring of their next kind */
  ARTIS_CTR_IDLE_CYCLES = 90,     /* ... waiting with every ring of their roles empty */
  ARTIS_CTR_ROW_FLUSHES = 91,     /* rows-on-demand tables: times the row arena was started afresh in the last transport */
  ARTIS_CTR_ROW_UNITS = 92,       /* ... 128-byte units of rows built in the last transport (all arena generations) */
  ARTIS_CTR_COUNT = 96
};

/* ---- life cycle --------------------------------------------------------------------------------------- */
/* Once, after input()/ratecoefficients_init()/grid_init(). `rank`/`nranks` are the shard ("MPI rank") numbers:
 * the Philox key of every packet is (seed, rank, packet.number) like the per-rank seeding at input.cc:1908. */
int artis_gpu_init(const ArtisStaticTables *tables, int device, int rank, int nranks, unsigned long long seed);
int artis_gpu_finalize(void);
/* Text of the last error on this thread-less (single host thread per shard) library. */
int artis_gpu_last_error(char *buf, size_t buflen);

/* After update_grid + broadcast (sn3d.cc:543-551). Also runs the per-cell device precompute. */
int artis_gpu_upload_cellstate(const ArtisCellState *state);
/* emissivities.cc:138 zero_estimators() equivalent for the device-resident E-list. */
int artis_gpu_zero_estimators(void);

/* ---- the hot path ------------------------------------------------------------------------------------- */
/* Drop-in for update_packets(my_rank, nts, packets) (update_packets.cc:234): H2D of the AoS packet array,
 * all passes of the transport, D2H of the packets. Postconditions as update_packets.cc:309:
 * every packet has type ESCAPE or prop_time == ts+tw. Packets come back in input order.
 * An ensemble that does not fit the device next to the tables is streamed through it in batches (the reference reads and
 * writes a rank's packets in one piece, sn3d.cc:387-398; packets are independent within a timestep and the estimators keep
 * accumulating): the batch size follows from the free device memory, ARTIS_BATCH_PACKETS sets it by hand. */
int artis_gpu_update_packets(int my_rank, int nts, ArtisPacket *host_packets, int npkts);
/* The same in three steps, so that a driver can keep the packets resident between timesteps
 *  and so that the device-resident rate can be measured separately from the copies. */
int artis_gpu_packets_upload(const ArtisPacket *host_packets, int npkts);
int artis_gpu_transport_resident(int nts);
int artis_gpu_packets_download(ArtisPacket *host_packets, int npkts);
/* milliseconds spent in the kernels of the last transport_resident call (CUDA events on the library's stream):
 * [0] total, [1] bucketing, [2] transport kernels, [3] cell_precompute of the last upload_cellstate, [4] the last
 * artis_gpu_bin_escaped kernel, [5] the NCCL group of the last artis_gpu_reduce_estimators / all-reduced bin_escaped;
 * returns number of values written */
int artis_gpu_last_timings(double *ms, int n);
/* Introspection for the measurement harness: kernels launched by the library since init, and bytes held by the
 * per-cell tables that replace the reference's cellhistory cache. */
long long artis_gpu_launch_count(void);
size_t artis_gpu_table_bytes(void);
/* Geometry of the persistent scheduler for the measurement harness: out[0] kernels per timestep (1: one kernel with every
 * role, 2: macro-atom jump kernel + every-other-role kernel), out[1..3] resident 128-thread blocks per SM of the single
 * kernel / the every-other-role kernel / the jump kernel. */
int artis_gpu_scheduler_geometry(int *out, int n);
/* Tunables (0 keeps the default): events a thread may process before it yields its packet for re-bucketing;
 * block_threads -3 / -4: one persistent kernel with every role / the two-kernel scheduler (default). block_threads <= -1000 sets the hang guard of the persistent scheduler instead: the timestep is
 * given up with ARTIS_ERR_TRANSPORT (device code 900) when no packet has finished for (-1000 - block_threads) seconds
 * (default 20; the guard looks at progress of the whole timestep, not at how long one warp has been idle). */
int artis_gpu_set_tuning(int max_events_per_pass, int block_threads);

/* ---- estimators --------------------------------------------------------------------------------------- */
/* Replaces mpi_reduce_estimators + radfield::reduce_estimators: ONE sum all-reduce over the contiguous FP64
 * estimator block and one over the integer block, on the transport stream (NCCL over NVLink). No-op for 1 rank. */
int artis_gpu_comm_unique_id(char id_out[128]);                 /* rank 0 creates the NCCL id */
int artis_gpu_comm_init(const char id[128], int rank, int nranks); /* every rank joins */
int artis_gpu_reduce_estimators(int nts);
int artis_gpu_download_estimators(ArtisEstimators *est);
/* Device address and length (in doubles) of the contiguous FP64 estimator block, for callers that run their own
 * collective on it (e.g. torch.distributed). */
int artis_gpu_estimator_block(void **dev_ptr, size_t *ndoubles);

/* ---- light curves and spectrum of the resident packets ------------------------------------ */
/* Replaces the packet loop and the MPI reductions of write_partial_lightcurve_spectra (spectrum.cc:641-700; call site
 * sn3d.cc:603) for a driver that keeps its packets on the device: every resident packet with type ESCAPE is binned by
 * add_to_lc_res (light_curve.cc:34-54) and, for escape_type RPKT, add_to_spec_res (spectrum.cc:602-615, 339-372) with
 * current_abin = -1 (all angles) and no emission/absorption resolution (WRITE_PARTIAL_EMISSIONABSORPTIONSPEC = false).
 *   tmax, nprocs  globals::tmax, globals::nprocs        nnubins  globals::nnubins (MNUBINS = 1000, exspec.h:5)
 *   allreduce     non-zero: sum over the shards of artis_gpu_comm_init before the download (spectrum.cc:688-699)
 *   lc            out [4][ntstep]: r-packet lum, r-packet lumcmf, gamma lum, gamma lumcmf (the arrays write_light_curve takes)
 *   spec_flux     out [ntstep][nnubins]: spectra->timesteps[nt].flux[nnu]
 *   stokes        out [3][ntstep][nnubins] (stokes_i/q/u of add_to_spec) or NULL
 * The frequency grid is init_spectra(nu_min_r, nu_max_r) (spectrum.cc:483-499), float delta_freq included. Sums are
 * accumulated with FP64 atomics, so a bin agrees with the reference's serial sum to rounding (not bit for bit). */
int artis_gpu_bin_escaped(double tmax, int nprocs, int nnubins, int allreduce, double *lc, double *spec_flux, double *stokes);
/* The resolved outputs of the same packet loop, for exspec-style products (what tests/setup_classicmode.sh switches on with
 * WRITE_PARTIAL_EMISSIONABSORPTIONSPEC) -- every resident packet that escaped as an r-packet:
 *   emission, trueemission  out [ntstep][nnubins][2 * nelements * max_nions + 1]: the angle-averaged spectrum resolved by the
 *                           process of the last / the true emission (add_to_spec with do_emission_res, spectrum.cc:373-386;
 *                           columns of columnindex_from_emissiontype, spectrum.cc:306-337: bound-bound per ion, bound-free per
 *                           ion, free-free)
 *   absorption              out [ntstep][nnubins][nelements * max_nions]: by the ion of the last line absorption, in the
 *                           frequency bin of the absorbed photon (spectrum.cc:410-421)
 *   lc_res                  out [MABINS = 100][ntstep]: light curve of each escape-direction bin (add_to_lc_res with
 *                           current_abin >= 0, light_curve.cc:55-62; get_escapedirectionbin, vectors.h:158-192)
 *   flux_res                out [100][ntstep][nnubins]: spectrum of each escape-direction bin (add_to_spec_res, spectrum.cc:616)
 *   syn_dir                 globals::syn_dir (unit vector to the observer) */
int artis_gpu_bin_escaped_resolved(double tmax, int nprocs, int nnubins, const double *syn_dir, double *emission, double *trueemission,
                                   double *absorption, double *lc_res, double *flux_res);
/* The escaped packets of the resident ensemble, packed in array order into out[0..*n_out): the subset that the host
 * loops over the packet array select with `type == TYPE_ESCAPE` (spectrum.cc:673, exspec.cc, packet.cc:152
 * write_packets), so that a driver with resident packets moves *n_out x 304 bytes instead of npkts x 304.
 * only_new != 0: only packets that no earlier only_new call since the last artis_gpu_packets_upload has delivered (an
 * escaped packet is never processed again, update_packets.cc:287), i.e. this timestep's escapes when called once per
 * timestep. If *n_out > capacity the call fails with ARTIS_ERR_ABI, *n_out holds the count and nothing is marked. */
int artis_gpu_packets_download_escaped(ArtisPacket *out, int capacity, int only_new, int *n_out);

/* Escaped-packet compaction that allocates nothing. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot reach cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). The same selection as artis_gpu_packets_download_escaped(only_new = 0) -- every packet with
 * type == TYPE_ESCAPE, in array order -- but on the CALLER's device arrays:
 *   dev_packets[npkts]  in,  device memory        dev_out[capacity]  out, device memory
 *   dev_n_out           out, device memory: the number of escaped packets (if it exceeds capacity only the first
 *                       `capacity` of them were written)
 *   cuda_stream         a cudaStream_t (NULL: the default stream); the kernels are enqueued there and the call returns
 *                       without synchronising -- results are valid in stream order
 * The per-chunk counts live in a file-scope device array of 64 KiB that every GPU loading the library holds until exit:
 * npkts above 4 194 304 (ARTIS_NOALLOC_MAX_PACKETS) is rejected with ARTIS_ERR_ABI. The call makes no cudaMalloc /
 * cudaFree and calls no allocating library routine; its first use on a device looks up the array's address
 * (cudaGetSymbolAddress), so make that first call outside a stream capture. All calls share the one array: calls that
 * overlap on the same GPU (two streams, two host threads) race on it and must be serialised by the caller. */
#define ARTIS_NOALLOC_MAX_PACKETS (1 << 22)
int artis_gpu_pack_escaped_noalloc(const ArtisPacket *dev_packets, int npkts, ArtisPacket *dev_out, int capacity,
                                   int *dev_n_out, void *cuda_stream);

/* ---- per-cell precompute on the device --------------------------------------------------------- */
/* calculate_cooling_rates (kpkt.cc:84-165, called by update_grid after the temperature solve, update_grid.cc:1205) for
 * every non-empty model cell of the cell state of the last artis_gpu_upload_cellstate, in the reference's order of
 * addition:
 *   totalcooling         out [npts_model + 1]                  grid::modelgrid[mgi].totalcooling
 *   cooling_contrib_ion  out [(npts_model + 1) * includedions] grid::modelgrid[mgi].cooling_contrib_ion[element][ion]
 * (entries of empty cells are 0). install != 0 also makes them the device-resident cell state's values and rebuilds the
 * cooling rows that start from them, for a driver that uploads its cell state without them. The other lazily computed
 * per-cell quantities of the reference (level populations, bound-free departure ratios, corrphotoioncoeff per
 * continuum, macro-atom and cooling rows) are built by artis_gpu_upload_cellstate itself. */
int artis_gpu_cooling_rates(double *totalcooling, double *cooling_contrib_ion, int install);

/* ---- estimator normalisation on the device ------------------------------------------------- */
/* What update_grid does to the raw estimators of a timestep before it uses them (update_grid.cc:1041-1134:
 * estimator_normfactor = 1 / deltaV / deltat / nprocs, radfield::normalise_J / normalise_nuJ with the 1/4pi, the
 * ffheating / colheating scaling), radfield::get_T_J_from_J (radfield.cc:1458-1475, clamped to [MINTEMP, MAXTEMP], T_R of the
 * cell if J is not finite) and normalise_grey (grey_emissivities.cc:124-139), for the estimators resident on the device and
 * the timestep of the last artis_gpu_upload_cellstate:
 *   vol_init_modelcell  in  [npts_model]  grid::vol_init_modelcell(mgi)       nprocs  globals::nprocs      maxtemp  MAXTEMP
 *   out_cells           out [npts_model][6]: J, nuJ, ffheatingestimator, colheatingestimator, rpkt_emiss (normalised), T_J;
 *                       zeros for model cells without propagation cells
 *   gamma_dep_pathint   out time_step[nts].gamma_dep_pathint
 * The device estimators themselves are left raw (the all-reduce and artis_gpu_download_estimators still see them). */
int artis_gpu_normalise_estimators(const double *vol_init_modelcell, int nprocs, double maxtemp, double *out_cells,
                                   double *gamma_dep_pathint);
/* The inverse of artis_gpu_download_estimators for the FP64 arrays and the counters (restart of a driver whose
 * estimators live on the device); NULL members are left as they are on the device. */
int artis_gpu_upload_estimators(const ArtisEstimators *est);

/* ---- packet files from the resident ensemble ------------------------------------------------ */
/* The driver's packet checkpoint and final packet file, written from / read into the resident ensemble a bounded chunk at
 * a time (a driver that keeps its packets on the device between timesteps never needs the whole array on the host):
 *   artis_gpu_packets_write_temp  = write_temp_packetsfile (sn3d.cc:387-398): the raw 304-byte records, byte for byte
 *   artis_gpu_packets_read_temp   = read_temp_packetsfile (packet.cc:198-209): such a file becomes the resident ensemble
 *   artis_gpu_packets_write_text  = write_packets (packet.cc:152-196): packets*.out, same header, columns and conversions
 * `path` is the file name the driver would build ("packets_%.4d_ts%d.tmp", "packets%.2d_%.4d.out"). */
int artis_gpu_packets_write_temp(const char *path);
int artis_gpu_packets_read_temp(const char *path, int npkts);
int artis_gpu_packets_write_text(const char *path);

/* ---- K10: deterministic probes (test entry points for SURVEY 8a A4-A8, A11, A12, A14) ------------------ */
/* Each probe runs the corresponding device function for n independent packet states on the GPU. */
/* boundary_cross (boundary.cc:101): out sdist[n], snext[n], last_cross_out[n] */
int artis_gpu_probe_boundary_cross(const ArtisPacket *pkts, int n, double *sdist, int *snext, int *last_cross_out);
/* calculate_kappa_rpkt_cont (rpkt.cc:1209): out kappa[n*5] = total, es, ff, bf, ffheating */
int artis_gpu_probe_kappa_cont(const ArtisPacket *pkts, int n, double *kappa);
/* closest_transition (rpkt.cc:26): out lineindex[n] */
int artis_gpu_probe_closest_transition(const double *nu_cmf, const int *next_trans, int n, int *lineindex);
/* get_event (rpkt.cc:67) with injected tau_rnd and abort_dist: out edist[n], eventtype[n] (-1 none, 1 bb, 2 cont),
 * next_trans_out[n], mastate_out[n*4] */
int artis_gpu_probe_get_event(const ArtisPacket *pkts, int n, const double *tau_rnd, const double *abort_dist,
                              double *edist, int *eventtype, int *next_trans_out, int *mastate_out);
/* calculate_macroatom_transitionrates (macroatom.cc:57) for (mgi, element, ion, level): out rates[n*9] */
int artis_gpu_probe_macroatom_rates(const int *mgi, const int *element, const int *ion, const int *level, int n,
                                    int nts, double *rates);
/* calculate_kpkt_rates_ion totals (kpkt.cc:167): cumulative cooling of one ion: out contrib_total[n] */
int artis_gpu_probe_cooling_ion(const int *mgi, const int *element, const int *ion, int n, double *contrib_total);
/* gamma opacities (gammapkt.cc:335, photo_electric.cc:10,51): out kap[n*3] = compton, photoelectric, pair */
int artis_gpu_probe_gamma_kappa(const ArtisPacket *pkts, int n, double *kap);
/* select_continuum_nu (ratecoeff.cc:628) with its random number injected (zrand = 1 - gsl_rng_uniform, ratecoeff.cc:643):
 * the free-bound emission frequency for recombination onto (element, lowerion, lower) from level `upperionlevel` of the ion
 * above at electron temperature T_e: out nu[n]. The reference integrates with adaptive GSL qag to
 * CONTINUUM_NU_INTEGRAL_ACCURACY (1e-2 / 1e-3); stated tolerance of the comparison: 1e-3 relative. */
int artis_gpu_probe_select_continuum_nu(const int *element, const int *lowerion, const int *lower, const int *upperionlevel,
                                        const float *T_e, const double *zrand, int n, double *nu);
/* get_levelpop (ltepop.cc:432): out pop[n] */
int artis_gpu_probe_levelpop(const int *mgi, const int *element, const int *ion, const int *level, int n, double *pop);

#ifdef __cplusplus
}
#endif
#endif /* ARTIS_GPU_H */
