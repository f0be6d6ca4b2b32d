// common.cuh -- device context, constants, Philox stream, vector helpers.
//
// Everything under artis_b200/csrc/device is written as `ARTIS_DEV` inline functions over a `const Ctx &`
// (the three C-ABI table structs with their pointers replaced by device addresses, plus derived index tables).
// The kernels in transport.cu include these headers; tests/hostsim compiles the same headers for the host as a
// debugging aid (never shipped, never imported by the package).
#pragma once
#include <math.h>
#include <stdint.h>

#include "artis_gpu.h"

#ifdef __CUDACC__
#define ARTIS_DEV __device__ __forceinline__
#define ARTIS_DEV_NOINLINE __device__ __noinline__
#else
#define ARTIS_DEV inline
#define ARTIS_DEV_NOINLINE inline
#endif

namespace artis {

// ---- packets on the device: hot / cold split --------------------------------------------------------------------------
// The ABI carries the reference's 304-byte `struct packet` (ArtisPacket, packet.h:28-73). On the device a packet is two
// records in two arrays:
//   PktHot  (128 B, one line)  what the transport roles read and write at EVERY visit, ordered so that each role touches
//                              whole 32-byte sectors: sector 0 = identity, scheduling and macro-atom state (all a
//                              macro-atom jump needs), sector 1 = Philox draw counter, e_cmf, nu_cmf, prop_time,
//                              sectors 2-3 = position, direction and the rest-frame pair. An r-packet step loads the
//                              line with eight 16-byte loads (the AoS record took 38 strided 8-byte loads and kept all 76
//                              words live in registers).
//   PktCold (176 B)            bookkeeping that only the rare events touch (emission / absorption / escape records, Stokes
//                              vector, pellet data); events write it in place through Pkt::cold, it never sits in registers.
// artis_gpu_packets_upload / _download transpose between the two forms on the device (k_unpack / k_pack); the
// transposition is lossless down to the padding bytes of the host struct.
struct alignas(128) PktHot {
  int where;
  unsigned type_cross;  // type | last_cross << 16   (two small enums: 16 bits each keep every valid value)
  unsigned ma_el_ion;   // mastate.element | mastate.ion << 16
  int ma_level, ma_activatingline, interactions, last_event, next_trans;
  unsigned rngctr;      // doubles drawn so far in this timestep (a packet draws ~1e5, far below 2^32)
  int number;
  double e_cmf, nu_cmf, prop_time;
  double pos[3], nu_rf;
  double dir[3], e_rf;
};
struct alignas(16) PktCold {
  double em_pos[3];
  int em_time, emissiontype;
  double stokes[3];
  int nscatterings, absorptiontype;
  double pol_dir[3];
  int escape_type, escape_time;
  double absorptiondir[3];
  int trueemissiontype, trueem_time;
  double absorptionfreq, tdecay;
  float trueemissionvelocity;
  int pellet_decaytype, pellet_nucindex;
  unsigned origin_bytes;  // originated_from_particlenotgamma and the three padding bytes behind it, as the host has them
  int scat_count, pad0, pad1, spare;
};
static_assert(sizeof(PktHot) == 128 && sizeof(PktCold) == 176, "device packet records");
// The register-resident working set of one packet: the hot fields by name (the physics reads and writes them like the
// members of struct packet) and the address of its cold record.
struct Pkt {
  int where, type, last_cross, interactions, last_event, next_trans, number;
  int ma_element, ma_ion, ma_level, ma_activatingline;
  double pos[3], dir[3], e_cmf, e_rf, nu_cmf, nu_rf, prop_time;
  PktCold *cold;
};
// A packet migrates between SMs in mid-timestep (the scheduler hands it from ring to ring), and L1 is not coherent across
// SMs: everything a role READS from a packet's records goes through L2 (ld.global.cg) -- ldcold for the cold record, the
// explicit 16-byte loads of transport.cu for the hot one -- so that no fence has to invalidate the SM's L1 (where the
// read-only atomic and per-cell tables live) when a packet arrives. Writes are write-through.
template <class T>
#ifdef __CUDACC__
__host__ __device__
#endif
    inline T
    ldcold(const T *addr) {
#ifdef __CUDA_ARCH__
  return __ldcg(addr);
#else
  return *addr;
#endif
}
#ifdef __CUDACC__
#define ARTIS_HD __host__ __device__ inline
#else
#define ARTIS_HD inline
#endif
ARTIS_HD unsigned pkt_pack2(int lo, int hi) { return ((unsigned)lo & 0xffffu) | ((unsigned)hi << 16); }
ARTIS_HD int pkt_lo16(unsigned w) { return (int)(short)(w & 0xffffu); }
ARTIS_HD int pkt_hi16(unsigned w) { return (int)(short)(w >> 16); }
// ABI record -> device records and back (k_unpack / k_pack, the probes, the host build)
ARTIS_HD void pkt_split(const ArtisPacket &a, PktHot &h, PktCold &c) {
  h.where = a.where;
  h.type_cross = pkt_pack2(a.type, a.last_cross);
  h.ma_el_ion = pkt_pack2(a.ma_element, a.ma_ion);
  h.ma_level = a.ma_level;
  h.ma_activatingline = a.ma_activatingline;
  h.interactions = a.interactions;
  h.last_event = a.last_event;
  h.next_trans = a.next_trans;
  h.rngctr = 0u;
  h.number = a.number;
  h.e_cmf = a.e_cmf;
  h.nu_cmf = a.nu_cmf;
  h.prop_time = a.prop_time;
  h.nu_rf = a.nu_rf;
  h.e_rf = a.e_rf;
  for (int d = 0; d < 3; d++) {
    h.pos[d] = a.pos[d];
    h.dir[d] = a.dir[d];
    c.em_pos[d] = a.em_pos[d];
    c.stokes[d] = a.stokes[d];
    c.pol_dir[d] = a.pol_dir[d];
    c.absorptiondir[d] = a.absorptiondir[d];
  }
  c.em_time = a.em_time;
  c.emissiontype = a.emissiontype;
  c.nscatterings = a.nscatterings;
  c.absorptiontype = a.absorptiontype;
  c.escape_type = a.escape_type;
  c.escape_time = a.escape_time;
  c.trueemissiontype = a.trueemissiontype;
  c.trueem_time = a.trueem_time;
  c.absorptionfreq = a.absorptionfreq;
  c.tdecay = a.tdecay;
  c.trueemissionvelocity = a.trueemissionvelocity;
  c.pellet_decaytype = a.pellet_decaytype;
  c.pellet_nucindex = a.pellet_nucindex;
  c.origin_bytes = (unsigned)a.originated_from_particlenotgamma | ((unsigned)a._pad2[0] << 8) | ((unsigned)a._pad2[1] << 16) |
                   ((unsigned)a._pad2[2] << 24);
  c.scat_count = a.scat_count;
  c.pad0 = a._pad0;
  c.pad1 = a._pad1;
  c.spare = 0;
}
ARTIS_HD void pkt_join(const PktHot &h, const PktCold &c, ArtisPacket &a) {
  a.where = h.where;
  a.type = pkt_lo16(h.type_cross);
  a.last_cross = pkt_hi16(h.type_cross);
  a.ma_element = pkt_lo16(h.ma_el_ion);
  a.ma_ion = pkt_hi16(h.ma_el_ion);
  a.ma_level = h.ma_level;
  a.ma_activatingline = h.ma_activatingline;
  a.interactions = h.interactions;
  a.last_event = h.last_event;
  a.next_trans = h.next_trans;
  a.number = h.number;
  a.e_cmf = h.e_cmf;
  a.nu_cmf = h.nu_cmf;
  a.prop_time = h.prop_time;
  a.nu_rf = h.nu_rf;
  a.e_rf = h.e_rf;
  for (int d = 0; d < 3; d++) {
    a.pos[d] = h.pos[d];
    a.dir[d] = h.dir[d];
    a.em_pos[d] = c.em_pos[d];
    a.stokes[d] = c.stokes[d];
    a.pol_dir[d] = c.pol_dir[d];
    a.absorptiondir[d] = c.absorptiondir[d];
  }
  a.em_time = c.em_time;
  a.emissiontype = c.emissiontype;
  a.nscatterings = c.nscatterings;
  a.absorptiontype = c.absorptiontype;
  a.escape_type = c.escape_type;
  a.escape_time = c.escape_time;
  a.trueemissiontype = c.trueemissiontype;
  a.trueem_time = c.trueem_time;
  a.absorptionfreq = c.absorptionfreq;
  a.tdecay = c.tdecay;
  a.trueemissionvelocity = c.trueemissionvelocity;
  a.pellet_decaytype = c.pellet_decaytype;
  a.pellet_nucindex = c.pellet_nucindex;
  a.originated_from_particlenotgamma = (uint8_t)(c.origin_bytes & 0xffu);
  a._pad2[0] = (uint8_t)((c.origin_bytes >> 8) & 0xffu);
  a._pad2[1] = (uint8_t)((c.origin_bytes >> 16) & 0xffu);
  a._pad2[2] = (uint8_t)(c.origin_bytes >> 24);
  a.scat_count = c.scat_count;
  a._pad0 = c.pad0;
  a._pad1 = c.pad1;
}
// working set <-> hot record. The Philox draw counter is handed over separately (it lives in the Rng while a lane holds the
// packet).
ARTIS_HD void pkt_from_hot(Pkt &p, const PktHot &h, PktCold *cold) {
  p.where = h.where;
  p.type = pkt_lo16(h.type_cross);
  p.last_cross = pkt_hi16(h.type_cross);
  p.ma_element = pkt_lo16(h.ma_el_ion);
  p.ma_ion = pkt_hi16(h.ma_el_ion);
  p.ma_level = h.ma_level;
  p.ma_activatingline = h.ma_activatingline;
  p.interactions = h.interactions;
  p.last_event = h.last_event;
  p.next_trans = h.next_trans;
  p.number = h.number;
  p.e_cmf = h.e_cmf;
  p.nu_cmf = h.nu_cmf;
  p.prop_time = h.prop_time;
  p.nu_rf = h.nu_rf;
  p.e_rf = h.e_rf;
  for (int d = 0; d < 3; d++) {
    p.pos[d] = h.pos[d];
    p.dir[d] = h.dir[d];
  }
  p.cold = cold;
}
ARTIS_HD void pkt_to_hot(PktHot &h, const Pkt &p, unsigned rngctr) {
  h.where = p.where;
  h.type_cross = pkt_pack2(p.type, p.last_cross);
  h.ma_el_ion = pkt_pack2(p.ma_element, p.ma_ion);
  h.ma_level = p.ma_level;
  h.ma_activatingline = p.ma_activatingline;
  h.interactions = p.interactions;
  h.last_event = p.last_event;
  h.next_trans = p.next_trans;
  h.rngctr = rngctr;
  h.number = p.number;
  h.e_cmf = p.e_cmf;
  h.nu_cmf = p.nu_cmf;
  h.prop_time = p.prop_time;
  h.nu_rf = p.nu_rf;
  h.e_rf = p.e_rf;
  for (int d = 0; d < 3; d++) {
    h.pos[d] = p.pos[d];
    h.dir[d] = p.dir[d];
  }
}

// ---- physical constants, values exactly as the reference's constants.h:3-37 ------------------------------
constexpr double CLIGHT = 2.99792458e+10;
constexpr double CLIGHT_PROP = CLIGHT;
constexpr double H = 6.6260755e-27;
constexpr double MH = 1.67352e-24;
constexpr double ME = 9.1093897e-28;
constexpr double PI = 3.1415926535987;  // sic (constants.h:12)
constexpr double EV = 1.6021772e-12;
constexpr double MEV = 1.6021772e-6;
constexpr double SIGMA_T = 6.6524e-25;
constexpr double THOMSON_LIMIT = 1e-2;
constexpr double KB = 1.38064852e-16;
constexpr double SAHACONST = 2.0706659e-16;
constexpr double CLIGHTSQUARED = 8.9875518e+20;
constexpr double TWOHOVERCLIGHTSQUARED = 1.4745007e-47;
constexpr double CLIGHTSQUAREDOVERTWOH = 6.7819570e+46;
constexpr double ONEOVERH = 1.509188961e+26;
constexpr double HOVERKB = 4.799243681748932e-11;
constexpr double HCLIGHTOVERFOURPI = 1.580764662876770e-17;
constexpr double H_ionpot = 13.5979996 * EV;
constexpr double FOURPI = 1.256637061600000e+01;
constexpr double C_0 = 5.465e-11;
constexpr double DBL_MAX_ = 1.7976931348623157e+308;

constexpr int COOLINGTYPE_FF = 880, COOLINGTYPE_FB = 881, COOLINGTYPE_COLLEXC = 882, COOLINGTYPE_COLLION = 883;
enum { MA_RADDEEXC = 0, MA_COLDEEXC, MA_RADRECOMB, MA_COLRECOMB, MA_INTERNALDOWNSAME, MA_INTERNALDOWNLOWER,
       MA_INTERNALUPSAME, MA_INTERNALUPHIGHER, MA_INTERNALUPHIGHERNT, MA_ACTION_COUNT };
enum { RPKT_EVENTTYPE_NONE = -1, RPKT_EVENTTYPE_BB = 550, RPKT_EVENTTYPE_CONT = 551 };

// ---- device context -----------------------------------------------------------------------------------------
// Per-cell tables filled by the cell_precompute kernels after every upload_cellstate (celltables.cuh). They replace
// the reference's per-thread `cellhistory` (globals.h:174-208): one row per non-empty model cell (index nonemptymgi)
// instead of one cached cell per host thread. Prefix sums are accumulated in the reference's own summation order, so
// a binary search over a row selects exactly what the reference's linear scans select.
// Static data of one bound-free continuum as the opacity sum reads it: one 32-byte record instead of five arrays.
// inv_nu_edge = RN(1/nu_edge) (an IEEE division on the host): with it nu/nu_edge is evaluated as
//   q = nu * inv; r = fma(-nu_edge, q, nu); q' = fma(r, inv, q)
// which is the correctly rounded quotient (Markstein's theorem: the reciprocal is correctly rounded and nu_edge's
// significand is not all ones -- checked at init, the plain division is used otherwise), i.e. the same bits as
// nu / nu_edge at a fifth of the instructions (no reciprocal seed, no Newton iterations).
struct alignas(16) ContRec {
  double nu_edge;
  double inv_nu_edge;   // 0 when the fast quotient is not provably exact for this divisor
  double probability;
  int tablerow;
  int ground_slot;      // index in the ground-continuum list if this is a ground-level continuum, else -1
};
// a / b given inv = RN(1/b), or the plain division when inv == 0
ARTIS_DEV double div_by_known(const double a, const double b, const double inv) {
  if (inv != 0.) {
    const double q = a * inv;
    const double r = fma(-b, q, a);
    return fma(r, inv, q);
  }
  return a / b;
}

struct CellTables {
  int ncells;                // nonempty_npts_model
  double *levelpop;          // [ncells][nlevels_total]        calculate_levelpop (ltepop.cc:417)
  // [ncells][nbfcontinua][2]: level pop of allcont[i] (-1 if the continuum is skipped) and its departure ratio
  // (ch_allcont_departureratios, rpkt.cc:1136-1145), interleaved so that one 16-byte load brings both
  double *bf_cell;
  // The LIVE continua of every cell, in list order (celltables.cuh bf_is_live): those whose term of the bound-free sum can be
  // non-zero at some frequency. bf_live_off[tc] .. bf_live_off[tc + 1] index the cell's entries of bf_live_pair (the bf_cell
  // pair) and bf_live_rec (a copy of the continuum's static record). nullptr: the sums run over the full list.
  const long long *bf_live_off;  // [ncells + 1]
  const double *bf_live_pair;    // [.][2]
  const ContRec *bf_live_rec;    // [.]
  const int *bf_live_rank;       // [ncells][nbfcontinua + 1]: live continua of the cell before list entry i
  // [ncells]: sum over the ions of Z^2 n_ion, the cell's factor of the free-free opacity (rpkt.cc:1040-1062): what every
  // evaluation of kappa_ff used to re-add from 30 ion populations
  const double *ff_zsum;
  double *ma_rates;          // [ncells][nlevels_total][9]     chlevels.processrates (macroatom.cc:57-159)
  double *ma_cum_raddeexc;   // [ncells][ndowntrans_total]     running sum of individ_rad_deexc
  double *ma_cum_intdown;    // [ncells][ndowntrans_total]     running sum of individ_internal_down_same
  double *ma_cum_intup;      // [ncells][nuptrans_total]       running sum of individ_internal_up_same
  double *cool_cum;          // [ncells][ncoolingterms]        cellhistory.cooling_contrib (kpkt.cc:167-308)
  double *collexc_cum;       // [ncells][nuptrans_total]       running `contrib` inside a COLLEXC term (kpkt.cc:706-724)
  // [ncells][nbfcontinua], NO_LUT_PHOTOION builds only: get_corrphotoioncoeff (ratecoeff.cc:1247) of allcont[i], i.e.
  // the bf-rate estimator of the previous timestep or the integral over the radiation field model
  double *corrphoto;
  // ---- rows on demand (large 3-D models: the five macro-atom / cooling arrays above would be ~1.5 MB per cell) ----
  // row_ptr != nullptr: ma_rates, ma_cum_*, cool_cum and collexc_cum are nullptr; the row of a (cell, level) or of a
  // (cell, ion) is built by the first event that needs it (celltables.cuh row_get: the analogue of cellhistory,
  // macroatom.cc:161-172, kpkt.cc:167-308) in a bounded arena, 128 bytes at a time so that no two rows share a cache line.
  unsigned *row_ptr;              // [ncells][nlevels_total + includedions]: 0 = not built, else 1 + offset in 128-byte units
  double *row_arena;              // [row_cap * 16]
  unsigned long long *row_top;    // units of 128 bytes handed out so far (device counter)
  unsigned long long row_cap;     // units the arena holds
  unsigned long long row_soft;    // above this fill the scheduler drains and the host starts the arena afresh
  int *row_stop;                  // set when row_top has passed row_soft
};

// VPKT_ON, deferred ray trace (vpkt.cuh): what an event leaves behind for its virtual packets -- the packet as
// it was at the event, 128 bytes. The persistent scheduler appends these to a list; k_vpkt_trace traces them afterwards.
struct alignas(16) VpktReq {
  int where, last_cross, next_trans, realtype;  // realtype 1 electron scattering, 2 k-packet -> r-packet, 3 macro-atom
  double pos[3], dir[3];                        // dir: the direction the packet ARRIVED with (only realtype 1 reads it)
  double e_cmf, nu_cmf, prop_time, stokes_q, stokes_u;
  double pad_[3];
};
static_assert(sizeof(VpktReq) == 128, "VpktReq is one 128-byte line");

struct Ctx {
  ArtisStaticTables st;   // pointers are device addresses
  ArtisCellState cs;      // 〃
  ArtisEstimators es;     // 〃 (non-const: accumulated with atomics)
  // derived at init on the host (transport.cu: build_derived)
  const int *line_upper_g;   // global level index of the line's upper level   [nlines]
  const int *line_lower_g;   // 〃 lower level
  const int *line_uion;      // unique ion index of the line                   [nlines]
  const int *cont_level_g;   // global level index of allcont[i] lower level   [nbfcontinua]
  const int *cont_upper_g;   // global level index of allcont[i] upper-ion level
  const int *cont_uion;      // unique ion index of allcont[i] lower ion
  const int *gcont_uion;     // unique ion index of groundcont[i]
  const ContRec *cont_rec;   // packed static data of a bf continuum          [nbfcontinua]
  double inv_nphixsnuincrement;  // RN(1 / NPHIXSNUINCREMENT), see ContRec
  const int *downtrans_target_g;  // global level index a downward transition ends in   [ndowntrans_total]
  const int *uptrans_target_g;    // global level index an upward transition ends in    [nuptrans_total]
  const int *lev_uion;       // unique ion index of a global level index      [nlevels_total]
  const int *uion_element;   // element index of a unique ion index           [includedions]
  const int *mgi_of_tabcell; // model grid index of a CellTables row           [tb.ncells]
  const int *allcont_of_lut; // allcont index of LUT continuum index (-1 - cont_index + phixstargetindex) [nbfcontinua]
  CellTables tb;
  unsigned long long seed;
  int rank;
  int nts;                   // timestep being transported
  double t2;                 // ts + tw of that timestep
  double ts_width;
  double t_mid;
  int *error_flag;           // set non-zero by a packet that hits a reference abort() branch
  // VPKT_ON with the barrier-free scheduler: the request list (nullptr: rays are traced where their events happen)
  VpktReq *vreq;
  unsigned long long *vreq_n;   // requests appended so far
  int *vreq_stop;               // set when the list has passed vreq_soft: the scheduler drains, the list is traced, and the
  unsigned long long vreq_cap;  // unfinished packets are launched again (engine.inc)
  unsigned long long vreq_soft;
};

// Build-mode switches as the device code tests them. They are run-time flags of the static tables (one library serves
// every artisoptions_*.h); the ARTIS_NO_* macros compile a family of them out for same-box A/B measurements of what the
// extra code paths cost the classic-mode kernels (profiles/r1_notes.md).
#ifdef ARTIS_CLASSIC_ONLY
#define ARTIS_NO_GRID1D
#define ARTIS_NO_NEBULAR
#endif
#ifdef ARTIS_NO_VPKT
#define ARTIS_OPT_VPKT(c) false
#else
#define ARTIS_OPT_VPKT(c) ((c).st.opt_vpkt_on != 0)
#endif
#ifdef ARTIS_NO_GRID1D
#define ARTIS_OPT_GRID1D(c) false
#else
#define ARTIS_OPT_GRID1D(c) ((c).st.opt_grid_type == 1)
#endif
#ifdef ARTIS_NO_NEBULAR
#define ARTIS_OPT_DETAILED_BF(c) false
#define ARTIS_OPT_MULTIBIN(c) false
#define ARTIS_OPT_NT_SF(c) false
#else
#define ARTIS_OPT_DETAILED_BF(c) ((c).st.opt_detailed_bf_estimators != 0)
#define ARTIS_OPT_MULTIBIN(c) ((c).st.opt_multibin_radfield != 0)
#define ARTIS_OPT_NT_SF(c) ((c).st.opt_nt_on != 0 && (c).st.opt_nt_solve_spencerfano != 0)
#endif

// ARTIS_CLASSIC_ONLY: every switch of artisoptions_classic.h a compile-time constant -- the experiment behind "per-mode
// libraries" (profiles/r2_notes.md): how much of the kernels is code of other build modes?
#ifdef ARTIS_CLASSIC_ONLY
#define ARTIS_OPT_NT(c) false
#define ARTIS_OPT_NLTE(c) false
#define ARTIS_OPT_NOLUT_PHOTOION(c) false
#define ARTIS_OPT_NOLUT_BFHEATING(c) false
#define ARTIS_OPT_RELDOPPLER(c) false
#define ARTIS_OPT_INSTANT_DEPOSITION(c) true
#else
#define ARTIS_OPT_NT(c) ((c).st.opt_nt_on != 0)
#define ARTIS_OPT_NLTE(c) ((c).st.opt_nlte_pops_on != 0)
#define ARTIS_OPT_NOLUT_PHOTOION(c) ((c).st.opt_no_lut_photoion != 0)
#define ARTIS_OPT_NOLUT_BFHEATING(c) ((c).st.opt_no_lut_bfheating != 0)
#define ARTIS_OPT_RELDOPPLER(c) ((c).st.opt_relativistic_doppler != 0)
#define ARTIS_OPT_INSTANT_DEPOSITION(c) ((c).st.opt_instant_particle_deposition != 0)
#endif

// ---- Philox4x32-10 counter-based stream (Salmon et al. 2011), one stream per packet and timestep -------------
// key = (seed_lo, seed_hi); counter = (draw block, packet.number, nts, rank): the seed and the packet number sit in
// different words, so two seeds never share a set of streams (with the number XORed into the key, seeds that differ in
// low bits gave the same streams permuted among the packets). The draw index restarts at 0 every timestep and is kept in
// a device array between the passes of one timestep, so nothing has to survive in the host's struct packet. Interval
// conventions of SURVEY Q17. One Philox call yields two doubles.
struct Rng {
  uint32_t k0, k1, c1, c2, c3;
  uint64_t ctr;      // number of doubles drawn so far
  double spare;      // second double of the last block
};
ARTIS_DEV uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}
// Out of line on the device (ARTIS_DEV_NOINLINE; arguments and result in registers): the generator is called from ~40
// places of the transport kernels, and the kernels' code footprint -- not their arithmetic -- is what the schedulers wait
// for (profiles/r2_notes.md: no-instruction stalls 8 cycles per issued instruction with 1.6 MB of SASS).
struct Philox4 {
  uint32_t x, y, z, w;
};
ARTIS_DEV_NOINLINE Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; r++) {
    const uint32_t hi0 = mulhi32(M0, c0), lo0 = M0 * c0;
    const uint32_t hi1 = mulhi32(M1, c2), lo1 = M1 * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0;
    c1 = n1;
    c2 = n2;
    c3 = n3;
    k0 += W0;
    k1 += W1;
  }
  Philox4 o;
  o.x = c0;
  o.y = c1;
  o.z = c2;
  o.w = c3;
  return o;
}
ARTIS_DEV Rng rng_init(const Ctx &c, const int number, const uint64_t ctr) {
  Rng r;
  r.k0 = (uint32_t)(c.seed & 0xffffffffu);
  r.k1 = (uint32_t)(c.seed >> 32);
  r.c1 = (uint32_t)number;
  r.c2 = (uint32_t)c.nts;
  r.c3 = (uint32_t)c.rank;
  r.ctr = ctr;
  r.spare = -1.;
  return r;
}
ARTIS_DEV double bits_to_unit(uint32_t hi, uint32_t lo) {
  const uint64_t bits = (((uint64_t)hi << 32) | (uint64_t)lo) >> 11;
  return (double)bits * (1.0 / 9007199254740992.0);
}
// [0,1) like gsl_rng_uniform. Draw 2n and 2n+1 come from Philox block n.
ARTIS_DEV double rng_uniform(Rng &r) {
  const uint64_t blk = r.ctr >> 1;
  const bool odd = (r.ctr & 1) != 0;
  r.ctr++;
  if (odd && r.spare >= 0.) {
    const double x = r.spare;
    r.spare = -1.;
    return x;
  }
  // a packet draws far fewer than 2^33 numbers per timestep; the high word of the block index is folded into the
  // timestep word so that even such a stream would stay unique
  const Philox4 o = philox4x32_10((uint32_t)(blk & 0xffffffffu), r.c1, r.c2 ^ ((uint32_t)(blk >> 32) << 16), r.c3, r.k0, r.k1);
  if (odd) return bits_to_unit(o.z, o.w);  // stream resumed on an odd draw (spare was lost between passes)
  r.spare = bits_to_unit(o.z, o.w);
  return bits_to_unit(o.x, o.y);
}
// (0,1) like gsl_rng_uniform_pos
ARTIS_DEV double rng_uniform_pos(Rng &r) {
  double x;
  do {
    x = rng_uniform(r);
  } while (x == 0.);
  return x;
}

// ---- vectors.h:15-144 -------------------------------------------------------------------------------------------
ARTIS_DEV double vec_len(const double x[3]) { return sqrt((x[0] * x[0]) + (x[1] * x[1]) + (x[2] * x[2])); }
ARTIS_DEV double dot3(const double x[3], const double y[3]) { return (x[0] * y[0]) + (x[1] * y[1]) + (x[2] * y[2]); }
ARTIS_DEV void vec_norm(const double in[3], double out[3]) {
  const double m = vec_len(in);
  out[0] = in[0] / m;
  out[1] = in[1] / m;
  out[2] = in[2] / m;
}
ARTIS_DEV void get_velocity(const double x[3], double y[3], const double t) {
  y[0] = x[0] / t;
  y[1] = x[1] / t;
  y[2] = x[2] / t;
}
ARTIS_DEV void cross_prod(const double a[3], const double b[3], double o[3]) {
  o[0] = (a[1] * b[2]) - (b[1] * a[2]);
  o[1] = (a[2] * b[0]) - (b[2] * a[0]);
  o[2] = (a[0] * b[1]) - (b[0] * a[1]);
}
ARTIS_DEV void vec_scale(double v[3], const double f) {
  v[0] *= f;
  v[1] *= f;
  v[2] *= f;
}
// aberration of angles (vectors.h:63-79)
ARTIS_DEV void angle_ab(const double dir1[3], const double vel[3], double dir2[3]) {
  const double vsqr = dot3(vel, vel) / CLIGHTSQUARED;
  const double gamma_rel = 1. / sqrt(1 - vsqr);
  const double ndotv = dot3(dir1, vel);
  const double fact1 = gamma_rel * (1 - (ndotv / CLIGHT));
  const double fact2 = (gamma_rel - (gamma_rel * gamma_rel * ndotv / (gamma_rel + 1) / CLIGHT)) / CLIGHT;
  for (int d = 0; d < 3; d++) dir2[d] = (dir1[d] - (vel[d] * fact2)) / fact1;
}
// vectors.h:81-106
ARTIS_DEV double doppler_nucmf_on_nurf(const Ctx &c, const double dir_rf[3], const double vel_rf[3]) {
  const double ndotv = dot3(dir_rf, vel_rf);
  double dopplerfactor = 1. - (ndotv / CLIGHT);
  if (ARTIS_OPT_RELDOPPLER(c)) {
    const double betasq = dot3(vel_rf, vel_rf) / CLIGHTSQUARED;
    dopplerfactor = dopplerfactor / sqrt(1 - betasq);
  }
  return dopplerfactor;
}
ARTIS_DEV double doppler_packet(const Ctx &c, const double pos[3], const double dir[3], const double prop_time) {
  double v[3];
  get_velocity(pos, v, prop_time);
  return doppler_nucmf_on_nurf(c, dir, v);
}
template <class P>
ARTIS_DEV double doppler_packet(const Ctx &c, const P &p) { return doppler_packet(c, p.pos, p.dir, p.prop_time); }
// the subset of struct packet that a dummy ("ray") packet needs in get_event (rpkt.cc:83-90)
struct Ray {
  double pos[3], dir[3], prop_time, nu_cmf, nu_rf, e_cmf, e_rf;
  int next_trans;
};
template <class P>
ARTIS_DEV Ray ray_of(const P &p) {
  Ray r;
  for (int d = 0; d < 3; d++) {
    r.pos[d] = p.pos[d];
    r.dir[d] = p.dir[d];
  }
  r.prop_time = p.prop_time;
  r.nu_cmf = p.nu_cmf;
  r.nu_rf = p.nu_rf;
  r.e_cmf = p.e_cmf;
  r.e_rf = p.e_rf;
  r.next_trans = p.next_trans;
  return r;
}
// vectors.h:115-144
template <class P>
ARTIS_DEV void move_pkt(const Ctx &c, P &p, const double distance) {
  p.pos[0] += (p.dir[0] * distance);
  p.pos[1] += (p.dir[1] * distance);
  p.pos[2] += (p.dir[2] * distance);
  const double dopplerfactor = doppler_packet(c, p);
  p.nu_cmf = p.nu_rf * dopplerfactor;
  p.e_cmf = p.e_rf * dopplerfactor;
}
template <class P>
ARTIS_DEV void move_pkt_withtime(const Ctx &c, P &p, const double distance) {
  const double nu_cmf_old = p.nu_cmf;
  p.prop_time += distance / CLIGHT_PROP;
  move_pkt(c, p, distance);
  if (p.nu_cmf > nu_cmf_old) p.nu_cmf = nu_cmf_old;
}
// vectors.cc:46-62
ARTIS_DEV void get_rand_isotropic_unitvec(Rng &r, double v[3]) {
  const double zrand = rng_uniform(r);
  const double zrand2 = rng_uniform(r);
  const double mu = -1 + (2. * zrand);
  const double phi = zrand2 * 2 * PI;
  const double sintheta = sqrt(1. - (mu * mu));
  v[0] = sintheta * cos(phi);
  v[1] = sintheta * sin(phi);
  v[2] = mu;
}
// vectors.cc:10-44
ARTIS_DEV void scatter_dir(Rng &r, const double dir_in[3], const double cos_theta, double dir_out[3]) {
  const double zrand = rng_uniform(r);
  const double phi = zrand * 2 * PI;
  const double sin_theta_sq = 1. - (cos_theta * cos_theta);
  const double sin_theta = sqrt(sin_theta_sq);
  const double zprime = cos_theta;
  const double xprime = sin_theta * cos(phi);
  const double yprime = sin_theta * sin(phi);
  const double norm1 = 1. / sqrt((dir_in[0] * dir_in[0]) + (dir_in[1] * dir_in[1]));
  const double norm2 = 1. / sqrt((dir_in[0] * dir_in[0]) + (dir_in[1] * dir_in[1]) + (dir_in[2] * dir_in[2]));
  const double r11 = dir_in[1] * norm1, r12 = -1 * dir_in[0] * norm1, r13 = 0.0;
  const double r21 = dir_in[0] * dir_in[2] * norm1 * norm2, r22 = dir_in[1] * dir_in[2] * norm1 * norm2;
  const double r23 = -1 * norm2 / norm1;
  const double r31 = dir_in[0] * norm2, r32 = dir_in[1] * norm2, r33 = dir_in[2] * norm2;
  dir_out[0] = (r11 * xprime) + (r21 * yprime) + (r31 * zprime);
  dir_out[1] = (r12 * xprime) + (r22 * yprime) + (r32 * zprime);
  dir_out[2] = (r13 * xprime) + (r23 * yprime) + (r33 * zprime);
}

// ---- estimator / counter accumulation ------------------------------------------------------------------------
#ifdef __CUDACC__
// FP64 estimator sums (safeadd, sn3d.h:104-127). With a 1D model every packet of the timestep adds to the same ~100
// cells' J / nuJ / heating sums, and the lanes of a warp are mostly in the same cell (packets are bucketed by cell):
// the lanes that are about to add to the SAME address are found with __match_any_sync, their values are summed with
// shuffles and the lowest lane issues one red.global.add.f64 for all of them. The lanes taking part are exactly those
// converged at the call (__activemask), so this is safe inside divergent event code. Returns the number of global
// reductions issued by this lane (1 for a group leader, else 0): ARTIS_CTR_ESTIM_ATOMICS counts what reaches L2.
ARTIS_DEV_NOINLINE int est_add(double *addr, double v) {  // one copy in a kernel: ~40 call sites, a shuffle loop each
  const unsigned act = __activemask();
  const unsigned peers = __match_any_sync(act, (unsigned long long)addr);
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(peers) - 1;
  double s = v;
  for (unsigned rest = peers & (peers - 1u); rest != 0u; rest &= rest - 1u)  // every peer walks the same list
    s += __shfl_sync(peers, v, __ffs(rest) - 1);
  if (lane == leader) atomicAdd(addr, s);
  return lane == leader ? 1 : 0;
}
ARTIS_DEV void ctr_add(long long *addr, long long v) { atomicAdd((unsigned long long *)addr, (unsigned long long)v); }
ARTIS_DEV void int_add(int *addr, int v) { atomicAdd(addr, v); }
ARTIS_DEV int int_fetch_add(int *addr, int v) { return atomicAdd(addr, v); }
#else
ARTIS_DEV int est_add(double *addr, double v) {
  *addr += v;
  return 1;
}
ARTIS_DEV void ctr_add(long long *addr, long long v) { *addr += v; }
ARTIS_DEV void int_add(int *addr, int v) { *addr += v; }
ARTIS_DEV int int_fetch_add(int *addr, int v) {
  const int old = *addr;
  *addr += v;
  return old;
}
#endif

// per-thread work counters, flushed once per thread at the end of a pass (keeps atomics off the hot loops)
struct Tally {
  int rpkt_steps, lines_scanned, lines_tau, bf_evaluated, estim_atomics, gamma_steps, ma_jumps, ma_trans_scanned,
      kpkt_events, cool_scanned, interactions, soft_repairs, kappa_evals, packet_updates, handoffs;
  int bf_listed;  // continua the reference's loop visits for the same sums (rpkt.cc:1095-1207): bf_evaluated counts the live ones
#ifndef __CUDACC__
  int stats[34];  // stats::eventcounters increments of this thread (stats.h:48-84); on the device they live in shared memory
#endif
  // VPKT_ON: a virtual-packet request left by the event that has just been processed (vpkt.cuh vpkt_flush_request runs it
  // at the one place per scheduler where events return): realtype 1-3 (0 = none) and, for an electron scattering, the
  // direction and Stokes Q, U the packet arrived with
  int vpkt_req;
  double vpkt_dir[3], vpkt_q, vpkt_u;
};
// stats::increment (stats.cc:151). The 34 event counters are bumped at most once or twice per event: on the device one
// copy per thread block in shared memory (an ATOMS each) instead of 34 registers per thread; a kernel that runs events
// clears the copy at its start (stats_block_begin) and adds it to the global counters at its end (stats_block_flush).
#ifdef __CUDACC__
ARTIS_DEV int *stats_block() {
  __shared__ int artis_stats_smem[34];
  return artis_stats_smem;
}
ARTIS_DEV void stat_inc(Tally &, int counter) { atomicAdd(stats_block() + counter, 1); }
ARTIS_DEV void stats_block_begin() {
  for (int i = threadIdx.x; i < 34; i += blockDim.x) stats_block()[i] = 0;
  __syncthreads();
}
ARTIS_DEV void stats_block_flush(const Ctx &c) {
  __syncthreads();
  for (int i = threadIdx.x; i < 34; i += blockDim.x) {
    const int v = stats_block()[i];
    if (v) atomicAdd((unsigned long long *)&c.es.counters[i], (unsigned long long)v);
  }
}
#else
ARTIS_DEV void stat_inc(Tally &t, int counter) { t.stats[counter]++; }
#endif
ARTIS_DEV void tally_flush(const Ctx &c, const Tally &t) {
  long long *k = c.es.counters;
  if (t.rpkt_steps) ctr_add(&k[ARTIS_CTR_RPKT_STEPS], t.rpkt_steps);
  if (t.lines_scanned) ctr_add(&k[ARTIS_CTR_LINES_SCANNED], t.lines_scanned);
  if (t.lines_tau) ctr_add(&k[ARTIS_CTR_LINES_TAU], t.lines_tau);
  if (t.bf_evaluated) ctr_add(&k[ARTIS_CTR_BF_EVALUATED], t.bf_evaluated);
  if (t.estim_atomics) ctr_add(&k[ARTIS_CTR_ESTIM_ATOMICS], t.estim_atomics);
  if (t.gamma_steps) ctr_add(&k[ARTIS_CTR_GAMMA_STEPS], t.gamma_steps);
  if (t.ma_jumps) ctr_add(&k[ARTIS_CTR_MA_JUMPS], t.ma_jumps);
  if (t.ma_trans_scanned) ctr_add(&k[ARTIS_CTR_MA_TRANS_SCANNED], t.ma_trans_scanned);
  if (t.kpkt_events) ctr_add(&k[ARTIS_CTR_KPKT_EVENTS], t.kpkt_events);
  if (t.cool_scanned) ctr_add(&k[ARTIS_CTR_COOL_SCANNED], t.cool_scanned);
  if (t.interactions) ctr_add(&k[ARTIS_CTR_INTERACTIONS], t.interactions);
  if (t.soft_repairs) ctr_add(&k[ARTIS_CTR_SOFT_REPAIRS], t.soft_repairs);
  if (t.kappa_evals) ctr_add(&k[ARTIS_CTR_KAPPA_EVALS], t.kappa_evals);
  if (t.packet_updates) ctr_add(&k[ARTIS_CTR_PACKET_UPDATES], t.packet_updates);
  if (t.handoffs) ctr_add(&k[ARTIS_CTR_HANDOFFS], t.handoffs);
  if (t.bf_listed) ctr_add(&k[ARTIS_CTR_BF_LISTED], t.bf_listed);
#ifndef __CUDACC__
  for (int i = 0; i < 34; i++)
    if (t.stats[i]) ctr_add(&k[i], t.stats[i]);
#endif
}
// Event routines out of line. A step of a packet is straight-line code that every warp iteration runs once; what it may
// END in -- a scattering with its Stokes rotations, an emission, a macro-atom de-activation, a k-packet's cooling-channel
// search -- is rare per step, large (FP64 trigonometry, rejection loops) and was inlined at several places each: the
// transport kernel was 1.6 MB of SASS and its warps spent 8 cycles per issued instruction waiting for instructions. Each
// such routine now exists once per kernel behind a real call. Its working state goes through memory for the call (a copy
// of the packet's working set, its Philox stream and the work tally: EvState), so that the caller's packet never has its
// address taken and stays in registers on the per-step path.
struct EvState {
  Pkt p;
  Rng rng;
  Tally tl;
};
#ifdef __CUDA_ARCH__
#define ARTIS_EVENT_OOL 1
#else
#define ARTIS_EVENT_OOL 0
#endif

// the FIRST fatal code of a timestep is the one reported (later ones are usually its consequences)
ARTIS_DEV void fatal(const Ctx &c, int code) {
  if (c.error_flag == nullptr) return;
#ifdef __CUDA_ARCH__
  atomicCAS(c.error_flag, 0, code);
#else
  if (*c.error_flag == 0) *c.error_flag = code;
#endif
}

// stats::eventcounters (stats.h:48-84)
enum {
  COUNTER_MA_STAT_ACTIVATION_COLLEXC = 0, COUNTER_MA_STAT_ACTIVATION_COLLION = 1,
  COUNTER_MA_STAT_ACTIVATION_NTCOLLEXC = 2, COUNTER_MA_STAT_ACTIVATION_NTCOLLION = 3,
  COUNTER_MA_STAT_ACTIVATION_BB = 4, COUNTER_MA_STAT_ACTIVATION_BF = 5, COUNTER_MA_STAT_ACTIVATION_FB = 6,
  COUNTER_MA_STAT_DEACTIVATION_COLLDEEXC = 7, COUNTER_MA_STAT_DEACTIVATION_COLLRECOMB = 8,
  COUNTER_MA_STAT_DEACTIVATION_BB = 9, COUNTER_MA_STAT_DEACTIVATION_FB = 10, COUNTER_MA_STAT_INTERNALUPHIGHER = 11,
  COUNTER_MA_STAT_INTERNALUPHIGHERNT = 12, COUNTER_MA_STAT_INTERNALDOWNLOWER = 13, COUNTER_K_STAT_TO_MA_COLLEXC = 14,
  COUNTER_K_STAT_TO_MA_COLLION = 15, COUNTER_K_STAT_TO_R_FF = 16, COUNTER_K_STAT_TO_R_FB = 17,
  COUNTER_K_STAT_TO_R_BB = 18, COUNTER_K_STAT_FROM_FF = 19, COUNTER_K_STAT_FROM_BF = 20,
  COUNTER_NT_STAT_FROM_GAMMA = 21, COUNTER_NT_STAT_TO_IONIZATION = 22, COUNTER_NT_STAT_TO_EXCITATION = 23,
  COUNTER_NT_STAT_TO_KPKT = 24, COUNTER_K_STAT_FROM_EARLIERDECAY = 25, COUNTER_ESCOUNTER = 26,
  COUNTER_RESONANCESCATTERINGS = 27, COUNTER_CELLCROSSINGS = 28, COUNTER_UPSCATTER = 29, COUNTER_DOWNSCATTER = 30,
  COUNTER_UPDATECELL = 31, COUNTER_COOLINGRATECALCCOUNTER = 32, COUNTER_NESC = 33
};

}  // namespace artis
