// kernels.cuh -- per-thread bodies of every kernel. transport.cu wraps each body in a __global__ grid-stride loop;
// tests/hostsim wraps the same bodies in plain host loops (debugging aid, never shipped).
#pragma once
#include "gamma.cuh"

namespace artis {

// ---- K8 cell_precompute: one thread per table entry group ------------------------------------------------
ARTIS_DEV void body_levelpop(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.nlevels_total);
  const int gl = (int)(idx % c.st.nlevels_total);
  build_levelpop(c, c.mgi_of_tabcell[tc], gl);
}
ARTIS_DEV bool cell_uses_tables(const Ctx &c, const int mgi) { return c.cs.thick[mgi] != 1; }
ARTIS_DEV void body_bf(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.nbfcontinua);
  const int mgi = c.mgi_of_tabcell[tc];
  if (cell_uses_tables(c, mgi)) build_bf(c, mgi, (int)(idx % c.st.nbfcontinua));
}
ARTIS_DEV void body_ff_zsum(const Ctx &c, const long long tc, double *out) { out[tc] = kappa_ff_zsum(c, c.mgi_of_tabcell[tc]); }
// the live continua of a table cell (rpkt.cuh bf_is_live), from the pairs body_bf has written
ARTIS_DEV bool body_bf_live_test(const Ctx &c, const int tc, const int i) {
  const int mgi = c.mgi_of_tabcell[tc];
  if (!cell_uses_tables(c, mgi)) return false;
  const double *cell = c.tb.bf_cell + 2 * ((size_t)tc * c.st.nbfcontinua + i);
  return bf_is_live(c, mgi, cell[0], cell[1], c.cont_rec[i].nu_edge);
}
ARTIS_DEV void body_bf_live_store(const Ctx &c, const int tc, const int i, const long long at, double *pair, ContRec *rec) {
  const double *cell = c.tb.bf_cell + 2 * ((size_t)tc * c.st.nbfcontinua + i);
  pair[2 * at] = cell[0];
  pair[2 * at + 1] = cell[1];
  rec[at] = c.cont_rec[i];
}
ARTIS_DEV void body_corrphoto(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.nbfcontinua);
  const int mgi = c.mgi_of_tabcell[tc];
  if (cell_uses_tables(c, mgi)) build_corrphoto(c, mgi, (int)(idx % c.st.nbfcontinua));
}
ARTIS_DEV void body_ma(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.nlevels_total);
  const int mgi = c.mgi_of_tabcell[tc];
  if (cell_uses_tables(c, mgi)) build_ma_level(c, mgi, (int)(idx % c.st.nlevels_total), c.t_mid);
}
// calculate_cooling_rates (kpkt.cc:84-165): idx over (table cell, ion) -> contrib[mgi][ion]; then idx over table cells ->
// total[mgi] = the ions' sums added in (element, ion) order
ARTIS_DEV void body_cooling_rates_ion(const Ctx &c, const long long idx, double *contrib) {
  const int tc = (int)(idx / c.st.includedions);
  const int ui = (int)(idx % c.st.includedions);
  const int mgi = c.mgi_of_tabcell[tc];
  contrib[(size_t)mgi * c.st.includedions + ui] = cooling_rates_ion(c, mgi, ui);
}
ARTIS_DEV void body_cooling_rates_total(const Ctx &c, const long long tc, const double *contrib, double *total) {
  const int mgi = c.mgi_of_tabcell[tc];
  double C_total = 0.;
  for (int ui = 0; ui < c.st.includedions; ui++) C_total += contrib[(size_t)mgi * c.st.includedions + ui];
  total[mgi] = C_total;
}
ARTIS_DEV void body_cool(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.includedions);
  const int mgi = c.mgi_of_tabcell[tc];
  if (cell_uses_tables(c, mgi)) build_cooling_ion(c, mgi, (int)(idx % c.st.includedions));
}

// ---- event kinds -------------------------------------------------------------------------
// Packets are grouped by the KIND of event they need next; each kind has its own ring of packet indices, and a warp only
// ever executes one kind of event: the cheap table-driven macro-atom jumps are never serialised behind a lane that is
// doing an r-packet step with a continuum-opacity evaluation and a line sweep, nor behind one that emits a packet
// (measured on v1/v2: ~1.8 active lanes per instruction, profiles/r1_v1_*, r1_v2_*; round 2: 31 % of the macro-atom
// kind's instructions ran at 6.6 active lanes because one lane in a hundred was de-activating).
//   kind 0  macro-atom EVENTS (de-activations, jumps to another ion), k-packets, non-thermal leptons
//           (macroatom.cc:222-414,766-876, kpkt.cc, nonthermal.cc:1877)
//   kind 1  r-packet steps in a non-grey cell                       (rpkt.cc:623)
//   kind 2  gamma-packet steps                                      (gammapkt.cc:533)
//   kind 3  pellets and non-thermal pre-deposition                  (update_packets.cc:16,71)
//   kind 4  r-packet steps in a grey or empty cell
//   kind 5  macro-atom JUMPS within an ion (macroatom.cc:174-220,815-835): 98 % of all macro-atom iterations, 32 bytes of
//           packet state, run by their own low-register kernel (transport.cu k_transport_majump)
constexpr int NKIND = 6;
enum { KIND_MAK = 0, KIND_RPKT_THIN = 1, KIND_GAMMA = 2, KIND_OTHER = 3, KIND_RPKT_GREY = 4, KIND_MAJUMP = 5 };
// r-packets are two kinds: in a non-grey cell a step evaluates the continuum opacity and sweeps lines (expensive), in a
// grey or empty cell it is a boundary distance and a Thomson-like scattering (cheap); mixed in one warp the cheap lanes
// would idle through the others' opacity loops.
ARTIS_DEV int kind_of(const Ctx &c, const int type, const int where) {
  switch (type) {
    case ARTIS_TYPE_MA:
      return KIND_MAJUMP;
    case ARTIS_TYPE_KPKT:
    case ARTIS_TYPE_NTLEPTON:
    case ARTIS_TYPE_PRE_KPKT:
    case ARTIS_TYPE_GAMMA_KPKT:
      return KIND_MAK;
    case ARTIS_TYPE_RPKT: {
      const int mgi = cell_mgi(c, where);
      return (mgi == c.st.npts_model || c.cs.thick[mgi] == 1) ? KIND_RPKT_GREY : KIND_RPKT_THIN;
    }
    case ARTIS_TYPE_GAMMA:
      return KIND_GAMMA;
    default:
      return KIND_OTHER;
  }
}
ARTIS_DEV int kind_of(const Ctx &c, const Pkt &p) { return kind_of(c, p.type, p.where); }
// One ring of packet indices per kind (transport.cu), plus two scheduler-internal ones for the thin-cell r-packet step, whose
// continuum-opacity sum runs in the light kernel: KIND_KAPPA = a step that has begun and waits for its opacity (consumed
// by the opacity role), KIND_THIN_RESUME = its opacity has been delivered (consumed by the thin role, before new steps).
constexpr int KIND_KAPPA = NKIND;
constexpr int KIND_THIN_RESUME = NKIND + 1;
constexpr int NRING = NKIND + 2;
// Ring heads, tails and the active-packet count take every warp's atomics for the whole step: one 128-byte line each, so
// that they sit in different L2 slices instead of queueing behind one another (words of 8 bytes).
#ifndef ARTIS_RING_STRIDE
#define ARTIS_RING_STRIDE 16
#endif
constexpr int RING_META_STRIDE = ARTIS_RING_STRIDE;
constexpr int RING_META_WORDS = (2 * NRING + 3) * RING_META_STRIDE;  // heads, tails, active count, start time, partner flag

struct QueueView {
  const int *in;        // indices to consume
  const int *n_in;      // how many
  int *head;            // consumption cursor (zeroed before the launch)
  int *out[NKIND];      // queues being produced, one per kind
  int *n_out[NKIND];    // their sizes
};

// K0 bucket_packets: counting sort of the active packets by (kind, model cell) into the initial queues; replaces the
// std::sort of update_packets.cc:204-256 (dense inner cells first, like the reference's comparator for 1D models). Reads
// and writes sector 0 and the draw counter of the hot record only.
ARTIS_DEV bool hot_active(const Ctx &c, const PktHot &h) { return pkt_lo16(h.type_cross) != ARTIS_TYPE_ESCAPE && h.prop_time < c.t2; }
ARTIS_DEV int bucket_key(const Ctx &c, const PktHot &h) {
  return kind_of(c, pkt_lo16(h.type_cross), h.where) * (c.st.npts_model + 1) + cell_mgi(c, h.where);
}
// `fresh`: the first launch of a timestep. A relaunch after the row arena has filled (rows on demand, engine.inc) buckets the
// packets that have not finished and leaves their interaction counts and random streams where they are.
ARTIS_DEV void body_bucket_hist(const Ctx &c, PktHot *hot, PktCold *cold, const int idx, int *counts, const bool fresh) {
  PktHot &h = hot[idx];
  if (fresh) {
    h.interactions = 0;  // update_packets.cc:282-285, for every packet including the escaped ones
    cold[idx].scat_count = 0;
    h.rngctr = 0u;       // each packet's Philox stream restarts with the timestep
  }
  if (hot_active(c, h)) int_add(&counts[bucket_key(c, h)], 1);
}
// exclusive prefix sum over the (kind, cell) counts, one thread (host build; the device has k_bucket_scan); also the size
// of each kind's queue
ARTIS_DEV void body_bucket_scan(const Ctx &c, int *counts, int *kind_sizes) {
  const int per = c.st.npts_model + 1;
  for (int k = 0; k < NKIND; k++) {
    int run = 0;
    for (int b = 0; b < per; b++) {
      const int cnt = counts[k * per + b];
      counts[k * per + b] = run;
      run += cnt;
    }
    kind_sizes[k] = run;
  }
}
ARTIS_DEV void body_bucket_scatter(const Ctx &c, const PktHot *hot, const int idx, int *offsets, int *const *queues) {
  const PktHot &h = hot[idx];
  if (hot_active(c, h)) queues[kind_of(c, pkt_lo16(h.type_cross), h.where)][int_fetch_add(&offsets[bucket_key(c, h)], 1)] = idx;
}

// Kernel rounds per kind: the scheduler of the HOST build of these headers (the oracle's debugging aid; one "thread"
// drains a whole queue). The device runs the barrier-free scheduler of transport.cu instead.
template <int KIND>
ARTIS_DEV void body_wavefront(const Ctx &c, PktHot *hot, PktCold *cold, const QueueView &qv, const long long max_events) {
  constexpr int kind = KIND;
  const double t2 = c.t2;
  const int n = *qv.n_in;
  Tally tl = {};
  KappaCache k;
  k.valid = false;
  bool have = false;
  int idx = 0;
  long long events = 0;
  Pkt p;
  Rng rng;
  while (true) {
    if (!have) {
      const int q = int_fetch_add(qv.head, 1);
      if (q >= n) break;
      idx = qv.in[q];
      pkt_from_hot(p, hot[idx], cold + idx);
      rng = rng_init(c, p.number, hot[idx].rngctr);
      k.valid = false;
      events = 0;
      have = true;
    }
    do_packet_kind<KIND>(c, p, t2, k, rng, tl);
    if (ARTIS_OPT_VPKT(c)) vpkt_flush_request(c, p, k, tl);
    events++;
    const bool active = (p.type != ARTIS_TYPE_ESCAPE) && (p.prop_time < t2);
    if (!(active && kind_of(c, p) == kind && events < max_events)) {
      pkt_to_hot(hot[idx], p, (unsigned)rng.ctr);
      if (active) {
        const int k2 = kind_of(c, p);
        qv.out[k2][int_fetch_add(qv.n_out[k2], 1)] = idx;
      }
      have = false;
    }
  }
  tally_flush(c, tl);
}

// a probe works on a copy of the caller's ABI record
struct ProbePkt {
  Pkt p;
  PktHot hot;
  PktCold cold;
};
ARTIS_DEV void probe_load(ProbePkt &w, const ArtisPacket &a) {
  pkt_split(a, w.hot, w.cold);
  pkt_from_hot(w.p, w.hot, &w.cold);
}

// ---- K10 deterministic probes --------------------------------------------------------------------------------
ARTIS_DEV void body_probe_boundary(const Ctx &c, const ArtisPacket *pkts, const int idx, double *sdist, int *snext,
                                   int *last_cross_out) {
  ProbePkt w;
  probe_load(w, pkts[idx]);
  int sn = 0;
  sdist[idx] = boundary_cross(c, w.p, &sn);
  snext[idx] = sn;
  last_cross_out[idx] = w.p.last_cross;
}
ARTIS_DEV void body_probe_kappa(const Ctx &c, const ArtisPacket *pkts, const int idx, double *kappa) {
  ProbePkt w;
  probe_load(w, pkts[idx]);
  KappaCache k;
  k.valid = false;
  Tally tl = {};
  calculate_kappa_rpkt_cont(c, w.p, cell_mgi(c, w.p.where), k, tl);
  kappa[idx * 5 + 0] = k.total;
  kappa[idx * 5 + 1] = k.es;
  kappa[idx * 5 + 2] = k.ff;
  kappa[idx * 5 + 3] = k.bf;
  kappa[idx * 5 + 4] = k.ffheating;
}
ARTIS_DEV void body_probe_closest(const Ctx &c, const double *nu_cmf, const int *next_trans, const int idx, int *lineindex) {
  lineindex[idx] = closest_transition(c, nu_cmf[idx], next_trans[idx]);
}
ARTIS_DEV void body_probe_get_event(const Ctx &c, const ArtisPacket *pkts, const int idx, const double *tau_rnd,
                                    const double *abort_dist, double *edist, int *eventtype, int *next_trans_out,
                                    int *mastate_out) {
  ProbePkt w;
  probe_load(w, pkts[idx]);
  Pkt &p = w.p;
  KappaCache k;
  k.valid = false;
  Tally tl = {};
  int et = RPKT_EVENTTYPE_NONE;
  edist[idx] = get_event(c, cell_mgi(c, p.where), p, &et, tau_rnd[idx], abort_dist[idx], k, tl);
  eventtype[idx] = (et == RPKT_EVENTTYPE_BB) ? 1 : (et == RPKT_EVENTTYPE_CONT ? 2 : -1);
  next_trans_out[idx] = p.next_trans;
  mastate_out[idx * 4 + 0] = p.ma_element;
  mastate_out[idx * 4 + 1] = p.ma_ion;
  mastate_out[idx * 4 + 2] = p.ma_level;
  mastate_out[idx * 4 + 3] = p.ma_activatingline;
}
ARTIS_DEV void body_probe_ma_rates(const Ctx &c, const int *mgi, const int *element, const int *ion, const int *level,
                                   const int idx, double *rates) {
  const int gl = glevel_of(c, element[idx], ion[idx], level[idx]);
  const double *pr = ma_rates_row(c, mgi[idx], tabcell(c, mgi[idx]), gl);
  for (int a = 0; a < MA_ACTION_COUNT; a++) rates[idx * MA_ACTION_COUNT + a] = pr[a];
}
ARTIS_DEV void body_probe_cooling(const Ctx &c, const int *mgi, const int *element, const int *ion, const int idx,
                                  double *contrib_total) {
  const int ui = uion_of(c, element[idx], ion[idx]);
  const int last = c.st.ion_coolingoffset[ui] + c.st.ion_ncoolingterms[ui] - 1;
  contrib_total[idx] = cool_row(c, mgi[idx], tabcell(c, mgi[idx]), ui).cool[last];
}
ARTIS_DEV void body_probe_gamma_kappa(const Ctx &c, const ArtisPacket *pkts, const int idx, double *kap) {
  ProbePkt w;
  probe_load(w, pkts[idx]);
  const Pkt &p = w.p;
  const int mgi = cell_mgi(c, p.where);
  kap[idx * 3 + 0] = sig_comp(c, p, mgi);
  kap[idx * 3 + 1] = sig_photo_electric(c, p, mgi);
  kap[idx * 3 + 2] = sig_pair_prod(c, p, mgi);
}
// select_continuum_nu (ratecoeff.cc:628) with the random number injected: (element, lower ion, lower level, level of the
// upper ion, T_e, zrand) -> frequency
ARTIS_DEV void body_probe_scnu(const Ctx &c, const int *element, const int *ion, const int *level, const int *upper, const float *T_e,
                               const double *zrand, const int idx, double *nu) {
  const int ui = uion_of(c, element[idx], ion[idx]);
  nu[idx] = select_continuum_nu_z(c, ui, c.st.ion_leveloffset[ui] + level[idx], upper[idx], T_e[idx], zrand[idx]);
}
ARTIS_DEV void body_probe_levelpop(const Ctx &c, const int *mgi, const int *element, const int *ion, const int *level,
                                   const int idx, double *pop) {
  pop[idx] = tab_levelpop(c, tabcell(c, mgi[idx]), glevel_of(c, element[idx], ion[idx], level[idx]));
}

}  // namespace artis
