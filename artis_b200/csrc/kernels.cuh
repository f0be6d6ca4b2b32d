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
ARTIS_DEV void body_ma(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.nlevels_total);
  const int mgi = c.mgi_of_tabcell[tc];
  if (cell_uses_tables(c, mgi)) build_ma_level(c, mgi, (int)(idx % c.st.nlevels_total), c.t_mid);
}
ARTIS_DEV void body_cool(const Ctx &c, const long long idx) {
  const int tc = (int)(idx / c.st.includedions);
  const int mgi = c.mgi_of_tabcell[tc];
  if (cell_uses_tables(c, mgi)) build_cooling_ion(c, mgi, (int)(idx % c.st.includedions));
}

// ---- K0 bucket_packets: replaces the std::sort of update_packets.cc:204-256 --------------------------------
// Counting sort of the ACTIVE packets by (model cell, packet kind). Dense inner cells come first, like the reference's
// comparator (density descending == model cell ascending for the 1D models; for 3D it is only a locality key), so the
// longest histories start first; packets that are escaped or already at t2 are left out of the queue.
ARTIS_DEV int bucket_key(const Ctx &c, const ArtisPacket &p) {
  int kind;
  switch (p.type) {
    case ARTIS_TYPE_KPKT:
    case ARTIS_TYPE_MA:
    case ARTIS_TYPE_NTLEPTON:
    case ARTIS_TYPE_PRE_KPKT:
    case ARTIS_TYPE_GAMMA_KPKT:
      kind = 0;
      break;
    case ARTIS_TYPE_RPKT:
      kind = 1;
      break;
    case ARTIS_TYPE_GAMMA:
      kind = 2;
      break;
    default:
      kind = 3;
  }
  return cell_mgi(c, p.where) * 4 + kind;
}
ARTIS_DEV bool packet_active(const Ctx &c, const ArtisPacket &p) { return p.type != ARTIS_TYPE_ESCAPE && p.prop_time < c.t2; }
ARTIS_DEV void body_bucket_hist(const Ctx &c, ArtisPacket *pkts, const int idx, const int pass, int *counts) {
  ArtisPacket &p = pkts[idx];
  if (pass == 0) {  // update_packets.cc:282-285, for every packet including the escaped ones
    p.interactions = 0;
    p.scat_count = 0;
  }
  if (packet_active(c, p)) int_add(&counts[bucket_key(c, p)], 1);
}
ARTIS_DEV void body_bucket_scatter(const Ctx &c, const ArtisPacket *pkts, const int idx, int *offsets, int *order) {
  const ArtisPacket &p = pkts[idx];
  if (packet_active(c, p)) order[int_fetch_add(&offsets[bucket_key(c, p)], 1)] = idx;
}
// exclusive prefix sum over the bucket counts (one thread; the bucket count is small next to the packet count)
ARTIS_DEV void body_bucket_scan(int *counts, const int nbuckets, int *n_active) {
  int run = 0;
  for (int b = 0; b < nbuckets; b++) {
    const int cnt = counts[b];
    counts[b] = run;
    run += cnt;
  }
  *n_active = run;
}

// ---- transport ------------------------------------------------------------------------------
// Persistent threads: every lane pulls the next packet from a global queue (`order`, packets bucketed by model cell;
// queue head = one atomic counter) as soon as its current packet has escaped, reached t2 or used up `max_events`
// events in this pass. One loop iteration = one event of one packet (do_packet: one r-packet step, one macro-atom
// jump, one k-packet / gamma / pellet event), so lanes holding the same kind of packet execute it in lock step and
// no lane idles while another works through a long history. Packets that stop on `max_events` are counted in
// *unfinished and picked up by the next pass (the passes of update_packets.cc:249-311).
ARTIS_DEV void body_transport(const Ctx &c, ArtisPacket *pkts, unsigned long long *rngctr, const int *order, const int n,
                              int *queue_head, const int pass, const long long max_events, int *unfinished) {
  const double t2 = c.t2;
  Tally tl = {};
  KappaCache k;
  k.valid = false;
  ArtisPacket p;
  Rng rng = {};
  int idx = -1;
  long long events = 0;
  bool have = false;
  while (true) {
    if (!have) {
      const int q = int_fetch_add(queue_head, 1);
      if (q >= n) break;
      idx = order ? order[q] : q;
      p = pkts[idx];
      if (pass == 0) {  // update_packets.cc:282-285
        p.interactions = 0;
        p.scat_count = 0;
      }
      if (p.type == ARTIS_TYPE_ESCAPE || !(p.prop_time < t2)) {
        if (pass == 0) pkts[idx] = p;
        continue;
      }
      rng = rng_init(c, p.number, pass == 0 ? 0ull : rngctr[idx]);
      k.valid = false;
      events = 0;
      have = true;
    }
    do_packet(c, p, t2, k, rng, tl);
    events++;
    const bool finished = (p.type == ARTIS_TYPE_ESCAPE) || !(p.prop_time < t2);
    if (finished || events >= max_events) {
      rngctr[idx] = rng.ctr;
      pkts[idx] = p;
      if (!finished) int_add(unfinished, 1);
      have = false;
    }
  }
  tally_flush(c, tl);
}

// ---- K10 deterministic probes --------------------------------------------------------------------------------
ARTIS_DEV void body_probe_boundary(const Ctx &c, const ArtisPacket *pkts, const int idx, double *sdist, int *snext,
                                   int *last_cross_out) {
  ArtisPacket p = pkts[idx];
  int sn = 0;
  sdist[idx] = boundary_cross(c, p, &sn);
  snext[idx] = sn;
  last_cross_out[idx] = p.last_cross;
}
ARTIS_DEV void body_probe_kappa(const Ctx &c, const ArtisPacket *pkts, const int idx, double *kappa) {
  const ArtisPacket p = pkts[idx];
  KappaCache k;
  k.valid = false;
  Tally tl = {};
  calculate_kappa_rpkt_cont(c, p, cell_mgi(c, p.where), k, tl);
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
  ArtisPacket p = pkts[idx];
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
  const double *pr = tab_ma_rates(c, tabcell(c, mgi[idx]), gl);
  for (int a = 0; a < MA_ACTION_COUNT; a++) rates[idx * MA_ACTION_COUNT + a] = pr[a];
}
ARTIS_DEV void body_probe_cooling(const Ctx &c, const int *mgi, const int *element, const int *ion, const int idx,
                                  double *contrib_total) {
  const int ui = uion_of(c, element[idx], ion[idx]);
  const int last = c.st.ion_coolingoffset[ui] + c.st.ion_ncoolingterms[ui] - 1;
  contrib_total[idx] = c.tb.cool_cum[(size_t)tabcell(c, mgi[idx]) * c.st.ncoolingterms + last];
}
ARTIS_DEV void body_probe_gamma_kappa(const Ctx &c, const ArtisPacket *pkts, const int idx, double *kap) {
  const ArtisPacket p = pkts[idx];
  const int mgi = cell_mgi(c, p.where);
  kap[idx * 3 + 0] = sig_comp(c, p, mgi);
  kap[idx * 3 + 1] = sig_photo_electric(c, p, mgi);
  kap[idx * 3 + 2] = sig_pair_prod(c, p, mgi);
}
ARTIS_DEV void body_probe_levelpop(const Ctx &c, const int *mgi, const int *element, const int *ion, const int *level,
                                   const int idx, double *pop) {
  pop[idx] = tab_levelpop(c, tabcell(c, mgi[idx]), glevel_of(c, element[idx], ion[idx], level[idx]));
}

}  // namespace artis
