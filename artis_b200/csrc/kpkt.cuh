// kpkt.cuh -- thermal k-packets: kpkt.cc:428-797 do_kpkt / do_kpkt_bb / sample_planck, and
// nonthermal::do_ntlepton (nonthermal.cc:1877-1978) for builds without the Spencer-Fano solver.
// The cumulative cooling list of the packet's cell is a row of CellTables.cool_cum; the reference fills the same
// numbers lazily per host thread (calculate_kpkt_rates_ion, kpkt.cc:167).
#pragma once
#include "macroatom.cuh"

namespace artis {

// kpkt.cc:428-446
ARTIS_DEV double sample_planck(const Ctx &c, const double T, Rng &rng) {
  const double nu_peak = 5.879e10 * T;
  const double B_peak = dbb(nu_peak, (float)T, 1.f);
  while (true) {
    const double zrand = rng_uniform(rng);
    const double zrand2 = rng_uniform(rng);
    const double nu = c.st.nu_min_r + zrand * (c.st.nu_max_r - c.st.nu_min_r);
    if (zrand2 * B_peak <= dbb(nu, (float)T, 1.f)) return nu;
  }
}

// kpkt.cc:448-475
ARTIS_DEV void do_kpkt_bb_body(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
  const int mgi = cell_mgi(c, p.where);
  const float T_e = c.cs.Te[mgi];
  p.nu_cmf = sample_planck(c, (double)T_e, rng);
  if (!isfinite(p.nu_cmf)) fatal(c, 301);
  emitt_rpkt(c, p, rng);
  p.next_trans = 0;
  stat_inc(tl, COUNTER_K_STAT_TO_R_BB);
  p.interactions++;
  tl.interactions++;
  p.last_event = 6;
  p.cold->emissiontype = -9999999;
  mark_emission(p);
}
ARTIS_DEV_NOINLINE void do_kpkt_bb_ool(const Ctx &c, EvState *s) { do_kpkt_bb_body(c, s->p, s->rng, s->tl); }
ARTIS_DEV void do_kpkt_bb(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  do_kpkt_bb_ool(c, &s);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  do_kpkt_bb_body(c, p, rng, tl);
#endif
}

// std::lower_bound over cum[lo .. hi]: first index with cum[i] >= value (hi + 1 if none)
ARTIS_DEV int lower_bound_idx(const double *cum, int lo, const int hi, const double value) {
  int end = hi + 1;
  while (lo < end) {
    const int mid = (lo + end) >> 1;
    if (cum[mid] < value)
      lo = mid + 1;
    else
      end = mid;
  }
  return lo;
}

// The ion whose cooling row do_kpkt is about to search, without touching the packet or its random stream (rows on demand: the
// scheduler has the warp build a missing row before the event, transport.cu): the first statements of do_kpkt_body on a COPY
// of the stream. -1: the event ends before it selects a cooling process.
ARTIS_DEV int kpkt_peek_cooling_ion(const Ctx &c, const Pkt &p, const double t2, const Rng &rng) {
  const double t1 = p.prop_time;
  const int mgi = cell_mgi(c, p.where);
  double deltat = 0.;
  if (c.nts < c.st.n_kpktdiffusion_timesteps) deltat = c.st.kpktdiffusion_timescale * c.ts_width;
  if (!(t1 + deltat <= t2)) return -1;
  Rng peek = rng;
  const double rndcool = rng_uniform(peek) * c.cs.totalcooling[mgi];
  double coolingsum = 0.;
  const int nions_total = c.st.includedions;
  for (int ui = 0; ui < nions_total; ui++) {
    coolingsum += c.cs.cooling_contrib_ion[(size_t)mgi * nions_total + ui];
    if (coolingsum > rndcool) return ui;
  }
  return -1;
}

// kpkt.cc:477-797
ARTIS_DEV void do_kpkt_body(const Ctx &c, Pkt &p, const double t2, Rng &rng, Tally &tl) {
  tl.kpkt_events++;
  const double t1 = p.prop_time;
  const int mgi = cell_mgi(c, p.where);
  const int tc = tabcell(c, mgi);
  const float T_e = c.cs.Te[mgi];
  double deltat = 0.;
  if (c.nts < c.st.n_kpktdiffusion_timesteps) deltat = c.st.kpktdiffusion_timescale * c.ts_width;
  const double t_current = t1 + deltat;
  if (!(t_current <= t2)) {
    vec_scale(p.pos, t2 / t1);
    p.prop_time = t2;
    return;
  }
  vec_scale(p.pos, t_current / t1);
  p.prop_time = t_current;

  const double zrand = rng_uniform(rng);
  const double rndcool = zrand * c.cs.totalcooling[mgi];
  double coolingsum = 0., oldcoolingsum = 0.;
  const int nions_total = c.st.includedions;
  int ui;
  for (ui = 0; ui < nions_total; ui++) {  // (element, ion) order == unique ion index order
    oldcoolingsum = coolingsum;
    coolingsum += c.cs.cooling_contrib_ion[(size_t)mgi * nions_total + ui];
    if (coolingsum > rndcool) break;
  }
  if (ui >= nions_total || tc < 0) {  // "do_kpkt: problem selecting a cooling process ... abort"
    fatal(c, 302);
    p.prop_time = t2;
    return;
  }
  const int ilow = c.st.ion_coolingoffset[ui];
  const int ihigh = ilow + c.st.ion_ncoolingterms[ui] - 1;
  const CoolRow crow = cool_row(c, mgi, tc, ui);
  const double *cool = crow.cool;
  int i = lower_bound_idx(cool, ilow, ihigh, rndcool);
  tl.cool_scanned += bsearch_probes(ihigh - ilow + 1) + ui + 1;  // list probes + ion totals scanned
  if (i > ihigh) {  // assert_always(i <= ihigh) in the reference
    fatal(c, 303);
    p.prop_time = t2;
    return;
  }
  const int type = c.st.cool_type[i];
  const int element = c.uion_element[ui];
  const int ion = ui - c.st.el_ionoffset[element];
  if (type == COOLINGTYPE_FF) {
    const double z = rng_uniform_pos(rng);
    p.nu_cmf = -KB * T_e / H * log(z);
    if (!isfinite(p.nu_cmf)) fatal(c, 304);
    emitt_rpkt(c, p, rng);
    p.next_trans = 0;
    stat_inc(tl, COUNTER_K_STAT_TO_R_FF);
    p.interactions += 1;
    tl.interactions++;
    p.last_event = 6;
    p.cold->emissiontype = -9999999;
    mark_emission(p);
    if (ARTIS_OPT_VPKT(c)) tl.vpkt_req = 2;  // kpkt.cc:631-634 (t_current == prop_time)
  } else if (type == COOLINGTYPE_FB) {
    const int level = c.st.cool_level[i];
    const int upper = c.st.cool_upperlevel[i];
    const int gl_lower = c.st.ion_leveloffset[ui] + level;
    p.nu_cmf = select_continuum_nu(c, ui, gl_lower, upper, T_e, rng);
    emitt_rpkt(c, p, rng);
    p.next_trans = 0;
    stat_inc(tl, COUNTER_K_STAT_TO_R_FB);
    p.interactions += 1;
    tl.interactions++;
    p.last_event = 7;
    p.cold->emissiontype = c.st.lev_cont_index[gl_lower] - phixstargetindex_of(c, gl_lower, upper);
    p.cold->trueemissiontype = p.cold->emissiontype;
    mark_emission(p);
    if (ARTIS_OPT_VPKT(c)) tl.vpkt_req = 2;  // kpkt.cc:691-694
  } else if (type == COOLINGTYPE_COLLEXC) {
    // kpkt.cc:695-759: walk the level's upward transitions until the running contribution reaches rndcool; the
    // running values are CellTables.collexc_cum (same arithmetic, same order), so this is a search for the first
    // entry >= rndcool
    const int level = c.st.cool_level[i];
    const int gl = c.st.ion_leveloffset[ui] + level;
    const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
    const double *cexc = crow.cexc;
    const int ii = lower_bound_idx(cexc, uoff, uoff + nup - 1, rndcool);
    if (ii >= uoff + nup) {  // "WARNING: Could not select an upper level" + abort
      fatal(c, 305);
      p.prop_time = t2;
      return;
    }
    const int lineindex = c.st.uptrans_lineindex[ii];
    p.ma_element = element;
    p.ma_ion = ion;
    p.ma_level = c.st.line_upper[lineindex];
    p.ma_activatingline = -99;
    p.type = ARTIS_TYPE_MA;
    stat_inc(tl, COUNTER_MA_STAT_ACTIVATION_COLLEXC);
    stat_inc(tl, COUNTER_K_STAT_TO_MA_COLLEXC);
    p.interactions += 1;
    tl.interactions++;
    p.last_event = 8;
    p.cold->trueemissiontype = -1;
    p.cold->trueemissionvelocity = -1;
  } else if (type == COOLINGTYPE_COLLION) {
    p.ma_element = element;
    p.ma_ion = ion + 1;
    p.ma_level = c.st.cool_upperlevel[i];
    p.ma_activatingline = -99;
    p.type = ARTIS_TYPE_MA;
    stat_inc(tl, COUNTER_MA_STAT_ACTIVATION_COLLION);
    stat_inc(tl, COUNTER_K_STAT_TO_MA_COLLION);
    p.interactions += 1;
    tl.interactions++;
    p.last_event = 9;
    p.cold->trueemissiontype = -1;
    p.cold->trueemissionvelocity = -1;
  } else {
    fatal(c, 306);
    p.prop_time = t2;
  }
}
ARTIS_DEV_NOINLINE void do_kpkt_ool(const Ctx &c, EvState *s, const double t2) { do_kpkt_body(c, s->p, t2, s->rng, s->tl); }
ARTIS_DEV void do_kpkt(const Ctx &c, Pkt &p, const double t2, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  do_kpkt_ool(c, &s, t2);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  do_kpkt_body(c, p, t2, rng, tl);
#endif
}

// nonthermal.cc:1877-1978 do_ntlepton. Without the Spencer-Fano solution (classic, kilonova_lte) every non-thermal
// lepton thermalises into a k-packet; with it (nebular NLTE) a fraction get_ntion_energyrate / deposition_rate_density
// activates a macro-atom by non-thermal ionisation (select_nt_ionization2, nonthermal.cc:1858-1875). NT_EXCITATION_ON
// is false in every shipped options header (frac_excitation = 0., nonthermal.cc:1898).
ARTIS_DEV void do_ntlepton_body(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
  est_add(&c.es.ts_scalars[ARTIS_TS_NT_ENERGY_DEPOSITED], p.e_cmf);
  const int mgi = cell_mgi(c, p.where);
  if (ARTIS_OPT_NT_SF(c) && c.cs.thick[mgi] != 1) {
    const double zrand = rng_uniform(rng);
    // get_ntion_energyrate (nonthermal.cc:1847-1856): the same sum serves frac_ionization and select_nt_ionization2
    double ratetotal = 0.;
    for (int el = 0; el < c.st.nelements; el++) {
      const int nions = c.st.el_nions[el];
      for (int ion = 0; ion < nions - 1; ion++)
        ratetotal += ion_ntion_energyrate(c, mgi, el, c.st.el_ionoffset[el] + ion, nions, ion);
    }
    const double frac_ionization = ratetotal / c.cs.nt_deposition_rate_density[mgi];
    if (zrand < frac_ionization) {
      const double zrand2 = rng_uniform(rng);
      double ratesum = 0.;
      int element = -1, lowerion = -1;
      for (int el = 0; el < c.st.nelements && element < 0; el++) {
        const int nions = c.st.el_nions[el];
        for (int ion = 0; ion < nions - 1; ion++) {
          ratesum += ion_ntion_energyrate(c, mgi, el, c.st.el_ionoffset[el] + ion, nions, ion);
          if (ratesum >= zrand2 * ratetotal) {
            element = el;
            lowerion = ion;
            break;
          }
        }
      }
      if (element < 0) {  // assert_always(false) in the reference
        fatal(c, 263);
        p.type = ARTIS_TYPE_KPKT;
        return;
      }
      const int upperion = lowerion + nt_random_upperion(c, mgi, c.st.el_ionoffset[element] + lowerion,
                                                         c.st.el_nions[element], lowerion, true, rng);
      p.ma_element = element;
      p.ma_ion = upperion;
      p.ma_level = 0;
      p.ma_activatingline = -99;
      p.type = ARTIS_TYPE_MA;
      stat_inc(tl, COUNTER_MA_STAT_ACTIVATION_NTCOLLION);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 20;
      p.cold->trueemissiontype = -1;
      p.cold->trueemissionvelocity = -1;
      stat_inc(tl, COUNTER_NT_STAT_TO_IONIZATION);
      return;
    }
  }
  p.last_event = 22;
  p.type = ARTIS_TYPE_KPKT;
  stat_inc(tl, COUNTER_NT_STAT_TO_KPKT);
}
ARTIS_DEV_NOINLINE void do_ntlepton_ool(const Ctx &c, EvState *s) { do_ntlepton_body(c, s->p, s->rng, s->tl); }
ARTIS_DEV void do_ntlepton(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  do_ntlepton_ool(c, &s);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  do_ntlepton_body(c, p, rng, tl);
#endif
}

}  // namespace artis
