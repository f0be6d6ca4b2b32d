// macroatom.cuh -- macro-atom state machine: macroatom.cc:174-901 and select_continuum_nu
// (ratecoeff.cc:628-684). Process rates and their running sums come from the per-cell tables (celltables.cuh); the
// reference's linear scans `rate += x[i]; if (zrand * total < rate) break;` become a binary search for the first
// prefix sum greater than zrand * total, which picks the same index because the prefix sums are the same FP64
// running sums.
#pragma once
#include "rpkt_events.cuh"

namespace artis {

// number of table entries a binary search over n entries reads (the work counters feed the roofline)
ARTIS_DEV int bsearch_probes(int n) {
  int k = 0;
  while (n > 0) {
    k++;
    n >>= 1;
  }
  return k;
}
// first index i in [0, n) with cum[i] > target, n if none
// `cum` is a prefix sum of non-negative rates (non-decreasing), so the answer is unique and the search may probe in any
// pattern. Eight-way: each round issues seven independent loads and keeps an eighth of the range, i.e. log8(n) round
// trips to L2 instead of log2(n) dependent ones -- a macro-atom jump is a chain of such round trips and nothing else.
ARTIS_DEV int first_greater(const double *cum, const int n, const double target) {
  int lo = 0, hi = n;  // the answer lies in [lo, hi]
  while (hi - lo > 7) {
    const int step = (hi - lo) >> 3;
    int below = 0;  // pivots lo + j*step - 1, j = 1..7, that are <= target: a prefix of them
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int j = 1; j <= 7; j++) below += (cum[lo + j * step - 1] <= target) ? 1 : 0;
    if (below < 7) hi = lo + (below + 1) * step - 1;
    lo += below * step;
  }
  int below = 0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
  for (int j = 0; j < 7; j++)
    if (lo + j < hi) below += (cum[lo + j] <= target) ? 1 : 0;
  return lo + below;
}

// ratecoeff.cc:264-280 alpha_sp_E_integrand_gsl
ARTIS_DEV double alpha_sp_E_integrand(const Ctx &c, const int tablerow, const double nu_edge, const float T, const double nu) {
  const float sigma_bf = (float)phixs_fromtable(c, tablerow, nu_edge, nu);
  return 2. / CLIGHTSQUARED * sigma_bf * pow(nu, 3) / nu_edge * exp(-HOVERKB * nu / T);
}
// 8-point Gauss-Legendre on [a, b]
ARTIS_DEV double gl8(const Ctx &c, const int tablerow, const double nu_edge, const float T, const double a, const double b) {
  const double x[4] = {0.1834346424956498, 0.5255324099163290, 0.7966664774136267, 0.9602898564975363};
  const double w[4] = {0.3626837833783620, 0.3137066458778873, 0.2223810344533745, 0.1012285362903763};
  const double h = 0.5 * (b - a), m = 0.5 * (a + b);
  double s = 0.;
  for (int i = 0; i < 4; i++)
    s += w[i] * (alpha_sp_E_integrand(c, tablerow, nu_edge, T, m - h * x[i]) +
                 alpha_sp_E_integrand(c, tablerow, nu_edge, T, m + h * x[i]));
  return s * h;
}
// ratecoeff.cc:628-684 select_continuum_nu. The reference evaluates up to NPHIXSPOINTS adaptive GK31 integrals of the
// tail [xlow_i, nu_max] to CONTINUUM_NU_INTEGRAL_ACCURACY (1e-2 classic / 1e-3 nebular); here the NPHIXSPOINTS piece
// integrals are evaluated once with a fixed 8-point rule per smooth sub-interval (pieces are split at the knots of
// the cross-section table) and the tails are their suffix sums. Selection rule and interpolation are the
// reference's, including the extrapolation when no piece is selected. Tolerance: |dnu/nu| <= 1e-3 (tests).
ARTIS_DEV_NOINLINE double select_continuum_nu_z(const Ctx &c, const int ui_lower, const int gl_lower, const int upperionlevel,
                                                const float T_e, const double zrand) {
  const int t = phixstargetindex_of(c, gl_lower, upperionlevel);
  if (t < 0) {
    fatal(c, 210);
    return 1e15;
  }
  const double E_threshold = phixs_threshold(c, ui_lower, gl_lower, t);
  const double nu_threshold = ONEOVERH * E_threshold;
  const double nu_max_phixs = nu_threshold * c.st.last_phixs_nuovernuedge;
  const int npieces = c.st.nphixspoints;
  const int tablerow = c.st.lev_phixstable[gl_lower];
  const double deltanu = (nu_max_phixs - nu_threshold) / npieces;
  const double knotstep = nu_threshold * c.st.nphixsnuincrement;
  // pass 1: total; pass 2: walk the tails from the blue end of the reference's loop
  double total = 0.;
  for (int pass = 0; pass < 2; pass++) {
    double done = 0.;  // integral over [nu_threshold, xlow_i)
    double alpha_sp_old = total, alpha_sp = total;
    int i = 1;
    for (int j = 0; j < npieces; j++) {
      const double a = nu_threshold + j * deltanu;
      const double b = (j == npieces - 1) ? nu_max_phixs : nu_threshold + (j + 1) * deltanu;
      // split [a, b] at table knots
      double piece = 0.;
      double lo = a;
      int kn = (int)floor((a - nu_threshold) / knotstep) + 1;
      while (true) {
        const double knot = nu_threshold + kn * knotstep;
        if (knot > lo && knot < b) {
          piece += gl8(c, tablerow, nu_threshold, T_e, lo, knot);
          lo = knot;
          kn++;
        } else if (knot <= lo) {
          kn++;
        } else {
          break;
        }
      }
      piece += gl8(c, tablerow, nu_threshold, T_e, lo, b);
      done += piece;
      if (pass == 1 && j + 1 < npieces) {
        i = j + 1;
        alpha_sp_old = alpha_sp;
        alpha_sp = total - done;
        if (alpha_sp < 0.) alpha_sp = 0.;
        if (zrand >= alpha_sp / total) {
          const double nuoffset = (total * zrand - alpha_sp_old) / (alpha_sp - alpha_sp_old) * deltanu;
          return nu_threshold + (i - 1) * deltanu + nuoffset;
        }
      }
    }
    if (pass == 0) {
      total = done;
    } else {
      // no break: i == npieces in the reference, alpha_sp(_old) hold the last two tails (ratecoeff.cc:676-677)
      const double nuoffset = (total * zrand - alpha_sp_old) / (alpha_sp - alpha_sp_old) * deltanu;
      return nu_threshold + (npieces - 1) * deltanu + nuoffset;
    }
  }
  return nu_threshold;
}
// the draw of ratecoeff.cc:643: zrand = 1 - uniform in (0, 1]
ARTIS_DEV double select_continuum_nu(const Ctx &c, const int ui_lower, const int gl_lower, const int upperionlevel, const float T_e,
                                     Rng &rng) {
  const double zrand = 1. - rng_uniform(rng);
  return select_continuum_nu_z(c, ui_lower, gl_lower, upperionlevel, T_e, zrand);
}

// common tail of every r-packet emission from a macro-atom / k-packet
ARTIS_DEV void mark_emission(Pkt &p) {
  p.cold->em_pos[0] = p.pos[0];
  p.cold->em_pos[1] = p.pos[1];
  p.cold->em_pos[2] = p.pos[2];
  p.cold->em_time = (int)p.prop_time;
  p.cold->nscatterings = 0;
}

// macroatom.cc:416-901 do_macroatom, ONE iteration of its `while (!end_packet)` loop per call: the macro-atom state
// (ion, level) lives in the packet's mastate between calls, so that all lanes of a warp that hold an activated
// macro-atom take their jumps in lock step and the scheduler can interleave packet kinds (the reference keeps the
// state in locals and never reads mastate again after the chain has started, so nothing observable changes).
//
// The iteration is written in two parts, because 98 % of them are jumps within the ion and need nothing but the
// macro-atom state, one row of process rates and one search over a prefix-sum row:
//   ma_jump   draws the action and, for the three list-driven actions, the transition; completes the jump when it stays
//             inside the ion (internal down / up: the new level is all that changes); otherwise returns the action.
//             Works on any packet view with where, ma_element, ma_ion, ma_level, interactions -- the scheduler's jump
//             role runs it on 32 bytes of the packet, at a small register budget and therefore many resident warps.
//   ma_event  completes the other actions (de-activations with their emission, jumps to another ion) on the full packet.
// Run back to back they draw the same numbers in the same order as the one-piece loop body of the reference.
constexpr int MA_ACTION_FATAL = 99;  // ma_jump hit one of the reference's abort() branches; ma_event parks the packet
template <class P>
ARTIS_DEV int ma_jump(const Ctx &c, P &p, Rng &rng, Tally &tl, int *sel_out) {
  const int mgi = cell_mgi(c, p.where);
  const int tc = tabcell(c, mgi);
  *sel_out = 0;
  if (c.cs.thick[mgi] == 1 || tc < 0) {
    fatal(c, 220);
    return MA_ACTION_FATAL;
  }
  const int ui = uion_of(c, p.ma_element, p.ma_ion);
  const int gl = c.st.ion_leveloffset[ui] + p.ma_level;
  tl.ma_jumps++;
  const MaRow row = ma_row(c, mgi, tc, gl);
  const double *processrates = row.pr;
  double pr[MA_ACTION_COUNT];
  double total_transitions = 0.;
  for (int a = 0; a < MA_ACTION_COUNT; a++) {
    pr[a] = processrates[a];
    total_transitions += pr[a];
  }
  const double randomrate = rng_uniform(rng) * total_transitions;
  double rate = 0., pr_sel = 0.;
  int selected_action = MA_ACTION_COUNT;
  for (int a = 0; a < MA_ACTION_COUNT; a++) {  // first action whose running sum exceeds the draw (macroatom.cc:470-478)
    rate += pr[a];
    if (selected_action == MA_ACTION_COUNT && rate > randomrate) {
      selected_action = a;
      pr_sel = pr[a];
    }
  }
  if (selected_action == MA_ACTION_COUNT) {  // reference: "[fatal] do_ma: problem with random numbers" + abort
    fatal(c, 221);
    return MA_ACTION_FATAL;
  }
  // The three list-driven actions (radiative de-excitation, internal jump down / up within the ion) share one draw and
  // one search over a per-cell prefix-sum row.
  const double *cum = nullptr;
  int ntrans = 0, toff = 0;
  if (selected_action == MA_RADDEEXC || selected_action == MA_INTERNALDOWNSAME) {
    ntrans = c.st.lev_ndowntrans[gl];
    toff = c.st.lev_downoffset[gl];
    cum = selected_action == MA_RADDEEXC ? row.cum_rad : row.cum_down;
  } else if (selected_action == MA_INTERNALUPSAME) {
    ntrans = c.st.lev_nuptrans[gl];
    toff = c.st.lev_upoffset[gl];
    cum = row.cum_up;
  }
  if (cum == nullptr) return selected_action;
  const double zrand = rng_uniform(rng);
  const int sel = first_greater(cum, ntrans, zrand * pr_sel);
  tl.ma_trans_scanned += bsearch_probes(ntrans);
  if (sel >= ntrans) {
    fatal(c, selected_action == MA_RADDEEXC ? 222 : (selected_action == MA_INTERNALDOWNSAME ? 223 : 227));
    return MA_ACTION_FATAL;
  }
  *sel_out = toff + sel;
  if (selected_action == MA_RADDEEXC) return selected_action;
  // internal jump within the ion (macroatom.cc:174-220, 815-835): remember where the macro-atom is
  p.interactions += 1;
  tl.interactions++;
  const int gl_new = selected_action == MA_INTERNALDOWNSAME ? c.downtrans_target_g[toff + sel] : c.uptrans_target_g[toff + sel];
  p.ma_level = gl_new - c.st.ion_leveloffset[ui];
  return -1;
}

// `action` and `selidx` are what ma_jump returned (selidx: index of the selected downward transition for MA_RADDEEXC)
ARTIS_DEV void ma_event_body(const Ctx &c, Pkt &p, const int action, const int selidx, Rng &rng, Tally &tl) {
  if (action == MA_ACTION_FATAL) {
    p.type = ARTIS_TYPE_KPKT;  // park the packet; the run is reported as failed
    return;
  }
  const int mgi = cell_mgi(c, p.where);
  const float T_e = c.cs.Te[mgi];
  const float nne = c.cs.nne[mgi];
  const int element = p.ma_element;
  int ui = uion_of(c, element, p.ma_ion);
  int gl = c.st.ion_leveloffset[ui] + p.ma_level;
  const int activatingline = p.ma_activatingline;
  const int level = p.ma_level;
  const double epsilon_current = c.st.lev_epsilon[gl];
  const double pr_sel = ma_rates_row(c, mgi, tabcell(c, mgi), gl)[action < MA_ACTION_COUNT ? action : 0];
  bool end_packet = false;
  switch (action) {
    case MA_RADDEEXC: {  // macroatom.cc:222-296
      const int linelistindex = c.st.downtrans_lineindex[selidx];
      if (c.st.opt_record_linestat) int_add(&c.es.ecounter[linelistindex], 1);
      const int gll = c.line_lower_g[linelistindex];
      const double epsilon_trans = epsilon_current - c.st.lev_epsilon[gll];
      const double oldnucmf = p.nu_cmf;
      p.nu_cmf = epsilon_trans / H;
      if (p.last_event == 1) stat_inc(tl, oldnucmf < p.nu_cmf ? COUNTER_UPSCATTER : COUNTER_DOWNSCATTER);
      stat_inc(tl, COUNTER_MA_STAT_DEACTIVATION_BB);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 0;
      emitt_rpkt(c, p, rng);
      if (linelistindex == activatingline) stat_inc(tl, COUNTER_RESONANCESCATTERINGS);
      p.next_trans = linelistindex + 1;
      p.cold->emissiontype = linelistindex;
      mark_emission(p);
      if (ARTIS_OPT_VPKT(c)) tl.vpkt_req = 3;  // macroatom.cc:292-295
      end_packet = true;
      break;
    }
    case MA_COLDEEXC: {
      stat_inc(tl, COUNTER_MA_STAT_DEACTIVATION_COLLDEEXC);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 10;
      p.type = ARTIS_TYPE_KPKT;
      end_packet = true;
      tl.estim_atomics += est_add(&c.es.colheatingestimator[mgi], p.e_cmf);
      break;
    }
    case MA_RADRECOMB: {  // macroatom.cc:298-380
      const int gl0_lower = c.st.ion_leveloffset[ui - 1];
      const int nlevels = c.st.ion_ionisinglevels[ui - 1];
      const double zrand = rng_uniform(rng);
      double rrate = 0.;
      int lower;
      for (lower = 0; lower < nlevels; lower++) {
        const double epsilon_trans = epsilon_current - c.st.lev_epsilon[gl0_lower + lower];
        const double R = rad_recombination_ratecoeff(c, T_e, nne, gl0_lower + lower, level);
        rrate += R * epsilon_trans;
        if (zrand * pr_sel < rrate) break;
      }
      tl.ma_trans_scanned += 3 * nlevels;  // LUT rows read per lower level
      if (zrand * pr_sel >= rrate) {
        fatal(c, 224);
        p.type = ARTIS_TYPE_KPKT;
        return;
      }
      const int gl_lower = gl0_lower + lower;
      p.nu_cmf = select_continuum_nu(c, ui - 1, gl_lower, level, T_e, rng);
      if (!isfinite(p.nu_cmf)) fatal(c, 225);
      stat_inc(tl, COUNTER_MA_STAT_DEACTIVATION_FB);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 2;
      emitt_rpkt(c, p, rng);
      p.next_trans = 0;
      p.cold->emissiontype = c.st.lev_cont_index[gl_lower] - phixstargetindex_of(c, gl_lower, level);
      mark_emission(p);
      if (ARTIS_OPT_VPKT(c)) tl.vpkt_req = 3;  // macroatom.cc:376-379
      ui -= 1;
      gl = gl_lower;
      end_packet = true;
      break;
    }
    case MA_COLRECOMB: {
      stat_inc(tl, COUNTER_MA_STAT_DEACTIVATION_COLLRECOMB);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 11;
      p.type = ARTIS_TYPE_KPKT;
      end_packet = true;
      tl.estim_atomics += est_add(&c.es.colheatingestimator[mgi], p.e_cmf);
      break;
    }
    case MA_INTERNALDOWNLOWER: {  // macroatom.cc:766-813
      p.interactions += 1;
      tl.interactions++;
      stat_inc(tl, COUNTER_MA_STAT_INTERNALDOWNLOWER);
      const double zrand = rng_uniform(rng);
      const int gl0_lower = c.st.ion_leveloffset[ui - 1];
      const int nlevels = c.st.ion_ionisinglevels[ui - 1];
      double rrate = 0.;
      int lower;
      for (lower = 0; lower < nlevels; lower++) {
        const double epsilon_target = c.st.lev_epsilon[gl0_lower + lower];
        const double epsilon_trans = epsilon_current - epsilon_target;
        const double R = rad_recombination_ratecoeff(c, T_e, nne, gl0_lower + lower, level);
        const double C = col_recombination_ratecoeff(c, mgi, ui, gl, gl0_lower + lower, epsilon_trans);
        rrate += (R + C) * epsilon_target;
        if (zrand * pr_sel < rrate) break;
      }
      tl.ma_trans_scanned += 3 * nlevels;  // LUT rows read per lower level
      if (lower >= nlevels) {
        fatal(c, 226);
        p.type = ARTIS_TYPE_KPKT;
        return;
      }
      ui -= 1;
      gl = gl0_lower + lower;
      break;
    }
    case MA_INTERNALUPHIGHER: {  // macroatom.cc:382-414 do_macroatom_ionisation
      p.interactions += 1;
      tl.interactions++;
      stat_inc(tl, COUNTER_MA_STAT_INTERNALUPHIGHER);
      const double zrand = rng_uniform(rng);
      const int ntargets = c.st.lev_nphixstargets[gl];
      double rrate = 0.;
      int upper = -1;
      for (int t = 0; t < ntargets; t++) {
        upper = c.st.phixstarget_level[c.st.lev_phixstargetoffset[gl] + t];
        const double epsilon_trans = phixs_threshold(c, ui, gl, t);
        const double R = get_corrphotoioncoeff(c, mgi, gl, t);
        const double C = col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans);
        rrate += (R + C) * epsilon_current;
        if (zrand * pr_sel < rrate) break;
      }
      if (zrand * pr_sel >= rrate || upper < 0) {
        fatal(c, 228);
        p.type = ARTIS_TYPE_KPKT;
        return;
      }
      ui += 1;
      gl = c.st.ion_leveloffset[ui] + upper;
      break;
    }
    case MA_INTERNALUPHIGHERNT: {  // macroatom.cc:856-876: non-thermal ionisation to the ground level of a higher ion
      p.interactions += 1;
      tl.interactions++;
      const int ion = ui - c.st.el_ionoffset[element];
      ui += nt_random_upperion(c, mgi, ui, c.st.el_nions[element], ion, false, rng);
      gl = c.st.ion_leveloffset[ui];
      stat_inc(tl, COUNTER_MA_STAT_INTERNALUPHIGHERNT);
      break;
    }
    default:
      fatal(c, 229);
      p.type = ARTIS_TYPE_KPKT;
      return;
  }
  if (!end_packet) {  // jump to another ion: remember where the macro-atom is
    p.ma_ion = ui - c.st.el_ionoffset[element];
    p.ma_level = gl - c.st.ion_leveloffset[ui];
    return;
  }
  if (ldcold(&p.cold->trueemissiontype) < 0) {  // macroatom.cc:894-898
    p.cold->trueemissiontype = ldcold(&p.cold->emissiontype);
    double ep[3] = {ldcold(&p.cold->em_pos[0]), ldcold(&p.cold->em_pos[1]), ldcold(&p.cold->em_pos[2])};
    const int em_time = ldcold(&p.cold->em_time);
    p.cold->trueemissionvelocity = (float)(vec_len(ep) / em_time);
    p.cold->trueem_time = em_time;
  }
}
ARTIS_DEV_NOINLINE void ma_event_ool(const Ctx &c, EvState *s, const int action, const int selidx) { ma_event_body(c, s->p, action, selidx, s->rng, s->tl); }
ARTIS_DEV void ma_event(const Ctx &c, Pkt &p, const int action, const int selidx, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  ma_event_ool(c, &s, action, selidx);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  ma_event_body(c, p, action, selidx, rng, tl);
#endif
}
ARTIS_DEV void do_macroatom(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
  int selidx;
  const int action = ma_jump(c, p, rng, tl, &selidx);
  if (action >= 0) ma_event(c, p, action, selidx, rng, tl);
}

}  // namespace artis
