// celltables.cuh -- rate coefficients (macroatom.h:52-150, macroatom.cc:922-1195) and the per-cell tables that
// replace the reference's per-thread cellhistory cache (hard part 7.2-1).
//
// The reference fills its cache lazily for the one cell a host thread is working on (update_grid.cc:659,
// macroatom.cc:161, kpkt.cc:561). Here every non-empty model cell gets its rows once per timestep, right after
// upload_cellstate, by the cell_precompute kernels (transport.cu): one thread per (cell, level) or (cell, ion),
// summing in the reference's order so that each prefix sum is bit-identical to the running sum of the reference's
// linear selection loops.
#pragma once
#include "atomicdata.cuh"

namespace artis {

ARTIS_DEV int tabcell(const Ctx &c, int mgi) { return c.st.nonemptymgi_of_mgi[mgi]; }
ARTIS_DEV double tab_levelpop(const Ctx &c, int tc, int gl) { return c.tb.levelpop[(size_t)tc * c.st.nlevels_total + gl]; }
ARTIS_DEV const double *tab_ma_rates(const Ctx &c, int tc, int gl) {
  return c.tb.ma_rates + ((size_t)tc * c.st.nlevels_total + gl) * MA_ACTION_COUNT;
}

// the members of linelist_entry the collisional formulae read
struct LineRef {
  float osc_strength, coll_str;
  int forbidden;
};
ARTIS_DEV LineRef lineref(const Ctx &c, int lineindex) {
  LineRef l;
  l.osc_strength = c.st.line_osc_strength[lineindex];
  l.coll_str = c.st.line_coll_str[lineindex];
  l.forbidden = c.st.line_forbidden[lineindex];
  return l;
}

// macroatom.h:52-103 (multiply by the upper level population to get a rate)
ARTIS_DEV double col_deexcitation_ratecoeff(const float T_e, const float nne, const double epsilon_trans, const LineRef &line,
                                            const double lowerstatweight, const double upperstatweight) {
  double C;
  const double coll_str_thisline = line.coll_str;
  if (coll_str_thisline < 0) {
    if (!line.forbidden) {
      const double eoverkt = epsilon_trans / (KB * T_e);
      const double g_bar = 0.2;
      const double gauntfac = (eoverkt > 0.33421) ? g_bar : 0.276 * exp(eoverkt) * (-0.5772156649 - log(eoverkt));
      const double g_ratio = lowerstatweight / upperstatweight;
      C = C_0 * 14.51039491 * nne * sqrtf(T_e) * line.osc_strength * pow(H_ionpot / epsilon_trans, 2) * eoverkt * g_ratio *
          gauntfac;
    } else {
      C = nne * 8.629e-6 * 0.01 * lowerstatweight / sqrtf(T_e);
    }
  } else {
    C = nne * 8.629e-6 * coll_str_thisline / upperstatweight / sqrtf(T_e);
  }
  return C;
}
// macroatom.h:105-150 (multiply by the lower level population to get a rate)
ARTIS_DEV double col_excitation_ratecoeff(const float T_e, const float nne, const LineRef &line, const double epsilon_trans,
                                          const double lowerstatweight, const double upperstatweight) {
  double C;
  const double coll_strength = line.coll_str;
  const double eoverkt = epsilon_trans / (KB * T_e);
  if (coll_strength < 0) {
    if (!line.forbidden) {
      const double g_bar = 0.2;
      const double exp_eoverkt = exp(eoverkt);
      const double test = 0.276 * exp_eoverkt * (-0.5772156649 - log(eoverkt));
      const double Gamma = g_bar > test ? g_bar : test;
      C = C_0 * nne * sqrtf(T_e) * 14.51039491 * line.osc_strength * pow(H_ionpot / epsilon_trans, 2) * eoverkt /
          exp_eoverkt * Gamma;
    } else {
      C = nne * 8.629e-6 * 0.01 * exp(-eoverkt) * upperstatweight / sqrtf(T_e);
    }
  } else {
    C = nne * 8.629e-6 * coll_strength * exp(-eoverkt) / lowerstatweight / sqrtf(T_e);
  }
  return C;
}

// Sobolev optical depth pieces shared by macroatom.cc:922-967 and :969-1062
struct Sobolev {
  double A_ul, B_ul, B_lu, tau;
};
ARTIS_DEV Sobolev sobolev_of(const Ctx &c, const int lineindex, const double nu_trans, const double g_u, const double g_l,
                             const double n_u, const double n_l, const double t_current) {
  Sobolev s;
  s.A_ul = (double)c.st.line_einstein_A[lineindex];
  s.B_ul = CLIGHTSQUAREDOVERTWOH / pow(nu_trans, 3) * s.A_ul;
  s.B_lu = g_u / g_l * s.B_ul;
  s.tau = (s.B_lu * n_l - s.B_ul * n_u) * HCLIGHTOVERFOURPI * t_current;
  return s;
}
// macroatom.cc:922-967
ARTIS_DEV double rad_deexcitation_ratecoeff(const Ctx &c, const int tc, const int gl_upper, const int gl_lower,
                                            const double epsilon_trans, const int lineindex, const double t_current) {
  const double n_u = tab_levelpop(c, tc, gl_upper);
  const double n_l = tab_levelpop(c, tc, gl_lower);
  const Sobolev s = sobolev_of(c, lineindex, epsilon_trans / H, (double)c.st.lev_statweight[gl_upper],
                               (double)c.st.lev_statweight[gl_lower], n_u, n_l, t_current);
  if (s.tau > 1e-100) {
    const double beta = 1.0 / s.tau * (-expm1(-s.tau));
    return s.A_ul * beta;
  }
  return 0.;
}
// macroatom.cc:969-1062 (DETAILED_LINE_ESTIMATORS_ON is false in every shipped options header)
ARTIS_DEV double rad_excitation_ratecoeff(const Ctx &c, const int mgi, const int tc, const int gl_lower, const int gl_upper,
                                          const double epsilon_trans, const int lineindex, const double t_current) {
  const double n_u = tab_levelpop(c, tc, gl_upper);
  const double n_l = tab_levelpop(c, tc, gl_lower);
  const double nu_trans = epsilon_trans / H;
  const Sobolev s = sobolev_of(c, lineindex, nu_trans, (double)c.st.lev_statweight[gl_upper],
                               (double)c.st.lev_statweight[gl_lower], n_u, n_l, t_current);
  double R = 0.;
  if (s.tau > 1e-100) {
    const double beta = 1.0 / s.tau * (-expm1(-s.tau));
    const double R_over_J_nu = n_l > 0. ? (s.B_lu - s.B_ul * n_u / n_l) * beta : s.B_lu * beta;
    R = R_over_J_nu * radfield_J(c, nu_trans, mgi);
  }
  // the reference aborts on R < 0 or non-finite R when the level is visited; here a bad R poisons the level's rates
  // and do_macroatom reports it when (and only when) a packet reaches that level
  if (R < 0) R = NAN;
  return R;
}

// phixs target index t of (lower level gl_lower) whose upper level is `upper` (atomic.cc:23-31), -1 if none
ARTIS_DEV int phixstargetindex_of(const Ctx &c, const int gl_lower, const int upper) {
  const int n = c.st.lev_nphixstargets[gl_lower];
  const int off = c.st.lev_phixstargetoffset[gl_lower];
  for (int t = 0; t < n; t++)
    if (c.st.phixstarget_level[off + t] == upper) return t;
  return -1;
}
ARTIS_DEV double get_spontrecombcoeff(const Ctx &c, int gl_lower, int t, float T_e) {
  return lut_interp(c, c.st.spontrecombcoeff, gl_lower, t, (double)T_e, true);
}
ARTIS_DEV double get_bfcoolingcoeff(const Ctx &c, int gl_lower, int t, float T_e) {
  return lut_interp(c, c.st.bfcooling_coeff, gl_lower, t, (double)T_e, false);
}
// macroatom.cc:1064-1092 (SEPARATE_STIMRECOMB false). `upper` is the level index within the upper ion.
ARTIS_DEV double rad_recombination_ratecoeff(const Ctx &c, const float T_e, const float nne, const int gl_lower, const int upper) {
  const int t = phixstargetindex_of(c, gl_lower, upper);
  if (t < 0) return 0.;
  return nne * get_spontrecombcoeff(c, gl_lower, t, T_e);
}
ARTIS_DEV double phixs_at_edge(const Ctx &c, const int gl) {
  return (double)c.st.phixs_tables[(size_t)c.st.lev_phixstable[gl] * c.st.nphixspoints];
}
ARTIS_DEV double colion_gaunt(const int ionstage) { return ionstage == 1 ? 0.1 : (ionstage == 2 ? 0.2 : 0.3); }
// macroatom.cc:1123-1163; gl_upper = global index of the recombining level, ui_upper its ion
ARTIS_DEV double col_recombination_ratecoeff(const Ctx &c, const int mgi, const int ui_upper, const int gl_upper,
                                             const int gl_lower, const double epsilon_trans) {
  const int upper = gl_upper - c.st.ion_leveloffset[ui_upper];
  const int t = phixstargetindex_of(c, gl_lower, upper);
  if (t < 0) return 0.;
  const float nne = c.cs.nne[mgi];
  const float T_e = c.cs.Te[mgi];
  const double fac1 = epsilon_trans / KB / T_e;
  const double g = colion_gaunt(c.st.ion_ionstage[ui_upper] - 1);
  const double sigma_bf = phixs_at_edge(c, gl_lower) * get_phixsprobability_g(c, gl_lower, t);
  const double sf = calculate_sahafact_g(c, gl_lower, gl_upper, (double)T_e, epsilon_trans);
  return nne * nne * sf * 1.55e13 * pow((double)T_e, -0.5) * g * sigma_bf * exp(-fac1) / fac1;
}
// macroatom.cc:1165-1195
ARTIS_DEV double col_ionization_ratecoeff(const Ctx &c, const float T_e, const float nne, const int ui, const int gl_lower,
                                          const int t, const double epsilon_trans) {
  const double g = colion_gaunt(c.st.ion_ionstage[ui]);
  const double fac1 = epsilon_trans / KB / T_e;
  const double sigma_bf = phixs_at_edge(c, gl_lower) * get_phixsprobability_g(c, gl_lower, t);
  return nne * 1.55e13 * pow((double)T_e, -0.5) * g * sigma_bf * exp(-fac1) / fac1;
}
// ratecoeff.cc:1247-1310 get_corrphotoioncoeff. LUT mode (NO_LUT_PHOTOION false): W * interpolated coefficient *
// renormalisation. NO_LUT_PHOTOION: the per-cell table filled by build_corrphoto below.
ARTIS_DEV double get_corrphotoioncoeff(const Ctx &c, const int mgi, const int gl_lower, const int t) {
  if (ARTIS_OPT_NOLUT_PHOTOION(c)) {
    const int i = c.allcont_of_lut[-1 - c.st.lev_cont_index[gl_lower] + t];
    return c.tb.corrphoto[(size_t)tabcell(c, mgi) * c.st.nbfcontinua + i];
  }
  const double W = (double)c.cs.W[mgi];
  const double T_R = (double)c.cs.TR[mgi];
  double gammacorr = W * lut_interp(c, c.st.corrphotoioncoeff, gl_lower, t, T_R, false);
  const int idx = c.st.lev_closestgroundlevelcont[gl_lower];
  if (idx >= 0) gammacorr *= c.cs.corrphotoionrenorm[(size_t)mgi * c.st.nelements * c.st.max_nions + idx];
  return gammacorr;
}
// threshold energy of phixs target t of level gl (atomic.cc:437-454); ui = the level's ion
ARTIS_DEV double phixs_threshold(const Ctx &c, const int ui, const int gl, const int t) {
  const int upper = c.st.phixstarget_level[c.st.lev_phixstargetoffset[gl] + t];
  return c.st.lev_epsilon[c.st.ion_leveloffset[ui + 1] + upper] - c.st.lev_epsilon[gl];
}
ARTIS_DEV bool ion_has_lower(const Ctx &c, const int ui) { return ui > c.st.el_ionoffset[c.uion_element[ui]]; }
ARTIS_DEV bool ion_has_upper(const Ctx &c, const int ui) {
  const int el = c.uion_element[ui];
  return ui < c.st.el_ionoffset[el] + c.st.el_nions[el] - 1;
}

// ---- non-thermal ionisation: nonthermal.cc:1567-1760 over the host solver's Spencer-Fano numbers ----
// nonthermal.cc:1567-1581 nt_ionization_ratecoeff_sf and :1684-1717 nt_ionization_ratecoeff (ui = unique ion index)
ARTIS_DEV double nt_ionization_ratecoeff(const Ctx &c, const int mgi, const int ui) {
  const size_t k = (size_t)mgi * c.st.includedions + ui;
  const double Y_nt_wfapprox = c.cs.nt_ratecoeff_wfapprox[k];
  if (!c.st.opt_nt_solve_spencerfano) return Y_nt_wfapprox;
  const double deposition_rate_density = c.cs.nt_deposition_rate_density[mgi];
  double Y_nt = 0.;
  if (deposition_rate_density > 0.) Y_nt = deposition_rate_density / c.cs.nntot[mgi] / c.cs.nt_eff_ionpot[k];
  if (!isfinite(Y_nt) || Y_nt <= 0) return Y_nt_wfapprox;
  return Y_nt;
}
// nonthermal.cc:1643-1655 nt_ionisation_maxupperion (ion indices within the element; nions of that element)
ARTIS_DEV int nt_ionisation_maxupperion(const Ctx &c, const int nions, const int lowerion) {
  int maxupper = lowerion + 1;
  if (c.st.opt_nt_solve_spencerfano) maxupper = lowerion + 1 + c.st.opt_nt_max_auger_electrons;
  if (maxupper > nions - 1) maxupper = nions - 1;
  return maxupper;
}
// nonthermal.cc:1584-1641 nt_ionization_upperion_probability; ui_lower = unique index of the lower ion,
// nupper = upperion - lowerion
ARTIS_DEV double nt_ionization_upperion_probability(const Ctx &c, const int mgi, const int ui_lower, const int nupper,
                                                    const bool energyweighted) {
  const int maxauger = c.st.opt_nt_max_auger_electrons;
  if (c.st.opt_nt_solve_spencerfano && maxauger > 0) {
    const int numaugerelec = nupper - 1;
    const float *tab = (energyweighted ? c.cs.nt_ionenfrac_num_auger : c.cs.nt_prob_num_auger) +
                       ((size_t)mgi * c.st.includedions + ui_lower) * (maxauger + 1);
    if (numaugerelec < maxauger) return (double)tab[numaugerelec];
    if (numaugerelec == maxauger) {
      double prob_remaining = 1.;
      for (int a = 0; a < maxauger; a++) prob_remaining -= (double)tab[a];
      if (!(fabs(prob_remaining - (double)tab[numaugerelec]) < 0.001)) fatal(c, 260);  // assert / abort in the reference
      return prob_remaining;
    }
    return 0.;
  }
  return (nupper == 1) ? 1.0 : 0.;
}
// nonthermal.cc:1657-1682 nt_random_upperion; returns upperion - lowerion
ARTIS_DEV int nt_random_upperion(const Ctx &c, const int mgi, const int ui_lower, const int nions, const int lowerion,
                                 const bool energyweighted, Rng &rng) {
  if (c.st.opt_nt_solve_spencerfano && c.st.opt_nt_max_auger_electrons > 0) {
    for (int attempt = 0; attempt < 1000; attempt++) {
      const double zrand = rng_uniform(rng);
      double prob_sum = 0.;
      const int maxupper = nt_ionisation_maxupperion(c, nions, lowerion);
      for (int upperion = lowerion + 1; upperion <= maxupper; upperion++) {
        prob_sum += nt_ionization_upperion_probability(c, mgi, ui_lower, upperion - lowerion, energyweighted);
        if (zrand <= prob_sum) return upperion - lowerion;
      }
      if (!(fabs(prob_sum - 1.0) < 1e-3)) {  // assert_always in the reference
        fatal(c, 261);
        return 1;
      }
    }
    fatal(c, 262);
    return 1;
  }
  return 1;
}
// nonthermal.cc:1827-1845 ion_ntion_energyrate
ARTIS_DEV double ion_ntion_energyrate(const Ctx &c, const int mgi, const int element, const int ui_lower, const int nions,
                                      const int lowerion) {
  const double nnlowerion = ionstagepop_ui(c, mgi, element, ui_lower);
  double enrate = 0.;
  const int maxupper = nt_ionisation_maxupperion(c, nions, lowerion);
  const double eps_lower = c.st.lev_epsilon[c.st.ion_leveloffset[ui_lower]];
  for (int upperion = lowerion + 1; upperion <= maxupper; upperion++) {
    const double upperionprobfrac = nt_ionization_upperion_probability(c, mgi, ui_lower, upperion - lowerion, false);
    const double epsilon_trans = c.st.lev_epsilon[c.st.ion_leveloffset[ui_lower + (upperion - lowerion)]] - eps_lower;
    enrate += nnlowerion * upperionprobfrac * epsilon_trans;
  }
  const double gamma_nt = nt_ionization_ratecoeff(c, mgi, ui_lower);
  return gamma_nt * enrate;
}

// ---- table builders (called by the cell_precompute kernels) ------------------------------------------------------
ARTIS_DEV void build_levelpop(const Ctx &c, const int mgi, const int gl) {
  const int ui = c.lev_uion[gl];
  c.tb.levelpop[(size_t)tabcell(c, mgi) * c.st.nlevels_total + gl] = levelpop_g(c, mgi, c.uion_element[ui], ui, gl);
}

// rpkt.cc:1096-1145: which continua enter the bf opacity of this cell, and their departure ratios
ARTIS_DEV void build_bf(const Ctx &c, const int mgi, const int i) {
  const int tc = tabcell(c, mgi);
  const int element = c.st.cont_element[i];
  const int ui = c.cont_uion[i];
  bool present;
  if (c.st.opt_detailed_bf_estimators) {
    present = elem_abundance(c, mgi, element) > 0;
  } else {
    present = (ionstagepop_ui(c, mgi, element, ui) / (double)c.cs.nnetot[mgi] > 1.e-6) || (c.st.cont_level[i] == 0);
  }
  double nnlevel = -1., dep = 0.;
  if (present) {
    const int gl = c.cont_level_g[i], glu = c.cont_upper_g[i];
    nnlevel = tab_levelpop(c, tc, gl);
    if (nnlevel > 0) {
      const double nnupperionlevel = tab_levelpop(c, tc, glu);
      const double sf = calculate_sahafact_g(c, gl, glu, (double)c.cs.Te[mgi], H * c.st.cont_nu_edge[i]);
      dep = nnupperionlevel / nnlevel * (double)c.cs.nne[mgi] * sf;
    }
  }
  c.tb.bf_cell[2 * ((size_t)tc * c.st.nbfcontinua + i)] = nnlevel;
  c.tb.bf_cell[2 * ((size_t)tc * c.st.nbfcontinua + i) + 1] = dep;
}

// ratecoeff.cc:1160-1183 integrand_corrphotoioncoeff_custom_radfield
ARTIS_DEV double corrphoto_integrand(const Ctx &c, const int mgi, const int tablerow, const double nu_edge, const float T_e,
                                     const double departure_ratio, const double nu) {
  double corrfactor = 1. - departure_ratio * exp(-HOVERKB * nu / T_e);
  if (corrfactor < 0) corrfactor = 0.;
  const float sigma_bf = (float)phixs_fromtable(c, tablerow, nu_edge, nu);
  const double Jnu = radfield_J(c, nu, mgi);
  return ONEOVERH * sigma_bf / nu * Jnu * corrfactor;
}
// ratecoeff.cc:1184-1245 calculate_corrphotoioncoeff_integral. The reference integrates [nu_threshold, nu_max_phixs]
// with adaptive GSL qag (GK61, epsrel 1e-3). The integrand is smooth between the knots of the cross-section table and
// the boundaries of the radiation-field bins and jumps at the latter, so here every such piece gets a fixed 8-point
// Gauss-Legendre rule: deterministic, and closer to the exact integral than the reference's own tolerance.
// Stated tolerance against the reference: 5e-3 relative (tests/probelib.py, macro-atom rate probe).
// Returns the sum over the knot intervals first, first + stride, ... (a single thread passes 0, 1; the lanes of a warp
// pass lane, 32 and add their parts, transport.cu k_corrphoto); the caller multiplies by FOURPI * phixsprobability.
ARTIS_DEV double corrphotoioncoeff_integral(const Ctx &c, const int mgi, const int tc, const int ui, const int gl, const int t,
                                            const int first = 0, const int stride = 1) {
  const double x8[4] = {0.1834346424956498, 0.5255324099163290, 0.7966664774136267, 0.9602898564975363};
  const double w8[4] = {0.3626837833783620, 0.3137066458778873, 0.2223810344533745, 0.1012285362903763};
  const double E_threshold = phixs_threshold(c, ui, gl, t);
  const double nu_threshold = ONEOVERH * E_threshold;
  const double nu_max_phixs = nu_threshold * c.st.last_phixs_nuovernuedge;
  const float T_e = c.cs.Te[mgi];
  const double nnlevel = tab_levelpop(c, tc, gl);
  const double nne = (double)c.cs.nne[mgi];
  const int glu = c.st.ion_leveloffset[ui + 1] + c.st.phixstarget_level[c.st.lev_phixstargetoffset[gl] + t];
  const double sf = calculate_sahafact_g(c, gl, glu, (double)T_e, H * nu_threshold);
  const double nnupperionlevel = tab_levelpop(c, tc, glu);
  double departure_ratio = nnlevel > 0. ? nnupperionlevel / nnlevel * nne * sf : 1.0;
  if (!isfinite(departure_ratio)) departure_ratio = 0.;
  const int tablerow = c.st.lev_phixstable[gl];
  const bool binned = c.st.opt_multibin_radfield && c.nts >= c.st.opt_first_nlte_radfield_timestep;
  const int nknots = c.st.nphixspoints;
  double total = 0.;
  for (int k = first; k + 1 < nknots; k += stride) {
    double lo = nu_threshold * (1. + k * c.st.nphixsnuincrement);
    const double hi = (k + 2 == nknots) ? nu_max_phixs : nu_threshold * (1. + (k + 1) * c.st.nphixsnuincrement);
    int bin = binned ? radfield_select_bin(c, lo) : -1;
    if (binned && bin == -2) bin = -3;  // below the first bin: the next boundary is the lower edge of bin 0
    while (lo < hi) {
      double b = hi;
      if (binned) {
        // next bin boundary above lo, if it falls inside the piece
        double edge = -1.;
        if (bin == -3) {
          edge = c.st.radfield_nu_lower_first;
        } else if (bin >= 0) {
          edge = c.st.radfieldbin_nu_upper[bin];
        }
        if (edge > lo && edge < hi) b = edge;
      }
      const double h = 0.5 * (b - lo), m = 0.5 * (lo + b);
      double sum = 0.;
      for (int q = 0; q < 4; q++)
        sum += w8[q] * (corrphoto_integrand(c, mgi, tablerow, nu_threshold, T_e, departure_ratio, m - h * x8[q]) +
                        corrphoto_integrand(c, mgi, tablerow, nu_threshold, T_e, departure_ratio, m + h * x8[q]));
      total += sum * h;
      if (b < hi) {  // crossed a bin boundary
        if (bin == -3) {
          bin = 0;
        } else {
          bin = (bin + 1 < c.st.opt_radfieldbincount) ? bin + 1 : -1;
        }
      }
      lo = b;
    }
  }
  return total;
}
// ratecoeff.cc:1247-1310 with NO_LUT_PHOTOION, for allcont[i] of one cell: the bf-rate estimator of the previous
// timestep where DETAILED_BF_ESTIMATORS_ON has one (radfield.cc:1440-1457), else the integral above
// `first`/`stride` as in corrphotoioncoeff_integral; *needs_sum is set when the returned value is a partial integral
// that the caller has to add up over its threads and scale with corrphoto_scale()
ARTIS_DEV double build_corrphoto_part(const Ctx &c, const int mgi, const int i, const int first, const int stride,
                                      bool *needs_sum) {
  const int tc = tabcell(c, mgi);
  double gammacorr = -1.;
  if (c.st.opt_detailed_bf_estimators && c.nts >= c.st.opt_detailed_bf_usefromtimestep)
    gammacorr = (double)c.cs.prev_bfrate_normed[(size_t)tc * c.st.nbfcontinua + i];
  *needs_sum = !(gammacorr > 0);
  if (!(gammacorr > 0)) {
    gammacorr = corrphotoioncoeff_integral(c, mgi, tc, c.cont_uion[i], c.cont_level_g[i], c.st.cont_phixstargetindex[i],
                                           first, stride);
  }
  return gammacorr;
}
ARTIS_DEV double corrphoto_scale(const Ctx &c, const int i) {
  return FOURPI * get_phixsprobability_g(c, c.cont_level_g[i], c.st.cont_phixstargetindex[i]);
}
ARTIS_DEV void build_corrphoto(const Ctx &c, const int mgi, const int i) {
  bool needs_sum;
  double gammacorr = build_corrphoto_part(c, mgi, i, 0, 1, &needs_sum);
  if (needs_sum) gammacorr *= corrphoto_scale(c, i);
  c.tb.corrphoto[(size_t)tabcell(c, mgi) * c.st.nbfcontinua + i] = gammacorr;
}

// macroatom.cc:57-159 calculate_macroatom_transitionrates for one (cell, level)
ARTIS_DEV void build_ma_level_into(const Ctx &c, const int mgi, const int gl, const double t_mid, double *out,
                                   double *cum_rad, double *cum_down, double *cum_up) {
  const int tc = tabcell(c, mgi);
  const int ui = c.lev_uion[gl];
  const int gl0 = c.st.ion_leveloffset[ui];
  const int level = gl - gl0;
  const float T_e = c.cs.Te[mgi];
  const float nne = c.cs.nne[mgi];
  const double epsilon_current = c.st.lev_epsilon[gl];
  const double statweight = (double)c.st.lev_statweight[gl];
  double pr[MA_ACTION_COUNT];
  for (int a = 0; a < MA_ACTION_COUNT; a++) pr[a] = 0.;

  const int ndown = c.st.lev_ndowntrans[gl], doff = c.st.lev_downoffset[gl];
  for (int i = 0; i < ndown; i++) {
    const int lineindex = c.st.downtrans_lineindex[doff + i];
    const int gll = c.line_lower_g[lineindex];
    const double epsilon_target = c.st.lev_epsilon[gll];
    const double epsilon_trans = epsilon_current - epsilon_target;
    const double R = rad_deexcitation_ratecoeff(c, tc, gl, gll, epsilon_trans, lineindex, t_mid);
    const double C = col_deexcitation_ratecoeff(T_e, nne, epsilon_trans, lineref(c, lineindex),
                                                (double)c.st.lev_statweight[gll], statweight);
    pr[MA_RADDEEXC] = fma(R, epsilon_trans, pr[MA_RADDEEXC]);
    pr[MA_COLDEEXC] = fma(C, epsilon_trans, pr[MA_COLDEEXC]);
    pr[MA_INTERNALDOWNSAME] = fma(R + C, epsilon_target, pr[MA_INTERNALDOWNSAME]);
    cum_rad[i] = pr[MA_RADDEEXC];
    cum_down[i] = pr[MA_INTERNALDOWNSAME];
  }

  if (ion_has_lower(c, ui) && level <= c.st.ion_maxrecombininglevel[ui]) {
    const int gl0_lower = c.st.ion_leveloffset[ui - 1];
    const int nlevels = c.st.ion_ionisinglevels[ui - 1];
    for (int lower = 0; lower < nlevels; lower++) {
      const int gll = gl0_lower + lower;
      const double epsilon_target = c.st.lev_epsilon[gll];
      const double epsilon_trans = epsilon_current - epsilon_target;
      const double R = rad_recombination_ratecoeff(c, T_e, nne, gll, level);
      const double C = col_recombination_ratecoeff(c, mgi, ui, gl, gll, epsilon_trans);
      pr[MA_INTERNALDOWNLOWER] = fma(R + C, epsilon_target, pr[MA_INTERNALDOWNLOWER]);
      pr[MA_RADRECOMB] = fma(R, epsilon_trans, pr[MA_RADRECOMB]);
      pr[MA_COLRECOMB] = fma(C, epsilon_trans, pr[MA_COLRECOMB]);
    }
  }

  const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
  for (int i = 0; i < nup; i++) {
    const int lineindex = c.st.uptrans_lineindex[uoff + i];
    const int glu = c.line_upper_g[lineindex];
    const double epsilon_trans = c.st.lev_epsilon[glu] - epsilon_current;
    const double R = rad_excitation_ratecoeff(c, mgi, tc, gl, glu, epsilon_trans, lineindex, t_mid);
    const double C = col_excitation_ratecoeff(T_e, nne, lineref(c, lineindex), epsilon_trans, statweight,
                                              (double)c.st.lev_statweight[glu]);
    const double NT = 0.;  // nonthermal::nt_excitation_ratecoeff: NT_EXCITATION_ON is false in every options header
    pr[MA_INTERNALUPSAME] = fma(R + C + NT, epsilon_current, pr[MA_INTERNALUPSAME]);
    cum_up[i] = pr[MA_INTERNALUPSAME];
  }

  if (ion_has_upper(c, ui) && level < c.st.ion_ionisinglevels[ui]) {
    if (ARTIS_OPT_NT(c)) pr[MA_INTERNALUPHIGHERNT] = nt_ionization_ratecoeff(c, mgi, ui) * epsilon_current;  // macroatom.cc:143-146
    const int ntargets = c.st.lev_nphixstargets[gl];
    for (int t = 0; t < ntargets; t++) {
      const double epsilon_trans = phixs_threshold(c, ui, gl, t);
      const double R = get_corrphotoioncoeff(c, mgi, gl, t);
      const double C = col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans);
      pr[MA_INTERNALUPHIGHER] = fma(R + C, epsilon_current, pr[MA_INTERNALUPHIGHER]);
    }
  }
  for (int a = 0; a < MA_ACTION_COUNT; a++) out[a] = pr[a];
}
// the dense tables (one row per level of every cell, built by the cell_precompute kernels)
ARTIS_DEV void build_ma_level(const Ctx &c, const int mgi, const int gl, const double t_mid) {
  const size_t tc = (size_t)tabcell(c, mgi);
  const int doff = c.st.lev_downoffset[gl], uoff = c.st.lev_upoffset[gl];
  build_ma_level_into(c, mgi, gl, t_mid, c.tb.ma_rates + (tc * c.st.nlevels_total + gl) * MA_ACTION_COUNT,
                      c.tb.ma_cum_raddeexc + tc * c.st.ndowntrans_total + doff,
                      c.tb.ma_cum_intdown + tc * c.st.ndowntrans_total + doff, c.tb.ma_cum_intup + tc * c.st.nuptrans_total + uoff);
}

// kpkt.cc:167-308 calculate_kpkt_rates_ion for one (cell, ion); `oldcoolingsum` = sum of cooling_contrib_ion of the
// ions before this one in (element, ion) order, accumulated like the selection loop of do_kpkt (kpkt.cc:527-536)
// `cool` is indexed with the global cooling-term index, `cexc` with the global upward-transition index.
// STORE = false, start from zero: calculate_cooling_rates' C_ion (kpkt.cc:310-365, the same template with
// update_cooling_contrib_list = false), returned.
template <bool STORE>
ARTIS_DEV double cooling_ion_sum(const Ctx &c, const int mgi, const int ui, double *cool, double *cexc) {
  const int tc = tabcell(c, mgi);
  const int element = c.uion_element[ui];
  const float nne = c.cs.nne[mgi];
  const float T_e = c.cs.Te[mgi];
  double oldcoolingsum = 0.;
  if (STORE)
    for (int u = 0; u < ui; u++) oldcoolingsum += c.cs.cooling_contrib_ion[(size_t)mgi * c.st.includedions + u];
  double contrib = oldcoolingsum;
  int i = c.st.ion_coolingoffset[ui];
  const int gl0 = c.st.ion_leveloffset[ui];
  const int nlevels = c.st.ion_nlevels[ui];
  const int ionisinglevels = c.st.ion_ionisinglevels[ui];
  const double nncurrention = ionstagepop_ui(c, mgi, element, ui);
  const int ioncharge = c.st.ion_ionstage[ui] - 1;
  if (ioncharge > 0) {
    contrib += 1.426e-27 * sqrtf(T_e) * pow((double)ioncharge, 2) * nncurrention * nne;
    if (STORE) cool[i] = contrib;
    i++;
  }
  const bool hasupper = ion_has_upper(c, ui);
  for (int level = 0; level < nlevels; level++) {
    const int gl = gl0 + level;
    const double epsilon_current = c.st.lev_epsilon[gl];
    const double nnlevel = tab_levelpop(c, tc, gl);
    const double statweight = (double)c.st.lev_statweight[gl];
    const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
    if (nup > 0) {
      for (int ii = 0; ii < nup; ii++) {
        const int lineindex = c.st.uptrans_lineindex[uoff + ii];
        const int glu = c.line_upper_g[lineindex];
        const double epsilon_trans = c.st.lev_epsilon[glu] - epsilon_current;
        const double C = nnlevel *
                         col_excitation_ratecoeff(T_e, nne, lineref(c, lineindex), epsilon_trans, statweight,
                                                  (double)c.st.lev_statweight[glu]) *
                         epsilon_trans;
        contrib += C;
        if (STORE) cexc[uoff + ii] = contrib;
      }
      if (STORE) cool[i] = contrib;
      i++;
    }
    if (hasupper && level < ionisinglevels) {
      const int ntargets = c.st.lev_nphixstargets[gl];
      for (int t = 0; t < ntargets; t++) {
        const double epsilon_trans = phixs_threshold(c, ui, gl, t);
        contrib += nnlevel * col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans) * epsilon_trans;
        if (STORE) cool[i] = contrib;
        i++;
      }
      for (int t = 0; t < ntargets; t++) {
        const double nnupperion = ionstagepop_ui(c, mgi, element, ui + 1);
        contrib += get_bfcoolingcoeff(c, gl, t, T_e) * nnupperion * nne;
        if (STORE) cool[i] = contrib;
        i++;
      }
    }
  }
  // fill the tail when levels >= ionisinglevels own (uncomputed) list slots: the reference leaves them untouched
  // and asserts i == offset + ncoolingterms (kpkt.cc:300); here a mismatch is reported as a fatal table error
  if (i != c.st.ion_coolingoffset[ui] + c.st.ion_ncoolingterms[ui]) fatal(c, 202);
  return contrib;
}
ARTIS_DEV void build_cooling_ion_into(const Ctx &c, const int mgi, const int ui, double *cool, double *cexc) {
  cooling_ion_sum<true>(c, mgi, ui, cool, cexc);
}
// kpkt.cc:84-165 calculate_cooling_rates for one (cell, ion): C_ion in the reference's own order of addition -- the
// free-free term, then the collisional-excitation sum of every level (get_cooling_ion_coll_exc, kpkt.cc:40-66, summed on
// its own), then level by level the collisional ionisations and the free-bound terms
ARTIS_DEV double cooling_rates_ion(const Ctx &c, const int mgi, const int ui) {
  const int tc = tabcell(c, mgi);
  const int element = c.uion_element[ui];
  const float nne = c.cs.nne[mgi];
  const float T_e = c.cs.Te[mgi];
  const int gl0 = c.st.ion_leveloffset[ui];
  const int nlevels = c.st.ion_nlevels[ui];
  const double nncurrention = ionstagepop_ui(c, mgi, element, ui);
  double C_ion = 0.;
  const int ioncharge = c.st.ion_ionstage[ui] - 1;
  if (ioncharge > 0) C_ion += 1.426e-27 * sqrtf(T_e) * pow((double)ioncharge, 2) * nncurrention * nne;
  double C_exc = 0.;
  for (int level = 0; level < nlevels; level++) {
    const int gl = gl0 + level;
    const double nnlevel = tab_levelpop(c, tc, gl);
    const double epsilon_current = c.st.lev_epsilon[gl];
    const double statweight = (double)c.st.lev_statweight[gl];
    const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
    for (int ii = 0; ii < nup; ii++) {
      const int lineindex = c.st.uptrans_lineindex[uoff + ii];
      const int glu = c.line_upper_g[lineindex];
      const double epsilon_trans = c.st.lev_epsilon[glu] - epsilon_current;
      C_exc += nnlevel *
               col_excitation_ratecoeff(T_e, nne, lineref(c, lineindex), epsilon_trans, statweight, (double)c.st.lev_statweight[glu]) *
               epsilon_trans;
    }
  }
  C_ion += C_exc;
  if (ion_has_upper(c, ui)) {
    const int ionisinglevels = c.st.ion_ionisinglevels[ui];
    for (int level = 0; level < ionisinglevels; level++) {
      const int gl = gl0 + level;
      const double nnlevel = tab_levelpop(c, tc, gl);
      const int ntargets = c.st.lev_nphixstargets[gl];
      for (int t = 0; t < ntargets; t++) {
        const double epsilon_trans = phixs_threshold(c, ui, gl, t);
        C_ion += nnlevel * col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans) * epsilon_trans;
      }
      for (int t = 0; t < ntargets; t++) {
        const double nnupperion = ionstagepop_ui(c, mgi, element, ui + 1);
        C_ion += get_bfcoolingcoeff(c, gl, t, T_e) * nnupperion * nne;
      }
    }
  }
  return C_ion;
}
ARTIS_DEV void build_cooling_ion(const Ctx &c, const int mgi, const int ui) {
  const size_t tc = (size_t)tabcell(c, mgi);
  build_cooling_ion_into(c, mgi, ui, c.tb.cool_cum + tc * c.st.ncoolingterms, c.tb.collexc_cum + tc * c.st.nuptrans_total);
}

// ---- rows on demand ----------------------------------------------------------------------------------------------------
#ifndef ARTIS_ROWS_ON_DEMAND
#define ARTIS_ROWS_ON_DEMAND 1   // 0: the accessors below read the dense tables only (A/B builds)
#endif
ARTIS_DEV bool rows_on_demand(const Ctx &c) { return ARTIS_ROWS_ON_DEMAND != 0 && c.tb.row_ptr != nullptr; }
// Upward transitions of an ion's levels are one contiguous range of the uptrans arrays (checked at init, engine.inc).
ARTIS_DEV int ion_upfirst(const Ctx &c, const int ui) { return c.st.lev_upoffset[c.st.ion_leveloffset[ui]]; }
ARTIS_DEV int ion_upcount(const Ctx &c, const int ui) {
  const int last = c.st.ion_leveloffset[ui] + c.st.ion_nlevels[ui] - 1;
  return c.st.lev_upoffset[last] + c.st.lev_nuptrans[last] - ion_upfirst(c, ui);
}
// doubles in the row of level gl: process rates | running rad. de-excitation | running internal down | running internal up
ARTIS_DEV int ma_row_doubles(const Ctx &c, const int gl) {
  return MA_ACTION_COUNT + 2 * c.st.lev_ndowntrans[gl] + c.st.lev_nuptrans[gl];
}
ARTIS_DEV int cool_row_doubles(const Ctx &c, const int ui) { return c.st.ion_ncoolingterms[ui] + ion_upcount(c, ui); }

// The row `key` (level gl, or nlevels_total + ion) of table cell tc: where it is if some event has built it already, else
// built here by the calling thread and published with one compare-and-swap. Two threads that miss at the same time both
// build (the values are identical: same inputs, same arithmetic) and the loser's copy is wasted arena space; nobody ever
// waits for anybody. Out of line: the builders are the cell_precompute code, not the event loop's.
ARTIS_DEV_NOINLINE const double *row_get(const Ctx &c, const int mgi, const int tc, const int key) {
  const CellTables &tb = c.tb;
  unsigned *slot = tb.row_ptr + (size_t)tc * (c.st.nlevels_total + c.st.includedions) + key;
#ifdef __CUDA_ARCH__
  unsigned have = __ldcg(slot);
#else
  unsigned have = *slot;
#endif
  if (have != 0u) return tb.row_arena + (size_t)(have - 1u) * 16;
  const bool is_level = key < c.st.nlevels_total;
  const int ui = key - c.st.nlevels_total;
  const int ndoubles = is_level ? ma_row_doubles(c, key) : cool_row_doubles(c, ui);
  const unsigned long long units = (unsigned long long)(ndoubles + 15) / 16;
#ifdef __CUDA_ARCH__
  const unsigned long long at = atomicAdd(tb.row_top, units);
#else
  // host build (one thread, and no caller holds a row while it asks for another): a full arena is started afresh on the spot
  if (*tb.row_top + units > tb.row_soft && *tb.row_top > 0) {
    const size_t slots = (size_t)tb.ncells * (c.st.nlevels_total + c.st.includedions);
    for (size_t i = 0; i < slots; i++) tb.row_ptr[i] = 0u;
    *tb.row_top = 0;
    *tb.row_stop += 1;  // counted by the engine
    slot = tb.row_ptr + (size_t)tc * (c.st.nlevels_total + c.st.includedions) + key;
  }
  const unsigned long long at = *tb.row_top;
  *tb.row_top += units;
#endif
#ifdef __CUDA_ARCH__
  if (at + units > tb.row_soft) *reinterpret_cast<volatile int *>(tb.row_stop) = 1;
#endif
  if (at + units > tb.row_cap) {  // the reserve above row_soft is sized so that this cannot happen (engine.inc)
    fatal(c, 240);
    return tb.row_arena;
  }
  double *row = tb.row_arena + (size_t)at * 16;
  if (is_level) {
    const int nd = c.st.lev_ndowntrans[key];
    build_ma_level_into(c, mgi, key, c.t_mid, row, row + MA_ACTION_COUNT, row + MA_ACTION_COUNT + nd,
                        row + MA_ACTION_COUNT + 2 * nd);
  } else {
    const int ncool = c.st.ion_ncoolingterms[ui];
    build_cooling_ion_into(c, mgi, ui, row - c.st.ion_coolingoffset[ui], row + ncool - ion_upfirst(c, ui));
  }
#ifdef __CUDA_ARCH__
  __threadfence();
  const unsigned prev = atomicCAS(slot, 0u, (unsigned)at + 1u);
  if (prev != 0u) return tb.row_arena + (size_t)(prev - 1u) * 16;
#else
  *slot = (unsigned)at + 1u;
#endif
  return row;
}

#ifdef __CUDACC__
// The row of level gl built by the 32 lanes of a warp (all of them call, with the same arguments). The lanes evaluate the
// rate coefficients of 32 transitions at a time -- the Sobolev escape probabilities, collision strengths and recombination
// coefficients are the work -- and then every lane adds the 32 contributions in list order from shuffled operands, so that
// each running sum sees the roundings of build_ma_level_into (one fused multiply-add per transition, in order): the row is
// bit for bit the row one thread builds, in a thirtieth of the time. Returns the row (also when another warp was faster).
// The arithmetic of build_ma_level_into by the 32 lanes of a warp: a lane per transition evaluates the rate coefficients (the
// Sobolev escape probabilities, collision strengths and recombination coefficients are the work), and every lane then adds the
// 32 contributions in list order from shuffled operands with the fused multiply-adds of the one-thread builder: the same bits.
// Writes the nine process rates to row[0..8] and the three running sums to cum_rad / cum_down / cum_up.
__device__ __forceinline__ void coop_build_ma_into(const Ctx &c, const int mgi, const int tc, const int gl, const int lane, double *row,
                                                   double *cum_rad, double *cum_down, double *cum_up) {
  const unsigned full = 0xffffffffu;
  const int ndown = c.st.lev_ndowntrans[gl], doff = c.st.lev_downoffset[gl];
  const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
  const int ui = c.lev_uion[gl];
  const int level = gl - c.st.ion_leveloffset[ui];
  const float T_e = c.cs.Te[mgi];
  const float nne = c.cs.nne[mgi];
  const double epsilon_current = c.st.lev_epsilon[gl];
  const double statweight = (double)c.st.lev_statweight[gl];
  const double t_mid = c.t_mid;
  double pr[MA_ACTION_COUNT];
  for (int a = 0; a < MA_ACTION_COUNT; a++) pr[a] = 0.;

  for (int base = 0; base < ndown; base += 32) {
    const int i = base + lane;
    double R = 0., C = 0., e_trans = 0., e_target = 0.;
    if (i < ndown) {
      const int lineindex = c.st.downtrans_lineindex[doff + i];
      const int gll = c.line_lower_g[lineindex];
      e_target = c.st.lev_epsilon[gll];
      e_trans = epsilon_current - e_target;
      R = rad_deexcitation_ratecoeff(c, tc, gl, gll, e_trans, lineindex, t_mid);
      C = col_deexcitation_ratecoeff(T_e, nne, e_trans, lineref(c, lineindex), (double)c.st.lev_statweight[gll], statweight);
    }
    const int cnt = ndown - base < 32 ? ndown - base : 32;
    double my_rad = 0., my_down = 0.;
    for (int j = 0; j < cnt; j++) {
      const double Rj = __shfl_sync(full, R, j), Cj = __shfl_sync(full, C, j);
      const double etj = __shfl_sync(full, e_trans, j), egj = __shfl_sync(full, e_target, j);
      pr[MA_RADDEEXC] = fma(Rj, etj, pr[MA_RADDEEXC]);
      pr[MA_COLDEEXC] = fma(Cj, etj, pr[MA_COLDEEXC]);
      pr[MA_INTERNALDOWNSAME] = fma(Rj + Cj, egj, pr[MA_INTERNALDOWNSAME]);
      if (lane == j) {
        my_rad = pr[MA_RADDEEXC];
        my_down = pr[MA_INTERNALDOWNSAME];
      }
    }
    if (i < ndown) {
      cum_rad[i] = my_rad;
      cum_down[i] = my_down;
    }
  }

  if (ion_has_lower(c, ui) && level <= c.st.ion_maxrecombininglevel[ui]) {
    const int gl0_lower = c.st.ion_leveloffset[ui - 1];
    const int nlevels = c.st.ion_ionisinglevels[ui - 1];
    for (int base = 0; base < nlevels; base += 32) {
      const int lower = base + lane;
      double R = 0., C = 0., e_trans = 0., e_target = 0.;
      if (lower < nlevels) {
        const int gll = gl0_lower + lower;
        e_target = c.st.lev_epsilon[gll];
        e_trans = epsilon_current - e_target;
        R = rad_recombination_ratecoeff(c, T_e, nne, gll, level);
        C = col_recombination_ratecoeff(c, mgi, ui, gl, gll, e_trans);
      }
      const int cnt = nlevels - base < 32 ? nlevels - base : 32;
      for (int j = 0; j < cnt; j++) {
        const double Rj = __shfl_sync(full, R, j), Cj = __shfl_sync(full, C, j);
        const double etj = __shfl_sync(full, e_trans, j), egj = __shfl_sync(full, e_target, j);
        pr[MA_INTERNALDOWNLOWER] = fma(Rj + Cj, egj, pr[MA_INTERNALDOWNLOWER]);
        pr[MA_RADRECOMB] = fma(Rj, etj, pr[MA_RADRECOMB]);
        pr[MA_COLRECOMB] = fma(Cj, etj, pr[MA_COLRECOMB]);
      }
    }
  }

  for (int base = 0; base < nup; base += 32) {
    const int i = base + lane;
    double RC = 0.;
    if (i < nup) {
      const int lineindex = c.st.uptrans_lineindex[uoff + i];
      const int glu = c.line_upper_g[lineindex];
      const double e_trans = c.st.lev_epsilon[glu] - epsilon_current;
      const double R = rad_excitation_ratecoeff(c, mgi, tc, gl, glu, e_trans, lineindex, t_mid);
      const double C = col_excitation_ratecoeff(T_e, nne, lineref(c, lineindex), e_trans, statweight, (double)c.st.lev_statweight[glu]);
      const double NT = 0.;
      RC = R + C + NT;
    }
    const int cnt = nup - base < 32 ? nup - base : 32;
    double my_up = 0.;
    for (int j = 0; j < cnt; j++) {
      pr[MA_INTERNALUPSAME] = fma(__shfl_sync(full, RC, j), epsilon_current, pr[MA_INTERNALUPSAME]);
      if (lane == j) my_up = pr[MA_INTERNALUPSAME];
    }
    if (i < nup) cum_up[i] = my_up;
  }

  if (ion_has_upper(c, ui) && level < c.st.ion_ionisinglevels[ui]) {  // a handful of targets: every lane the same sums
    if (ARTIS_OPT_NT(c)) pr[MA_INTERNALUPHIGHERNT] = nt_ionization_ratecoeff(c, mgi, ui) * epsilon_current;
    const int ntargets = c.st.lev_nphixstargets[gl];
    for (int t = 0; t < ntargets; t++) {
      const double e_trans = phixs_threshold(c, ui, gl, t);
      const double R = get_corrphotoioncoeff(c, mgi, gl, t);
      const double C = col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, e_trans);
      pr[MA_INTERNALUPHIGHER] = fma(R + C, epsilon_current, pr[MA_INTERNALUPHIGHER]);
    }
  }
  if (lane < MA_ACTION_COUNT) {
    double v = pr[0];
#pragma unroll
    for (int a = 1; a < MA_ACTION_COUNT; a++)
      if (lane == a) v = pr[a];
    row[lane] = v;
  }
}

// cooling_ion_sum<true> by the 32 lanes of a warp: the collisional-excitation terms -- one per upward transition of every level
// of the ion, thousands for an iron-group ion -- are evaluated a lane per transition and added in list order from shuffled
// operands (every lane keeps the same running sum, the lane of a transition stores its value); the handful of other terms
// per level are evaluated by every lane alike and stored by lane 0. Same additions in the same order as the one-thread loop.
__device__ __forceinline__ void coop_cooling_ion_into(const Ctx &c, const int mgi, const int ui, const int lane, double *cool,
                                                      double *cexc) {
  const unsigned full = 0xffffffffu;
  const int tc = tabcell(c, mgi);
  const int element = c.uion_element[ui];
  const float nne = c.cs.nne[mgi];
  const float T_e = c.cs.Te[mgi];
  double contrib = 0.;
  for (int u = 0; u < ui; u++) contrib += c.cs.cooling_contrib_ion[(size_t)mgi * c.st.includedions + u];
  int i = c.st.ion_coolingoffset[ui];
  const int gl0 = c.st.ion_leveloffset[ui];
  const int nlevels = c.st.ion_nlevels[ui];
  const int ionisinglevels = c.st.ion_ionisinglevels[ui];
  const double nncurrention = ionstagepop_ui(c, mgi, element, ui);
  const int ioncharge = c.st.ion_ionstage[ui] - 1;
  if (ioncharge > 0) {
    contrib += 1.426e-27 * sqrtf(T_e) * pow((double)ioncharge, 2) * nncurrention * nne;
    if (lane == 0) cool[i] = contrib;
    i++;
  }
  const bool hasupper = ion_has_upper(c, ui);
  for (int level = 0; level < nlevels; level++) {
    const int gl = gl0 + level;
    const double epsilon_current = c.st.lev_epsilon[gl];
    const double nnlevel = tab_levelpop(c, tc, gl);
    const double statweight = (double)c.st.lev_statweight[gl];
    const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
    if (nup > 0) {
      for (int base = 0; base < nup; base += 32) {
        const int ii = base + lane;
        double C = 0.;
        if (ii < nup) {
          const int lineindex = c.st.uptrans_lineindex[uoff + ii];
          const int glu = c.line_upper_g[lineindex];
          const double epsilon_trans = c.st.lev_epsilon[glu] - epsilon_current;
          C = nnlevel *
              col_excitation_ratecoeff(T_e, nne, lineref(c, lineindex), epsilon_trans, statweight, (double)c.st.lev_statweight[glu]) *
              epsilon_trans;
        }
        const int cnt = nup - base < 32 ? nup - base : 32;
        double mine = 0.;
        for (int j = 0; j < cnt; j++) {
          contrib += __shfl_sync(full, C, j);
          if (lane == j) mine = contrib;
        }
        if (ii < nup) cexc[uoff + ii] = mine;
      }
      if (lane == 0) cool[i] = contrib;
      i++;
    }
    if (hasupper && level < ionisinglevels) {
      const int ntargets = c.st.lev_nphixstargets[gl];
      for (int t = 0; t < ntargets; t++) {
        const double epsilon_trans = phixs_threshold(c, ui, gl, t);
        contrib += nnlevel * col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans) * epsilon_trans;
        if (lane == 0) cool[i] = contrib;
        i++;
      }
      for (int t = 0; t < ntargets; t++) {
        const double nnupperion = ionstagepop_ui(c, mgi, element, ui + 1);
        contrib += get_bfcoolingcoeff(c, gl, t, T_e) * nnupperion * nne;
        if (lane == 0) cool[i] = contrib;
        i++;
      }
    }
  }
  if (i != c.st.ion_coolingoffset[ui] + c.st.ion_ncoolingterms[ui]) fatal(c, 202);
}

// rows on demand: the cooling row of (cell, ion) built by the whole warp (the event role's k-packets; transport.cu). Same
// protocol as coop_build_ma_row below: space from the arena, the row written, its pointer published with a compare-and-swap.
__device__ __noinline__ void coop_build_cool_row(const Ctx &c, const int mgi, const int tc, const int ui, const int lane) {
  const unsigned full = 0xffffffffu;
  const CellTables &tb = c.tb;
  unsigned *slot = tb.row_ptr + (size_t)tc * (c.st.nlevels_total + c.st.includedions) + c.st.nlevels_total + ui;
  unsigned long long at = 0ull;
  unsigned have = 0u;
  int bad = 0;
  if (lane == 0) {
    have = __ldcg(slot);
    if (have == 0u) {
      const unsigned long long units = (unsigned long long)(cool_row_doubles(c, ui) + 15) / 16;
      at = atomicAdd(tb.row_top, units);
      if (at + units > tb.row_soft) *reinterpret_cast<volatile int *>(tb.row_stop) = 1;
      if (at + units > tb.row_cap) {
        fatal(c, 240);
        bad = 1;
      }
    }
  }
  have = __shfl_sync(full, have, 0);
  if (have != 0u || __shfl_sync(full, bad, 0) != 0) return;
  at = __shfl_sync(full, at, 0);
  double *row = tb.row_arena + (size_t)at * 16;
  coop_cooling_ion_into(c, mgi, ui, lane, row - c.st.ion_coolingoffset[ui], row + c.st.ion_ncoolingterms[ui] - ion_upfirst(c, ui));
  __threadfence();
  __syncwarp(full);
  if (lane == 0) atomicCAS(slot, 0u, (unsigned)at + 1u);
}

__device__ __noinline__ const double *coop_build_ma_row(const Ctx &c, const int mgi, const int tc, const int gl, const int lane) {
  const unsigned full = 0xffffffffu;
  const CellTables &tb = c.tb;
  unsigned *slot = tb.row_ptr + (size_t)tc * (c.st.nlevels_total + c.st.includedions) + gl;
  const int ndown = c.st.lev_ndowntrans[gl], doff = c.st.lev_downoffset[gl];
  const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
  unsigned long long at = 0ull;
  unsigned have = 0u;
  int bad = 0;
  if (lane == 0) {
    have = __ldcg(slot);
    if (have == 0u) {
      const unsigned long long units = (unsigned long long)(MA_ACTION_COUNT + 2 * ndown + nup + 15) / 16;
      at = atomicAdd(tb.row_top, units);
      if (at + units > tb.row_soft) *reinterpret_cast<volatile int *>(tb.row_stop) = 1;
      if (at + units > tb.row_cap) {
        fatal(c, 240);
        bad = 1;
      }
    }
  }
  have = __shfl_sync(full, have, 0);
  if (have != 0u) return tb.row_arena + (size_t)(have - 1u) * 16;
  if (__shfl_sync(full, bad, 0) != 0) return tb.row_arena;
  at = __shfl_sync(full, at, 0);
  double *row = tb.row_arena + (size_t)at * 16;
  double *cum_rad = row + MA_ACTION_COUNT, *cum_down = cum_rad + ndown, *cum_up = cum_down + ndown;
  coop_build_ma_into(c, mgi, tc, gl, lane, row, cum_rad, cum_down, cum_up);
  __threadfence();
  __syncwarp(full);
  unsigned prev = 0u;
  if (lane == 0) prev = atomicCAS(slot, 0u, (unsigned)at + 1u);
  prev = __shfl_sync(full, prev, 0);
  return prev != 0u ? tb.row_arena + (size_t)(prev - 1u) * 16 : row;
}
#endif

struct MaRow {
  const double *pr;        // [MA_ACTION_COUNT]
  const double *cum_rad;   // [lev_ndowntrans]
  const double *cum_down;  // [lev_ndowntrans]
  const double *cum_up;    // [lev_nuptrans]
};
ARTIS_DEV MaRow ma_row(const Ctx &c, const int mgi, const int tc, const int gl) {
  MaRow r;
  if (rows_on_demand(c)) {
    const double *row = row_get(c, mgi, tc, gl);
    const int nd = c.st.lev_ndowntrans[gl];
    r.pr = row;
    r.cum_rad = row + MA_ACTION_COUNT;
    r.cum_down = row + MA_ACTION_COUNT + nd;
    r.cum_up = row + MA_ACTION_COUNT + 2 * nd;
  } else {
    const int doff = c.st.lev_downoffset[gl], uoff = c.st.lev_upoffset[gl];
    r.pr = tab_ma_rates(c, tc, gl);
    r.cum_rad = c.tb.ma_cum_raddeexc + (size_t)tc * c.st.ndowntrans_total + doff;
    r.cum_down = c.tb.ma_cum_intdown + (size_t)tc * c.st.ndowntrans_total + doff;
    r.cum_up = c.tb.ma_cum_intup + (size_t)tc * c.st.nuptrans_total + uoff;
  }
  return r;
}
ARTIS_DEV const double *ma_rates_row(const Ctx &c, const int mgi, const int tc, const int gl) {
  return rows_on_demand(c) ? row_get(c, mgi, tc, gl) : tab_ma_rates(c, tc, gl);
}
struct CoolRow {
  const double *cool;  // indexed with the global cooling-term index (ion_coolingoffset[ui] ...)
  const double *cexc;  // indexed with the global upward-transition index (lev_upoffset[...] ...)
};
ARTIS_DEV CoolRow cool_row(const Ctx &c, const int mgi, const int tc, const int ui) {
  CoolRow r;
  if (rows_on_demand(c)) {
    const double *row = row_get(c, mgi, tc, c.st.nlevels_total + ui);
    r.cool = row - c.st.ion_coolingoffset[ui];
    r.cexc = row + c.st.ion_ncoolingterms[ui] - ion_upfirst(c, ui);
  } else {
    r.cool = c.tb.cool_cum + (size_t)tc * c.st.ncoolingterms;
    r.cexc = c.tb.collexc_cum + (size_t)tc * c.st.nuptrans_total;
  }
  return r;
}

}  // namespace artis
