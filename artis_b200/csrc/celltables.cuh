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
// ratecoeff.cc:1247-1310, LUT mode (NO_LUT_PHOTOION false): W * interpolated coefficient * renormalisation
ARTIS_DEV double get_corrphotoioncoeff(const Ctx &c, const int mgi, const int gl_lower, const int t) {
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

// macroatom.cc:57-159 calculate_macroatom_transitionrates for one (cell, level)
ARTIS_DEV void build_ma_level(const Ctx &c, const int mgi, const int gl, const double t_mid) {
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
  double *cum_rad = c.tb.ma_cum_raddeexc + (size_t)tc * c.st.ndowntrans_total + doff;
  double *cum_down = c.tb.ma_cum_intdown + (size_t)tc * c.st.ndowntrans_total + doff;
  for (int i = 0; i < ndown; i++) {
    const int lineindex = c.st.downtrans_lineindex[doff + i];
    const int gll = c.line_lower_g[lineindex];
    const double epsilon_target = c.st.lev_epsilon[gll];
    const double epsilon_trans = epsilon_current - epsilon_target;
    const double R = rad_deexcitation_ratecoeff(c, tc, gl, gll, epsilon_trans, lineindex, t_mid);
    const double C = col_deexcitation_ratecoeff(T_e, nne, epsilon_trans, lineref(c, lineindex),
                                                (double)c.st.lev_statweight[gll], statweight);
    pr[MA_RADDEEXC] += R * epsilon_trans;
    pr[MA_COLDEEXC] += C * epsilon_trans;
    pr[MA_INTERNALDOWNSAME] += (R + C) * epsilon_target;
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
      pr[MA_INTERNALDOWNLOWER] += (R + C) * epsilon_target;
      pr[MA_RADRECOMB] += R * epsilon_trans;
      pr[MA_COLRECOMB] += C * epsilon_trans;
    }
  }

  const int nup = c.st.lev_nuptrans[gl], uoff = c.st.lev_upoffset[gl];
  double *cum_up = c.tb.ma_cum_intup + (size_t)tc * c.st.nuptrans_total + uoff;
  for (int i = 0; i < nup; i++) {
    const int lineindex = c.st.uptrans_lineindex[uoff + i];
    const int glu = c.line_upper_g[lineindex];
    const double epsilon_trans = c.st.lev_epsilon[glu] - epsilon_current;
    const double R = rad_excitation_ratecoeff(c, mgi, tc, gl, glu, epsilon_trans, lineindex, t_mid);
    const double C = col_excitation_ratecoeff(T_e, nne, lineref(c, lineindex), epsilon_trans, statweight,
                                              (double)c.st.lev_statweight[glu]);
    const double NT = 0.;  // nonthermal::nt_excitation_ratecoeff: NT_EXCITATION_ON is false in every options header
    pr[MA_INTERNALUPSAME] += (R + C + NT) * epsilon_current;
    cum_up[i] = pr[MA_INTERNALUPSAME];
  }

  if (ion_has_upper(c, ui) && level < c.st.ion_ionisinglevels[ui]) {
    // MA_INTERNALUPHIGHERNT: nonthermal::nt_ionization_ratecoeff needs the Spencer-Fano solution (NT_ON builds only)
    const int ntargets = c.st.lev_nphixstargets[gl];
    for (int t = 0; t < ntargets; t++) {
      const double epsilon_trans = phixs_threshold(c, ui, gl, t);
      const double R = get_corrphotoioncoeff(c, mgi, gl, t);
      const double C = col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans);
      pr[MA_INTERNALUPHIGHER] += (R + C) * epsilon_current;
    }
  }
  double *out = c.tb.ma_rates + ((size_t)tc * c.st.nlevels_total + gl) * MA_ACTION_COUNT;
  for (int a = 0; a < MA_ACTION_COUNT; a++) out[a] = pr[a];
}

// kpkt.cc:167-308 calculate_kpkt_rates_ion for one (cell, ion); `oldcoolingsum` = sum of cooling_contrib_ion of the
// ions before this one in (element, ion) order, accumulated like the selection loop of do_kpkt (kpkt.cc:527-536)
ARTIS_DEV void build_cooling_ion(const Ctx &c, const int mgi, const int ui) {
  const int tc = tabcell(c, mgi);
  const int element = c.uion_element[ui];
  const float nne = c.cs.nne[mgi];
  const float T_e = c.cs.Te[mgi];
  double oldcoolingsum = 0.;
  for (int u = 0; u < ui; u++) oldcoolingsum += c.cs.cooling_contrib_ion[(size_t)mgi * c.st.includedions + u];
  double contrib = oldcoolingsum;
  int i = c.st.ion_coolingoffset[ui];
  double *cool = c.tb.cool_cum + (size_t)tc * c.st.ncoolingterms;
  double *cexc = c.tb.collexc_cum + (size_t)tc * c.st.nuptrans_total;
  const int gl0 = c.st.ion_leveloffset[ui];
  const int nlevels = c.st.ion_nlevels[ui];
  const int ionisinglevels = c.st.ion_ionisinglevels[ui];
  const double nncurrention = ionstagepop_ui(c, mgi, element, ui);
  const int ioncharge = c.st.ion_ionstage[ui] - 1;
  if (ioncharge > 0) {
    contrib += 1.426e-27 * sqrtf(T_e) * pow((double)ioncharge, 2) * nncurrention * nne;
    cool[i++] = contrib;
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
        cexc[uoff + ii] = contrib;
      }
      cool[i++] = contrib;
    }
    if (hasupper && level < ionisinglevels) {
      const int ntargets = c.st.lev_nphixstargets[gl];
      for (int t = 0; t < ntargets; t++) {
        const double epsilon_trans = phixs_threshold(c, ui, gl, t);
        contrib += nnlevel * col_ionization_ratecoeff(c, T_e, nne, ui, gl, t, epsilon_trans) * epsilon_trans;
        cool[i++] = contrib;
      }
      for (int t = 0; t < ntargets; t++) {
        const double nnupperion = ionstagepop_ui(c, mgi, element, ui + 1);
        contrib += get_bfcoolingcoeff(c, gl, t, T_e) * nnupperion * nne;
        cool[i++] = contrib;
      }
    }
  }
  // fill the tail when levels >= ionisinglevels own (uncomputed) list slots: the reference leaves them untouched
  // and asserts i == offset + ncoolingterms (kpkt.cc:300); here a mismatch is reported as a fatal table error
  if (i != c.st.ion_coolingoffset[ui] + c.st.ion_ncoolingterms[ui]) fatal(c, 202);
}

}  // namespace artis
