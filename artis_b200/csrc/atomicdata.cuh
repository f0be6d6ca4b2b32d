// atomicdata.cuh -- device mirrors of the read-only accessors on the path:
// level populations (ltepop.cc:307-564), photoionisation cross sections (atomic.cc:87), temperature LUTs
// (ratecoeff.cc:686,1026; kpkt.cc:69), radiation field (radfield.cc:898, radfield.h:44).
//
// The reference's per-thread `cellhistory` cache (globals.h:174-208) has no device analogue: level populations
// are evaluated on the fly from groundlevelpop[ion] * g/g0 * exp(-dE/kT_exc) (one FP64 exp per lookup, same
// arithmetic as ltepop.cc:329-347), which costs less than a random 32-byte HBM sector per lookup.
#pragma once
#include "common.cuh"

namespace artis {

ARTIS_DEV int uion_of(const Ctx &c, int element, int ion) { return c.st.el_ionoffset[element] + ion; }
ARTIS_DEV int glevel_of(const Ctx &c, int element, int ion, int level) {
  return c.st.ion_leveloffset[uion_of(c, element, ion)] + level;
}
ARTIS_DEV int get_nions(const Ctx &c, int element) { return c.st.el_nions[element]; }
ARTIS_DEV int get_ionstage(const Ctx &c, int element, int ion) { return c.st.ion_ionstage[uion_of(c, element, ion)]; }
ARTIS_DEV double epsilon(const Ctx &c, int element, int ion, int level) {
  return c.st.lev_epsilon[glevel_of(c, element, ion, level)];
}
ARTIS_DEV double stat_weight(const Ctx &c, int element, int ion, int level) {
  return (double)c.st.lev_statweight[glevel_of(c, element, ion, level)];
}
ARTIS_DEV float elem_abundance(const Ctx &c, int mgi, int element) {
  return c.cs.elem_abundance[(size_t)mgi * c.st.nelements + element];
}

// ltepop.cc:307-326 get_groundlevelpop
ARTIS_DEV double groundlevelpop_ui(const Ctx &c, int mgi, int element, int ui) {
  const double nn = (double)c.cs.groundlevelpop[(size_t)mgi * c.st.includedions + ui];
  if (nn < c.st.opt_minpop) {
    return (elem_abundance(c, mgi, element) > 0) ? c.st.opt_minpop : 0.;
  }
  return nn;
}
ARTIS_DEV double get_groundlevelpop(const Ctx &c, int mgi, int element, int ion) {
  return groundlevelpop_ui(c, mgi, element, uion_of(c, element, ion));
}
// LTEPOP_EXCITATIONTEMPERATURE (artisoptions_classic.h:32 / _nltenebular.h:36)
ARTIS_DEV double T_exc_of(const Ctx &c, int mgi) { return (double)(c.st.opt_ltepop_use_TJ ? c.cs.TJ[mgi] : c.cs.Te[mgi]); }

// ltepop.cc:329-347 calculate_levelpop_lte, with global level indices (gl0 = the ion's ground level)
ARTIS_DEV double levelpop_lte_g(const Ctx &c, int mgi, int element, int ui, int gl0, int gl) {
  const double nnground = groundlevelpop_ui(c, mgi, element, ui);
  if (gl == gl0) return nnground;
  const double T_exc = T_exc_of(c, mgi);
  const double W = 1.;
  const double E_level = c.st.lev_epsilon[gl];
  const double E_ground = c.st.lev_epsilon[gl0];
  return (nnground * W * (double)c.st.lev_statweight[gl] / (double)c.st.lev_statweight[gl0] *
          exp(-(E_level - E_ground) / KB / T_exc));
}

// ltepop.cc:349-430 calculate_levelpop (= get_levelpop: the cellhistory copy holds the same number,
// update_grid.cc:685-697). NLTE_POPS_ON branches follow ltepop.cc:362-409.
ARTIS_DEV double levelpop_g(const Ctx &c, int mgi, int element, int ui, int gl) {
  const int gl0 = c.st.ion_leveloffset[ui];
  const int level = gl - gl0;
  double nn;
  if (level > 0 && ARTIS_OPT_NLTE(c) && c.st.ion_nlevels_nlte[ui] > 0) {
    const int nlev_nlte = c.st.ion_nlevels_nlte[ui];
    const double *pops = c.cs.nlte_pops + (size_t)mgi * c.cs.total_nlte_levels;
    if (level <= nlev_nlte) {  // LEVEL_IS_NLTE (artisoptions_nltenebular.h): levels 1..nlevels_nlte
      const double nltepop_over_rho = pops[c.st.ion_first_nlte[ui] + level - 1];
      if (nltepop_over_rho >= -0.9) return nltepop_over_rho * (double)c.cs.rho[mgi];
      nn = levelpop_lte_g(c, mgi, element, ui, gl0, gl);
    } else {
      const double superlevelpop_over_rho = pops[c.st.ion_first_nlte[ui] + nlev_nlte];
      if (superlevelpop_over_rho >= -0.9) {
        // nltepop.cc:1543-1554 superlevel_boltzmann: relative to the first level of the superlevel
        const int gls = gl0 + nlev_nlte + 1;
        const double T_exc = T_exc_of(c, mgi);
        return superlevelpop_over_rho * (double)c.cs.rho[mgi] *
               ((double)c.st.lev_statweight[gl] / (double)c.st.lev_statweight[gls] *
                exp(-(c.st.lev_epsilon[gl] - c.st.lev_epsilon[gls]) / KB / T_exc));
      }
      nn = levelpop_lte_g(c, mgi, element, ui, gl0, gl);
    }
  } else {
    nn = levelpop_lte_g(c, mgi, element, ui, gl0, gl);
  }
  if (nn < c.st.opt_minpop) {
    nn = (elem_abundance(c, mgi, element) > 0) ? c.st.opt_minpop : 0.;
  }
  return nn;
}
ARTIS_DEV double get_levelpop(const Ctx &c, int mgi, int element, int ion, int level) {
  const int ui = uion_of(c, element, ion);
  return levelpop_g(c, mgi, element, ui, c.st.ion_leveloffset[ui] + level);
}
// ltepop.cc:558-564
ARTIS_DEV double ionstagepop_ui(const Ctx &c, int mgi, int element, int ui) {
  return groundlevelpop_ui(c, mgi, element, ui) * (double)c.cs.partfunct[(size_t)mgi * c.st.includedions + ui] /
         (double)c.st.lev_statweight[c.st.ion_leveloffset[ui]];
}
ARTIS_DEV double ionstagepop(const Ctx &c, int mgi, int element, int ion) {
  return ionstagepop_ui(c, mgi, element, uion_of(c, element, ion));
}
// ltepop.cc:539-556
ARTIS_DEV double calculate_sahafact_g(const Ctx &c, int gl_lower, int gl_upper, double T, double E_threshold) {
  const double g_lower = (double)c.st.lev_statweight[gl_lower];
  const double g_upper = (double)c.st.lev_statweight[gl_upper];
  return SAHACONST * g_lower / g_upper * pow(T, -1.5) * exp(E_threshold / KB / T);
}

// atomic.cc:87-137 photoionization_crosssection_fromtable (phixsdata_v2 branch)
// inv_nu_edge: RN(1/nu_edge) when that gives an exact quotient (ContRec), or 0 for the plain division
ARTIS_DEV double phixs_fromtable(const Ctx &c, int tablerow, double nu_edge, double nu, double inv_nu_edge = 0.) {
  const float *xs = c.st.phixs_tables + (size_t)tablerow * c.st.nphixspoints;
  const double ireal = div_by_known(div_by_known(nu, nu_edge, inv_nu_edge) - 1.0, c.st.nphixsnuincrement,
                                    inv_nu_edge != 0. ? c.inv_nphixsnuincrement : 0.);
  const int i = (int)floor(ireal);
  float sigma_bf;
  if (i < 0) {
    sigma_bf = 0.0f;
  } else if (i < c.st.nphixspoints - 1) {
    const double a = xs[i];
    const double b = xs[i + 1];
    const double factor_b = ireal - i;
    sigma_bf = (float)(((1. - factor_b) * a) + (factor_b * b));
  } else {
    const double nu_max_phixs = nu_edge * c.st.last_phixs_nuovernuedge;
    sigma_bf = (float)(xs[c.st.nphixspoints - 1] * pow(nu_max_phixs / nu, 3));
  }
  return (double)sigma_bf;
}

// phixs targets of a level
ARTIS_DEV int get_nphixstargets_g(const Ctx &c, int gl) { return c.st.lev_nphixstargets[gl]; }
ARTIS_DEV int get_phixsupperlevel_g(const Ctx &c, int gl, int t) { return c.st.phixstarget_level[c.st.lev_phixstargetoffset[gl] + t]; }
ARTIS_DEV double get_phixsprobability_g(const Ctx &c, int gl, int t) { return c.st.phixstarget_prob[c.st.lev_phixstargetoffset[gl] + t]; }
// sn3d.h:64-69 get_bflutindex
ARTIS_DEV int bflutindex(const Ctx &c, int tempindex, int gl, int phixstargetindex) {
  const int contindex = -1 - c.st.lev_cont_index[gl] + phixstargetindex;
  return tempindex * c.st.nbfcontinua + contindex;
}
// shared shape of ratecoeff.cc:686-708 (spontrecomb), :1026-1043 (corrphotoion), kpkt.cc:69-82 (bfcooling)
ARTIS_DEV double lut_interp(const Ctx &c, const double *table, int gl, int phixstargetindex, double T, bool assert_low) {
  const int lowerindex = (int)floor(log(T / c.st.opt_mintemp) / c.st.T_step_log);
  if (lowerindex < c.st.opt_tablesize - 1) {
    const int upperindex = lowerindex + 1;
    const double T_lower = c.st.opt_mintemp * exp(lowerindex * c.st.T_step_log);
    const double T_upper = c.st.opt_mintemp * exp(upperindex * c.st.T_step_log);
    const double f_upper = table[bflutindex(c, upperindex, gl, phixstargetindex)];
    const double f_lower = table[bflutindex(c, lowerindex, gl, phixstargetindex)];
    return (f_lower + (f_upper - f_lower) / (T_upper - T_lower) * (T - T_lower));
  }
  return table[bflutindex(c, c.st.opt_tablesize - 1, gl, phixstargetindex)];
}

// radfield.h:44 dbb with float T, W as passed by radfield::radfield (radfield.cc:930-934)
ARTIS_DEV double dbb(double nu, float T, float W) { return W * TWOHOVERCLIGHTSQUARED * pow(nu, 3) / expm1(HOVERKB * nu / T); }
// radfield.cc:596-610 select_bin: -2 below the first bin, -1 above the last one, else the first bin whose upper
// boundary is greater than nu (std::upper_bound)
ARTIS_DEV int radfield_select_bin(const Ctx &c, const double nu) {
  if (nu < c.st.radfield_nu_lower_first) return -2;
  int lo = 0, hi = c.st.opt_radfieldbincount;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (nu < c.st.radfieldbin_nu_upper[mid]) {
      hi = mid;
    } else {
      lo = mid + 1;
    }
  }
  return lo >= c.st.opt_radfieldbincount ? -1 : lo;
}
// radfield.cc:898-943 radfield(): the bin's dilute black body from FIRST_NLTE_RADFIELD_TIMESTEP on when
// MULTIBIN_RADFIELD_MODEL_ON (zero outside the bins and in bins without a fit), else the full-spectrum fit
ARTIS_DEV double radfield_J(const Ctx &c, double nu, int mgi) {
  if (c.st.opt_multibin_radfield && c.nts >= c.st.opt_first_nlte_radfield_timestep) {
    const int binindex = radfield_select_bin(c, nu);
    if (binindex >= 0) {
      const size_t mgibinindex = (size_t)c.st.nonemptymgi_of_mgi[mgi] * c.st.opt_radfieldbincount + binindex;
      const float W = c.cs.radfieldbin_W[mgibinindex];
      if (W >= 0.f) return dbb(nu, c.cs.radfieldbin_TR[mgibinindex], W);
    }
    return 0.;
  }
  return dbb(nu, c.cs.TR[mgi], c.cs.W[mgi]);
}

// atomic.cc:437-454
ARTIS_DEV double get_phixs_threshold_g(const Ctx &c, int gl_lower, int gl_upperion_ground, int t) {
  const int upperlevel = get_phixsupperlevel_g(c, gl_lower, t);
  return c.st.lev_epsilon[gl_upperion_ground + upperlevel] - c.st.lev_epsilon[gl_lower];
}

}  // namespace artis
