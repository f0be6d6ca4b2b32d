// rpkt.cuh -- r-packet propagation: boundary crossing, continuum opacity, Sobolev line
// sweep, event handlers, estimators. Mirrors rpkt.cc / boundary.cc / polarization.cc of the reference;
// each function cites the lines it follows.
#pragma once
#include "celltables.cuh"

namespace artis {

constexpr int MAX_GROUNDCONT = 96;  // per-thread scratch for groundcont_gamma_contr (checked at init)

// per-thread continuum-opacity cache == globals::kappa_rpkt_cont[tid] + phixslist[tid].groundcont_gamma_contr
// (globals.h:162-172, 72-80). Lives for one pass of one packet (reuse if same cell and |dnu/nu|<1e-4).
struct KappaCache {
  double nu, total, es, ff, bf, ffheating;
  int mgi;
  bool valid;
  // DETAILED_BF_ESTIMATORS_ON: a scheduler that sets defer_bfest gets the bf-rate estimator sweep of update_estimators
  // (radfield.cc:764-829) left pending here -- frequency of the step's midpoint, distance_e_cmf / nu * doppler, cell row --
  // and runs it with a whole warp right after the step (transport.cu coop_bfest_flush); bfest_factor == 0: nothing pending
  bool defer_bfest = false;
  int bfest_tc = 0;
  double bfest_nu = 0., bfest_factor = 0.;
  double gamma_contr[MAX_GROUNDCONT];
};

// ---- grid geometry (grid.cc:76-200, 2086) -----------------------------------------------------------------------
ARTIS_DEV int cell_mgi(const Ctx &c, int cellindex) { return c.st.cell_mgi[cellindex]; }
ARTIS_DEV int ncoordgrid(const Ctx &c, int d) { return d == 0 ? c.st.ncoordgrid_x : (d == 1 ? c.st.ncoordgrid_y : c.st.ncoordgrid_z); }
ARTIS_DEV double coordmax(const Ctx &c, int d) { return d == 0 ? c.st.coordmax_x : (d == 1 ? c.st.coordmax_y : c.st.coordmax_z); }
ARTIS_DEV int cellcoordpointnum(const Ctx &c, int cellindex, int d) {
  if (d == 0) return cellindex % c.st.ncoordgrid_x;
  if (d == 1) return (cellindex / c.st.ncoordgrid_x) % c.st.ncoordgrid_y;
  return (cellindex / (c.st.ncoordgrid_x * c.st.ncoordgrid_y)) % c.st.ncoordgrid_z;
}
ARTIS_DEV int coordcellindexincrement(const Ctx &c, int d) {
  return d == 0 ? 1 : (d == 1 ? c.st.ncoordgrid_x : c.st.ncoordgrid_x * c.st.ncoordgrid_y);
}
// grid.cc:2086: pos_min = -coordmax + (2 * n * coordmax / ncoordgrid), evaluated in the same order
ARTIS_DEV double cellcoordmin(const Ctx &c, int cellindex, int d) {
  const int n = cellcoordpointnum(c, cellindex, d);
  return -coordmax(c, d) + (2 * n * coordmax(c, d) / ncoordgrid(c, d));
}
ARTIS_DEV double wid_init(const Ctx &c) { return 2 * c.st.coordmax_x / c.st.ncoordgrid_x; }

// ---- boundary.cc:14-99 get_shellcrossdist: closest forward distance to an expanding spherical shell, -1 if none ----
ARTIS_DEV double get_shellcrossdist(const Ctx &c, const double pos[3], const double dir[3], const double shellradius,
                                    const bool isinnerboundary, const double tstart) {
  const double speed = vec_len(dir) * CLIGHT_PROP;
  const double q = shellradius / tstart / speed;
  const double a = dot3(dir, dir) - q * q;                                       // pow(x, 2) in the reference
  const double b = 2 * (dot3(dir, pos) - shellradius * shellradius / tstart / speed);
  const double cc = dot3(pos, pos) - shellradius * shellradius;
  const double discriminant = b * b - 4 * a * cc;
  if (discriminant < 0) {
    if (!(shellradius < vec_len(pos))) fatal(c, 110);  // assert_always in the reference
    return -1;
  } else if (discriminant > 0) {
    double d1 = (-b + sqrt(discriminant)) / 2 / a;
    double d2 = (-b - sqrt(discriminant)) / 2 / a;
    double posfinal1[3], posfinal2[3];
    for (int d = 0; d < 3; d++) {
      posfinal1[d] = pos[d] + d1 * dir[d];
      posfinal2[d] = pos[d] + d2 * dir[d];
    }
    const double shellradiusfinal1 = shellradius / tstart * (tstart + d1 / speed);
    const double shellradiusfinal2 = shellradius / tstart * (tstart + d2 / speed);
    if (!(fabs(vec_len(posfinal1) / shellradiusfinal1 - 1.) < 1e-3) || !(fabs(vec_len(posfinal2) / shellradiusfinal2 - 1.) < 1e-3))
      fatal(c, 111);  // assert_always in the reference
    // a solution that enters the boundary from the wrong radial direction is not a crossing
    if (isinnerboundary) {
      if (dot3(posfinal1, dir) > 0.) d1 = -1;
      if (dot3(posfinal2, dir) > 0.) d2 = -1;
    } else {
      if (dot3(posfinal1, dir) < 0.) d1 = -1;
      if (dot3(posfinal2, dir) < 0.) d2 = -1;
    }
    if (d1 < 0 && d2 < 0) return -1;
    if (d2 < 0) return d1;
    if (d1 < 0) return d2;
    return fmin(d1, d2);
  }
  if (!(shellradius <= vec_len(pos))) fatal(c, 112);
  return -1.;
}
// grid.cc:76-95 wid_init and :131-139 get_cellcoordmax for GRID_SPHERICAL1D (cellindex == shell number)
ARTIS_DEV double wid_init_1d(const Ctx &c, const int cellindex) {
  const int mgi = c.st.cell_mgi[cellindex];
  const double v_inner = mgi > 0 ? c.st.vout_model[mgi - 1] : 0.;
  return (c.st.vout_model[mgi] - v_inner) * c.st.tmin;
}
// ---- boundary.cc:101-330 boundary_cross, GRID_SPHERICAL1D branch: one radial coordinate, inner and outer shell ------
template <class P>
ARTIS_DEV double boundary_cross_1d(const Ctx &c, P &p, int *snext) {
  const double tstart = p.prop_time;
  const int cellindex = p.where;
  const double tmin = c.st.tmin;
  const double initpos = vec_len(p.pos);
  const double cmin = c.st.cell_posmin_x[cellindex];
  const double cmax = cmin + wid_init_1d(c, cellindex);
  const double vel = dot3(p.pos, p.dir) / vec_len(p.pos) * CLIGHT_PROP;
  const int ncoord = c.st.ncoordgrid_x;
  // 10 cm tolerance repair path (boundary.cc:165-222) on the radial coordinate
  for (int flip = 0; flip < 2; flip++) {
    const int direction = flip ? ARTIS_POS_X : ARTIS_NEG_X;
    const int invdirection = !flip ? ARTIS_POS_X : ARTIS_NEG_X;
    const int stride = flip ? -1 : 1;
    const bool isoutside = flip ? (initpos < (cmin / tmin * tstart - 10.)) : (initpos > (cmax / tmin * tstart + 10.));
    if (isoutside && (p.last_cross != direction)) {
      if ((vel - (initpos / tstart)) > 0) {
        if ((cellindex == (ncoord - 1) && stride > 0) || (cellindex == 0 && stride < 0)) {
          *snext = -99;
          return 0;
        }
        *snext = p.where + stride;
        p.last_cross = invdirection;
        return 0;
      }
      // "pretending last_cross is direction" has no effect here: boundary.cc:235 resets last_cross for this grid
    }
  }
  // invalid directions are handled by negative distances (boundary.cc:235-245)
  const double r_inner = cmin * tstart / tmin;
  const double d_inner = (r_inner > 0.) ? get_shellcrossdist(c, p.pos, p.dir, r_inner, true, tstart) : -1.;
  const double t_min = d_inner / CLIGHT_PROP;
  const double r_outer = cmax * tstart / tmin;
  const double d_outer = get_shellcrossdist(c, p.pos, p.dir, r_outer, false, tstart);
  const double t_max = d_outer / CLIGHT_PROP;
  double time = 1.e99;
  if ((t_max > 0) && (t_max < time)) {
    time = t_max;
    if (cellindex == (ncoord - 1)) {
      *snext = -99;
    } else {
      *snext = p.where + 1;
      p.last_cross = ARTIS_POS_X;
    }
  }
  if ((t_min > 0) && (t_min < time)) {
    time = t_min;
    if (cellindex == 0) {
      *snext = -99;
    } else {
      *snext = p.where - 1;
      p.last_cross = ARTIS_NEG_X;
    }
  }
  return CLIGHT_PROP * time;
}
// rpkt.cc:659-661 / gammapkt.cc:551-553: the largest credible distance to a cell boundary
template <class P>
ARTIS_DEV double max_sdist(const Ctx &c, const P &p, const double sdist) {
  return ARTIS_OPT_GRID1D(c) ? 2 * c.st.rmax * (p.prop_time + sdist / CLIGHT_PROP) / c.st.tmin
                                 : c.st.rmax * p.prop_time / c.st.tmin;
}

// ---- boundary.cc:101-330 boundary_cross ----------------------------------------------------------------------------
template <class P>
ARTIS_DEV double boundary_cross(const Ctx &c, P &p, int *snext) {
  if (ARTIS_OPT_GRID1D(c)) return boundary_cross_1d(c, p, snext);
  const double tstart = p.prop_time;
  const int cellindex = p.where;
  const double tmin = c.st.tmin;
  double initpos[3], cmin[3], cmax[3], vel[3];
  const double wid = wid_init(c);
  for (int d = 0; d < 3; d++) {
    initpos[d] = p.pos[d];
    cmin[d] = cellcoordmin(c, cellindex, d);
    cmax[d] = cmin[d] + wid;
    vel[d] = p.dir[d] * CLIGHT_PROP;
  }
  int last_cross = p.last_cross;
  const int negdirections[3] = {ARTIS_NEG_X, ARTIS_NEG_Y, ARTIS_NEG_Z};
  const int posdirections[3] = {ARTIS_POS_X, ARTIS_POS_Y, ARTIS_POS_Z};

  // 10 cm tolerance repair path (boundary.cc:160-222)
  for (int d = 0; d < 3; d++) {
    for (int flip = 0; flip < 2; flip++) {
      const int direction = flip ? posdirections[d] : negdirections[d];
      const int invdirection = !flip ? posdirections[d] : negdirections[d];
      const int stride = flip ? -coordcellindexincrement(c, d) : coordcellindexincrement(c, d);
      bool isoutside;
      if (flip) {
        isoutside = initpos[d] < (cmin[d] / tmin * tstart - 10.);
      } else {
        isoutside = initpos[d] > (cmax[d] / tmin * tstart + 10.);
      }
      if (isoutside && (last_cross != direction)) {
        if ((vel[d] - (initpos[d] / tstart)) > 0) {
          const int pn = cellcoordpointnum(c, cellindex, d);
          if ((pn == (ncoordgrid(c, d) - 1) && stride > 0) || (pn == 0 && stride < 0)) {
            *snext = -99;
            return 0;
          }
          *snext = p.where + stride;
          p.last_cross = invdirection;
          return 0;
        }
        last_cross = direction;
      }
    }
  }

  double t_max[3], t_min[3];
  for (int d = 0; d < 3; d++) {
    t_max[d] = ((initpos[d] - (vel[d] * tstart)) / ((cmax[d]) - (vel[d] * tmin)) * tmin) - tstart;
    t_min[d] = ((initpos[d] - (vel[d] * tstart)) / ((cmin[d]) - (vel[d] * tmin)) * tmin) - tstart;
  }
  int choice = 0;
  double time = 1.e99;
  for (int d = 0; d < 3; d++) {
    if ((t_max[d] > 0) && (t_max[d] < time) && (last_cross != negdirections[d])) {
      choice = posdirections[d];
      time = t_max[d];
      if (cellcoordpointnum(c, cellindex, d) == (ncoordgrid(c, d) - 1)) {
        *snext = -99;
      } else {
        *snext = p.where + coordcellindexincrement(c, d);
        p.last_cross = posdirections[d];
      }
    }
    if ((t_min[d] > 0) && (t_min[d] < time) && (last_cross != posdirections[d])) {
      choice = negdirections[d];
      time = t_min[d];
      if (cellcoordpointnum(c, cellindex, d) == 0) {
        *snext = -99;
      } else {
        *snext = p.where - coordcellindexincrement(c, d);
        p.last_cross = negdirections[d];
      }
    }
  }
  (void)choice;  // choice == 0 ("didn't find anything") yields 1e99*c and is caught by the maxsdist check
  return CLIGHT_PROP * time;
}

// boundary.cc:332-357 change_cell
ARTIS_DEV void change_cell(const Ctx &c, Pkt &p, int snext, Tally &tl) {
  if (snext == -99) {
    p.cold->escape_type = p.type;
    p.cold->escape_time = (int)p.prop_time;  // int field
    p.type = ARTIS_TYPE_ESCAPE;
    stat_inc(tl, COUNTER_NESC);
  } else {
    p.where = snext;
    stat_inc(tl, COUNTER_CELLCROSSINGS);
  }
}

// ---- rpkt.cc:26-65 closest_transition -----------------------------------------------------------------------------
// first index i in [lo, nlines) with line_nu[i] <= nu_cmf (list sorted by decreasing nu): std::lower_bound with
// the comparator of rpkt.cc:24
ARTIS_DEV int linelist_lower_bound(const Ctx &c, int lo, double nu_cmf) {
  int hi = c.st.nlines;
  while (lo < hi) {
    const int mid = lo + ((hi - lo) >> 1);
    if (!(c.st.line_nu[mid] <= nu_cmf))
      lo = mid + 1;
    else
      hi = mid;
  }
  return lo;
}
ARTIS_DEV int closest_transition(const Ctx &c, const double nu_cmf, const int next_trans) {
  const int left = next_trans;
  const int right = c.st.nlines - 1;
  if (nu_cmf < c.st.line_nu[right]) return -1;
  if (left > right) return -1;
  if (left > 0) return left;  // index-driven
  if (nu_cmf >= c.st.line_nu[0]) return 0;
  return linelist_lower_bound(c, next_trans, nu_cmf);
}
// rpkt.cc:511-555 closest_transition_empty. Where the reference reads past the end of the line list
// (left > right) or calls lower_bound on an inverted range, every outcome means "no more lines"; nlines+1 is
// what the reference stores in those cases.
ARTIS_DEV void closest_transition_empty(const Ctx &c, Pkt &p) {
  const int left = p.next_trans;
  const int right = c.st.nlines - 1;
  if (left > right) {
    p.next_trans = c.st.nlines + 1;
  } else if (p.nu_cmf >= c.st.line_nu[left]) {
    p.next_trans = left;
  } else if (p.nu_cmf < c.st.line_nu[right]) {
    p.next_trans = c.st.nlines + 1;
  } else {
    p.next_trans = linelist_lower_bound(c, left, p.nu_cmf);
  }
}

// ---- rpkt.cc:1027-1073 calculate_kappa_ff -------------------------------------------------------------------------
// the frequency-independent part: sum over the ions of Z^2 n_ion, in the reference's order of addition
ARTIS_DEV double kappa_ff_zsum(const Ctx &c, const int mgi) {
  double kappa_ff = 0.;
  int ui = 0;
  for (int element = 0; element < c.st.nelements; element++) {
    const int nions = c.st.el_nions[element];
    for (int ion = 0; ion < nions; ion++, ui++) {
      const int Z = c.st.ion_ionstage[ui] - 1;
      if (Z > 0) {
        const double nnion = ionstagepop_ui(c, mgi, element, ui);
        kappa_ff += Z * Z * 1 * nnion;
      }
    }
  }
  return kappa_ff;
}
ARTIS_DEV double calculate_kappa_ff(const Ctx &c, const int mgi, const double nu) {
  const float nne = c.cs.nne[mgi];
  const float T_e = c.cs.Te[mgi];
  // the ion sum is a property of the cell: cell_precompute has it (CellTables.ff_zsum, the same additions in the same order)
  const int tc = (c.tb.ff_zsum != nullptr) ? tabcell(c, mgi) : -1;
  double kappa_ff = (tc >= 0) ? c.tb.ff_zsum[tc] : kappa_ff_zsum(c, mgi);
  // sqrt(T_e) is evaluated in float in the reference (float argument, rpkt.cc:1064)
  kappa_ff *= 3.69255e8 / sqrtf(T_e) * pow(nu, -3) * nne * (1 - exp(-HOVERKB * nu / T_e));
  return kappa_ff;
}

// ---- rpkt.cc:1075-1207 calculate_kappa_bf_gammacontr ---------------------------------------------------------------
// The reference walks the continua in order of nu_edge, skips those of absent elements / unimportant ions before looking
// at the frequency, and stops at the first remaining one with nu < nu_edge. Because the list is sorted, every term at or
// beyond i0 = (first i with nu < nu_edge[i]) is either skipped or stops the loop, so the sum is exactly the sum over
// i < i0; that bound comes from a binary search here and the terms become independent.
ARTIS_DEV int bf_stop_index(const Ctx &c, const double nu) {
  int lo = 0, hi = c.st.nbfcontinua;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (nu < c.st.cont_nu_edge[mid]) {
      hi = mid;
    } else {
      lo = mid + 1;
    }
  }
  return lo;
}
// One term of the sum for a continuum with nu >= nu_edge: returns kappa_bf_contr at frequency nu (0 if skipped);
// *gcontr receives sigma_bf*probability*corrfactor (the gamma_contr of rpkt.cc:1158-1166) or 0.
// the arithmetic of one term, given its per-cell pair (level population, departure ratio) and its static record
ARTIS_DEV double kappa_bf_term_core(const Ctx &c, const double nnlevel, const double depratio, const ContRec &r, const double nu,
                                    const double expfac, double *gcontr) {
  const double sigma_bf = phixs_fromtable(c, r.tablerow, r.nu_edge, nu, r.inv_nu_edge);
  const double nu_max_phixs = r.nu_edge * c.st.last_phixs_nuovernuedge;
  // nnlevel < 0: element absent / unimportant ion (rpkt.cc:1101-1104); otherwise rpkt.cc:1128
  const bool counts = (nnlevel > 0.) && (nu <= nu_max_phixs);
  const double stimfactor = depratio * expfac;  // expfac = exp(-h nu / k T_e)
  double corrfactor = 1 - stimfactor;
  if (corrfactor < 0) corrfactor = 0.;
  *gcontr = counts ? sigma_bf * r.probability * corrfactor : 0.;
  return counts ? nnlevel * sigma_bf * r.probability * corrfactor : 0.;
}
// The same term as straight-line code for the cooperative sum (transport.cu), whose loop should neither branch nor outgrow
// the instruction buffers: both quotients by the two-FMA form (exact: ContRec), the table index clamped instead of
// tested, selects instead of branches. The clamped form equals atomic.cc:87-137 wherever 0 <= i < NPHIXSPOINTS - 1; a lane
// outside that range or without an exact reciprocal (never, for the data sets seen so far) takes kappa_bf_term_core.
ARTIS_DEV double kappa_bf_term_fast(const Ctx &c, const double nnlevel, const double depratio, const ContRec &r, const double nu,
                                    const double expfac, double *gcontr) {
  const double inv = r.inv_nu_edge, invinc = c.inv_nphixsnuincrement;
  const double q = nu * inv;
  const double x = fma(fma(-r.nu_edge, q, nu), inv, q);  // nu / nu_edge, correctly rounded
  const double y = x - 1.0;
  const double q2 = y * invinc;
  const double ireal = fma(fma(-c.st.nphixsnuincrement, q2, y), invinc, q2);  // (nu / nu_edge - 1) / NPHIXSNUINCREMENT
  const int i = (int)floor(ireal);
  const int npts = c.st.nphixspoints;
  const bool fast = (inv != 0.) && (invinc != 0.) && (i >= 0) && (i < npts - 1);
  if (!fast) return kappa_bf_term_core(c, nnlevel, depratio, r, nu, expfac, gcontr);
  const float *xs = c.st.phixs_tables + (size_t)r.tablerow * npts;
  const double a = xs[i];
  const double b = xs[i + 1];
  const double factor_b = ireal - i;
  const double sigma_bf = (double)(float)(((1. - factor_b) * a) + (factor_b * b));
  const double nu_max_phixs = r.nu_edge * c.st.last_phixs_nuovernuedge;
  const bool counts = (nnlevel > 0.) && (nu <= nu_max_phixs);
  double corrfactor = 1 - depratio * expfac;
  if (corrfactor < 0) corrfactor = 0.;
  *gcontr = counts ? sigma_bf * r.probability * corrfactor : 0.;
  return counts ? nnlevel * sigma_bf * r.probability * corrfactor : 0.;
}
ARTIS_DEV double kappa_bf_term(const Ctx &c, const int tc, const int i, const double nu, const double expfac, double *gcontr,
                               int *ground_slot = nullptr) {
  // every load is issued before the first use, and sigma_bf is interpolated even for a term that ends up skipped: the
  // dependent nnlevel -> branch -> table row -> cross-section chain was the latency of the whole sum
  const double *cell = c.tb.bf_cell + 2 * ((size_t)tc * c.st.nbfcontinua + i);
  const double nnlevel = cell[0];
  const double depratio = cell[1];
  const ContRec r = c.cont_rec[i];
  if (ground_slot != nullptr) *ground_slot = r.ground_slot;
  return kappa_bf_term_core(c, nnlevel, depratio, r, nu, expfac, gcontr);
}
// ground-continuum slot that term i feeds (rpkt.cc:1158), or -1
ARTIS_DEV int kappa_bf_ground_slot(const Ctx &c, const int i) { return c.cont_rec[i].ground_slot; }
// A continuum also contributes nothing above the end of its cross-section table, nu > nu_edge * last_phixs_nuovernuedge
// (rpkt.cc:1128): first index whose table still reaches nu. Everything below it is an exact zero and is not visited.
ARTIS_DEV int bf_start_index(const Ctx &c, const double nu, const int i0) {
  int lo = 0, hi = i0;
  const double reach = c.st.last_phixs_nuovernuedge;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (nu <= c.st.cont_nu_edge[mid] * reach) {
      hi = mid;
    } else {
      lo = mid + 1;
    }
  }
  return lo;
}

// ---- live continua --------------------------------------------------------------------------------------------------------------
// A term of the sum is n_level * sigma_bf(nu) * probability * max(0, 1 - departure_ratio * exp(-h nu / k T_e)) for
// nu_edge <= nu <= nu_edge * last_phixs_nuovernuedge. In the dilute radiation fields of a supernova most excited levels sit
// far below their Saha-Boltzmann population at T_e, the stimulated-recombination factor exceeds one over the whole
// frequency range of the cross-section table, and the reference adds an exact zero -- for five terms out of six of the
// classic model at maximum light, a further share because the ion is absent from the cell (rpkt.cc:1101-1104). Whether a
// continuum can EVER contribute in a cell does not depend on the packet: exp(-h nu / k T_e) is smallest at the top of the
// table, so departure_ratio * exp(-h nu_max / k T_e) >= 1 (with a margin of 1e-9 against the last bits of exp) means the
// clamp gives zero at every frequency. cell_precompute lists the other continua per cell, in list order
// (CellTables.bf_live_*), and the sums below run over that list: the same non-zero terms, so the same value up to the
// rounding of the additions (the tiles are counted along the live list), which is what already separates this sum from
// the reference's strictly serial one.
ARTIS_DEV bool bf_is_live(const Ctx &c, const int mgi, const double nnlevel, const double depratio, const double nu_edge) {
  if (!(nnlevel > 0.)) return false;
  const double nu_top = nu_edge * c.st.last_phixs_nuovernuedge;
  return depratio * exp(-HOVERKB * nu_top / (double)c.cs.Te[mgi]) < 1. + 1e-9;
}
// the list a cell's sum runs over: the live continua of the cell, or every continuum (no live lists built)
struct BfList {
  const double *pair;   // [n][2] level population (or -1) and departure ratio
  const ContRec *rec;   // [n]
  int n;
};
ARTIS_DEV BfList bf_list(const Ctx &c, const int tc) {
  BfList l;
  if (tc < 0) {  // a cell without a table row (empty or grey: the transport never asks; the probes may)
    l.pair = nullptr;
    l.rec = nullptr;
    l.n = 0;
  } else if (c.tb.bf_live_off != nullptr) {
    const long long o = c.tb.bf_live_off[tc];
    l.pair = c.tb.bf_live_pair + 2 * o;
    l.rec = c.tb.bf_live_rec + o;
    l.n = (int)(c.tb.bf_live_off[tc + 1] - o);
  } else {
    l.pair = c.tb.bf_cell + 2 * ((size_t)tc * c.st.nbfcontinua);
    l.rec = c.cont_rec;
    l.n = c.st.nbfcontinua;
  }
  return l;
}
// The range [*ilo, *i0) of the list that frequency nu reaches -- bf_start_index / bf_stop_index of the full list (two binary
// searches over the static edge frequencies, which every cell shares and the L1 keeps), mapped onto the cell's live list
// through the count of live continua before each list entry. Also returns what the reference's loop visits: *listed.
ARTIS_DEV void bf_list_range(const Ctx &c, const int tc, const double nu, int *ilo, int *i0, int *listed) {
  const int f0 = bf_stop_index(c, nu);
  const int flo = bf_start_index(c, nu, f0);
  *listed = f0 - flo;
  if (tc < 0) {
    *i0 = 0;
    *ilo = 0;
  } else if (c.tb.bf_live_off != nullptr) {
    const int *rank = c.tb.bf_live_rank + (size_t)tc * (c.st.nbfcontinua + 1);
    *i0 = rank[f0];
    *ilo = rank[flo];
  } else {
    *i0 = f0;
    *ilo = flo;
  }
}

// Summation order: the terms are added in tiles of BF_TILE consecutive continua (counted from the first visited one),
// each tile reduced as a balanced binary tree over adjacent pairs ((t0+t1)+(t2+t3))+..., and the tile sums accumulated
// in order. A warp produces exactly this tree with shuffle-down reductions of offsets 1, 2, 4, 8, 16
// (transport.cu coop_kappa_warm); a single thread produces it with a binary counter of partial sums (PairTree below:
// five doubles in registers instead of a 32-entry scratch array in local memory, which was a quarter of the hot
// per-thread stack). Host and device use the same code, so both give the same bits; the result differs from the
// reference's strictly serial sum by rounding only (~1e-16 relative, far inside the 1e-10 tolerance on kappa_cont;
// checked against the reference probes in tests/).
constexpr int BF_TILE = 32;
struct PairTree {
  double s0, s1, s2, s3, s4;  // pending partial sums of 1, 2, 4, 8, 16 consecutive terms
  // add term number j (0-based within the tile); returns true when the tile is complete, with its sum in *tile_sum
  ARTIS_DEV bool push(const int j, double v, double *tile_sum) {
    if ((j & 1) == 0) {
      s0 = v;
      return false;
    }
    v = s0 + v;
    if ((j & 2) == 0) {
      s1 = v;
      return false;
    }
    v = s1 + v;
    if ((j & 4) == 0) {
      s2 = v;
      return false;
    }
    v = s2 + v;
    if ((j & 8) == 0) {
      s3 = v;
      return false;
    }
    v = s3 + v;
    if ((j & 16) == 0) {
      s4 = v;
      return false;
    }
    *tile_sum = s4 + v;
    return true;
  }
};
ARTIS_DEV double calculate_kappa_bf_gammacontr(const Ctx &c, const int mgi, const double nu, KappaCache &k, Tally &tl) {
  double kappa_bf_sum = 0.;
  const int ng = c.st.nbfcontinua_ground;
  for (int g = 0; g < ng; g++) k.gamma_contr[g] = 0.;
  const int tc = tabcell(c, mgi);
  const BfList l = bf_list(c, tc);
  const double expfac = exp(-HOVERKB * nu / (double)c.cs.Te[mgi]);
  int i0, ilo, listed;
  bf_list_range(c, tc, nu, &ilo, &i0, &listed);
  tl.bf_evaluated += i0 - ilo;
  tl.bf_listed += listed;
  // whole tiles: the last one is padded with zero terms, exactly like the idle lanes of a warp
  const int iend = ilo + ((i0 - ilo + BF_TILE - 1) / BF_TILE) * BF_TILE;
  PairTree tree;
  // the terms are independent: four in flight per thread so that their table loads overlap
  for (int base = ilo; base < iend; base += 4) {
    double t[4];
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int u = 0; u < 4; u++) {
      const int i = base + u;
      t[u] = 0.;
      if (i < i0) {
        double gcontr;
        const ContRec r = l.rec[i];
        const int slot = r.ground_slot;
        t[u] = kappa_bf_term_core(c, l.pair[2 * i], l.pair[2 * i + 1], r, nu, expfac, &gcontr);
        if (gcontr != 0. && slot >= 0) k.gamma_contr[slot] += gcontr;
      }
    }
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int u = 0; u < 4; u++) {
      double tile_sum;
      if (tree.push((base + u - ilo) & (BF_TILE - 1), t[u], &tile_sum)) kappa_bf_sum += tile_sum;
    }
  }
  return kappa_bf_sum;
}

// rpkt.cc:1209-1309 calculate_kappa_rpkt_cont, in pieces: cache test, the cheap terms, the bf sum, the bookkeeping
ARTIS_DEV bool kappa_cache_hit(const KappaCache &k, const int mgi, const double nu_cmf) {
  return k.valid && (mgi == k.mgi) && (fabs(k.nu / nu_cmf - 1.0) < 1e-4);
}
ARTIS_DEV void kappa_store(const Ctx &c, KappaCache &k, const int mgi, const double nu_cmf, const double sigma, const double kappa_ff,
                           const double kappa_bf, const double kappa_ffheating, Tally &tl) {
  k.nu = nu_cmf;
  k.mgi = mgi;
  k.valid = true;
  k.total = sigma + kappa_bf + kappa_ff;
  k.es = sigma;
  k.ff = kappa_ff;
  k.bf = kappa_bf;
  k.ffheating = kappa_ffheating;
  if (!isfinite(k.total)) {  // rpkt.cc:1280-1294 soft fallback
    if (isfinite(k.es)) {
      k.ff = 0.;
      k.bf = 0.;
      k.total = k.es;
      tl.soft_repairs++;
    } else {
      fatal(c, 101);
    }
  }
}
template <class P>
ARTIS_DEV void calculate_kappa_rpkt_cont(const Ctx &c, const P &p, const int mgi, KappaCache &k, Tally &tl) {
  const double nu_cmf = p.nu_cmf;
  if (kappa_cache_hit(k, mgi, nu_cmf)) return;
  tl.kappa_evals++;
  const float nne = c.cs.nne[mgi];
  double sigma = 0.0, kappa_ff = 0., kappa_bf = 0., kappa_ffheating = 0.;
  if (c.st.do_r_lc) {
    if (c.st.opacity_case == 4) {
      sigma = SIGMA_T * nne;
      kappa_ff = calculate_kappa_ff(c, mgi, nu_cmf);
      kappa_ffheating = kappa_ff;
      kappa_bf = calculate_kappa_bf_gammacontr(c, mgi, nu_cmf, k, tl);
    } else {
      sigma = 0.;
      kappa_ff = 1e5 * calculate_kappa_ff(c, mgi, nu_cmf);
      kappa_bf = 0.;
    }
  }
  kappa_store(c, k, mgi, nu_cmf, sigma, kappa_ff, kappa_bf, kappa_ffheating, tl);
}

// ---- rpkt.cc:67-328 get_event ---------------------------------------------------------------------------------------
// `p` is only modified where the reference modifies pkt_ptr: next_trans and mastate.
ARTIS_DEV double get_event(const Ctx &c, const int mgi, Pkt &p, int *rpkt_eventtype, const double tau_rnd,
                           const double abort_dist, KappaCache &k, Tally &tl) {
  double tau = 0., dist = 0., edist = 0.;
  Ray abortray = ray_of(p);
  move_pkt_withtime(c, abortray, abort_dist / 2.);
  move_pkt_withtime(c, abortray, abort_dist / 2.);
  const double nu_cmf_abort = abortray.nu_cmf;
  Ray d = ray_of(p);
  calculate_kappa_rpkt_cont(c, p, mgi, k, tl);
  const double kap_cont = k.total * doppler_packet(c, p);
  const int tc = tabcell(c, mgi);
  while (true) {
    const int lineindex = closest_transition(c, d.nu_cmf, d.next_trans);
    if (lineindex >= 0) {
      tl.lines_scanned++;
      const double nu_trans = c.st.line_nu[lineindex];
      d.next_trans = lineindex + 1;
      double ldist;
      if (d.nu_cmf <= nu_trans) {
        ldist = 0;
      } else if (!ARTIS_OPT_RELDOPPLER(c)) {
        ldist = CLIGHT * d.prop_time * (d.nu_cmf / nu_trans - 1);
      } else {
        const double nu_r = nu_trans / d.nu_rf;
        const double ct = CLIGHT * d.prop_time;
        const double r = vec_len(d.pos);
        const double mu = dot3(d.dir, d.pos) / r;
        ldist = -mu * r + (ct - nu_r * nu_r * sqrt(ct * ct - (1 + r * r * (1 - mu * mu) * (1 + pow(nu_r, -2))))) /
                              (1 + nu_r * nu_r);
      }
      if (ldist < 0.) {
        if (!(ldist >= -100.)) fatal(c, 102);
        ldist = 0.;
        tl.soft_repairs++;
      }
      const double tau_cont = kap_cont * ldist;
      if (tau_rnd - tau > tau_cont) {
        if (nu_trans < nu_cmf_abort) {
          d.next_trans -= 1;
          p.next_trans = d.next_trans;
          return DBL_MAX_;
        }
        tl.lines_tau++;
        const int element = c.st.line_element[lineindex];
        const int glu = c.line_upper_g[lineindex];
        const int gll = c.line_lower_g[lineindex];
        const double A_ul = (double)c.st.line_einstein_A[lineindex];
        const double B_ul = CLIGHTSQUAREDOVERTWOH / pow(nu_trans, 3) * A_ul;
        const double B_lu = (double)c.st.lev_statweight[glu] / (double)c.st.lev_statweight[gll] * B_ul;
        const double n_u = tab_levelpop(c, tc, glu);
        const double n_l = tab_levelpop(c, tc, gll);
        double tau_line = (B_lu * n_l - B_ul * n_u) * HCLIGHTOVERFOURPI * d.prop_time;
        if (tau_line < 0) tau_line = 0.;
        if (tau_rnd - tau > tau_cont + tau_line) {
          dist = dist + ldist;
          tau += tau_cont + tau_line;
          move_pkt_withtime(c, d, ldist);
          // radfield::update_lineestimator: DETAILED_LINE_ESTIMATORS_ON is false in every shipped header
        } else {
          p.ma_element = element;
          p.ma_ion = c.st.line_ion[lineindex];
          p.ma_level = c.st.line_upper[lineindex];
          p.ma_activatingline = lineindex;
          edist = dist + ldist;
          if (edist >= abort_dist) {  // rpkt.cc:249-262
            edist = abort_dist * (1 - 2e-8);
            tl.soft_repairs++;
          }
          *rpkt_eventtype = RPKT_EVENTTYPE_BB;
          p.next_trans = d.next_trans;
          return edist;
        }
      } else {
        edist = dist + (tau_rnd - tau) / kap_cont;
        d.next_trans -= 1;
        *rpkt_eventtype = RPKT_EVENTTYPE_CONT;
        p.next_trans = d.next_trans;
        return edist;
      }
    } else {
      d.next_trans = c.st.nlines + 1;
      const double tau_cont = kap_cont * (abort_dist - dist);
      if (tau_rnd - tau > tau_cont) {
        edist = DBL_MAX_;
      } else {
        edist = dist + (tau_rnd - tau) / kap_cont;
        *rpkt_eventtype = RPKT_EVENTTYPE_CONT;
      }
      p.next_trans = d.next_trans;
      return edist;
    }
  }
}

}  // namespace artis
