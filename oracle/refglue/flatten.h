// flatten.h -- host glue: flattens the reference's process globals into the C-ABI tables of include/artis_gpu.h.
//
// TEST INFRASTRUCTURE here (it is compiled into oracle/_ref/<mode>/sn3d_ref to dump snapshots), and at the same
// time the worked example of the ~200-line glue TU a maintainer adds to the reference: it is
// compiled with the reference's own headers and the same artisoptions.h, reads only public globals plus the few
// file-static arrays exposed by the wrapper TUs (refglue.h), and never touches reference sources.
#ifndef REFGLUE_FLATTEN_H
#define REFGLUE_FLATTEN_H

#include <map>
#include <vector>

#include "artis_gpu.h"
#include "atomic.h"
#include "decay.h"
#include "globals.h"
#include "grid.h"
#include "kpkt.h"
#include "nonthermal.h"
#include "radfield.h"
#include "ratecoeff.h"
#include "refglue.h"
#include "sn3d.h"
#include "stats.h"
#include "vpkt.h"

namespace grid {
extern __managed__ struct gridcell *cell;  // grid.cc:47 (non-static, not declared in grid.h)
extern __managed__ double *vout_model;     // grid.cc:37 (non-static, not declared in grid.h)
}
namespace nonthermal {
extern __managed__ double nt_energy_deposited;  // nonthermal.cc:115 (non-static, not declared in nonthermal.h)
}

namespace refglue {

// Owns the storage behind the pointers of one table struct.
struct Storage {
  std::vector<std::vector<int>> ints;
  std::vector<std::vector<float>> floats;
  std::vector<std::vector<double>> doubles;
  std::vector<std::vector<long long>> longs;
  int *I(size_t n) {
    ints.emplace_back(n ? n : 1, 0);
    return ints.back().data();
  }
  float *F(size_t n) {
    floats.emplace_back(n ? n : 1, 0.f);
    return floats.back().data();
  }
  double *D(size_t n) {
    doubles.emplace_back(n ? n : 1, 0.);
    return doubles.back().data();
  }
  long long *L(size_t n) {
    longs.emplace_back(n ? n : 1, 0);
    return longs.back().data();
  }
};

inline void flatten_static(ArtisStaticTables *T, Storage *st) {
  ArtisStaticTables &t = *T;
  memset(&t, 0, sizeof(t));
  t.abi_version = ARTIS_GPU_ABI_VERSION;
  t.opt_grid_type = (GRID_TYPE == GRID_SPHERICAL1D) ? 1 : 0;
  t.opt_nlte_pops_on = NLTE_POPS_ON;
  {
    // LTEPOP_EXCITATIONTEMPERATURE is a macro over `modelgridindex`: detect which getter it names by probing
    const int modelgridindex = 0;
    const float save = grid::modelgrid[0].TJ;
    grid::modelgrid[0].TJ = -12345.f;
    t.opt_ltepop_use_TJ = ((float)(LTEPOP_EXCITATIONTEMPERATURE) == -12345.f) ? 1 : 0;
    grid::modelgrid[0].TJ = save;
  }
#ifdef POL_ON
  t.opt_pol_on = 1;
#endif
#ifdef DIPOLE
  t.opt_dipole = 1;
#endif
#ifdef VPKT_ON
  t.opt_vpkt_on = 1;
#endif
  t.opt_no_lut_photoion = NO_LUT_PHOTOION;
  t.opt_no_lut_bfheating = NO_LUT_BFHEATING;
  t.opt_multibin_radfield = MULTIBIN_RADFIELD_MODEL_ON;
  t.opt_radfieldbincount = RADFIELDBINCOUNT;
  t.opt_first_nlte_radfield_timestep = FIRST_NLTE_RADFIELD_TIMESTEP;
  t.opt_detailed_bf_estimators = DETAILED_BF_ESTIMATORS_ON;
  t.opt_detailed_bf_usefromtimestep = DETAILED_BF_ESTIMATORS_USEFROMTIMESTEP;
  t.opt_nt_on = NT_ON;
  t.opt_nt_solve_spencerfano = NT_SOLVE_SPENCERFANO;
  t.opt_nt_max_auger_electrons = NT_MAX_AUGER_ELECTRONS;
  t.opt_relativistic_doppler = USE_RELATIVISTIC_DOPPLER_SHIFT;
  t.opt_instant_particle_deposition = INSTANT_PARTICLE_DEPOSITION;
#ifdef RECORD_LINESTAT
  t.opt_record_linestat = 1;
#endif
  t.opt_tablesize = TABLESIZE;
  t.opt_mintemp = MINTEMP;
  t.opt_maxtemp = MAXTEMP;
  t.opt_minpop = MINPOP;
  t.T_step_log = T_step_log;
  t.nu_min_r = globals::nu_min_r;
  t.nu_max_r = globals::nu_max_r;
  t.opacity_case = globals::opacity_case;
  t.gamma_grey = globals::gamma_grey;
  t.do_r_lc = globals::do_r_lc;
  t.n_kpktdiffusion_timesteps = globals::n_kpktdiffusion_timesteps;
  t.kpktdiffusion_timescale = globals::kpktdiffusion_timescale;

  // ---- geometry
  t.ncoordgrid_x = grid::ncoordgrid[0];
  t.ncoordgrid_y = grid::ncoordgrid[1];
  t.ncoordgrid_z = grid::ncoordgrid[2];
  t.ngrid = grid::ngrid;
  t.npts_model = grid::get_npts_model();
  t.nonempty_npts_model = grid::get_nonempty_npts_model();
  t.coordmax_x = globals::coordmax[0];
  t.coordmax_y = globals::coordmax[1];
  t.coordmax_z = globals::coordmax[2];
  t.tmin = globals::tmin;
  t.rmax = globals::rmax;
  t.vmax = globals::vmax;
  {
    int *cm = st->I(t.ngrid);
    double *px = st->D(t.ngrid);
    for (int n = 0; n < t.ngrid; n++) {
      cm[n] = grid::get_cell_modelgridindex(n);
      px[n] = grid::cell[n].pos_min[0];
    }
    t.cell_mgi = cm;
    t.cell_posmin_x = px;
    const int nv = (GRID_TYPE == GRID_SPHERICAL1D) ? grid::get_npts_model() : 0;
    double *vo = st->D(nv);
    for (int m = 0; m < nv; m++) vo[m] = grid::vout_model[m];
    t.vout_model = vo;
    int *ne = st->I(t.npts_model + 1);
    for (int m = 0; m < t.npts_model; m++)
      ne[m] = (grid::get_numassociatedcells(m) > 0) ? grid::get_modelcell_nonemptymgi(m) : -1;
    ne[t.npts_model] = -1;
    t.nonemptymgi_of_mgi = ne;
  }
  // ---- time steps
  t.ntstep = globals::ntstep;
  {
    double *a = st->D(t.ntstep), *b = st->D(t.ntstep), *c = st->D(t.ntstep);
    for (int n = 0; n < t.ntstep; n++) {
      a[n] = globals::time_step[n].start;
      b[n] = globals::time_step[n].width;
      c[n] = globals::time_step[n].mid;
    }
    t.ts_start = a;
    t.ts_width = b;
    t.ts_mid = c;
  }
  // ---- elements / ions / levels
  t.nelements = get_nelements();
  t.max_nions = get_max_nions();
  t.includedions = get_includedions();
  int nlev = 0, ndown = 0, nup = 0, ntarg = 0, ntables = 0;
  for (int el = 0; el < t.nelements; el++)
    for (int ion = 0; ion < get_nions(el); ion++)
      for (int lev = 0; lev < get_nlevels(el, ion); lev++) {
        const levellist_entry &L = globals::elements[el].ions[ion].levels[lev];
        nlev++;
        ndown += L.ndowntrans;
        nup += L.nuptrans;
        ntarg += L.nphixstargets;
        if (L.photoion_xs != NULL) ntables++;
      }
  t.nlevels_total = nlev;
  t.ndowntrans_total = ndown;
  t.nuptrans_total = nup;
  t.nphixstargets_total = ntarg;
  t.nphixstables = ntables;
  t.nphixspoints = globals::NPHIXSPOINTS;
  t.nphixsnuincrement = globals::NPHIXSNUINCREMENT;
  t.last_phixs_nuovernuedge = last_phixs_nuovernuedge;
  int *el_anumber = st->I(t.nelements), *el_nions = st->I(t.nelements), *el_ionoffset = st->I(t.nelements);
  const int NI = t.includedions;
  int *ion_ionstage = st->I(NI), *ion_nlevels = st->I(NI), *ion_ionisinglevels = st->I(NI), *ion_maxrec = st->I(NI),
      *ion_leveloffset = st->I(NI), *ion_cooloff = st->I(NI), *ion_ncool = st->I(NI), *ion_firstnlte = st->I(NI),
      *ion_nlevnlte = st->I(NI);
  double *ion_ionpot = st->D(NI);
  double *lev_eps = st->D(nlev), *lev_thr = st->D(nlev);
  float *lev_g = st->F(nlev);
  int *lev_nd = st->I(nlev), *lev_nu = st->I(nlev), *lev_do = st->I(nlev), *lev_uo = st->I(nlev), *lev_nt = st->I(nlev),
      *lev_to = st->I(nlev), *lev_tab = st->I(nlev), *lev_ci = st->I(nlev), *lev_cg = st->I(nlev);
  int *down = st->I(ndown), *up = st->I(nup), *targ_level = st->I(ntarg);
  double *targ_prob = st->D(ntarg);
  float *tables = st->F((size_t)ntables * t.nphixspoints);
  std::map<const float *, int> tablerow;
  int il = 0, id = 0, iu = 0, it = 0, itab = 0, ui = 0;
  for (int el = 0; el < t.nelements; el++) {
    el_anumber[el] = get_element(el);
    el_nions[el] = get_nions(el);
    el_ionoffset[el] = ui;
    for (int ion = 0; ion < get_nions(el); ion++, ui++) {
      const ionlist_entry &I = globals::elements[el].ions[ion];
      assert_always(get_uniqueionindex(el, ion) == ui);
      ion_ionstage[ui] = I.ionstage;
      ion_nlevels[ui] = I.nlevels;
      ion_ionisinglevels[ui] = I.ionisinglevels;
      ion_maxrec[ui] = I.maxrecombininglevel;
      ion_leveloffset[ui] = il;
      ion_cooloff[ui] = I.coolingoffset;
      ion_ncool[ui] = I.ncoolingterms;
      ion_firstnlte[ui] = I.first_nlte;
      ion_nlevnlte[ui] = I.nlevels_nlte;
      ion_ionpot[ui] = I.ionpot;
      for (int lev = 0; lev < I.nlevels; lev++, il++) {
        const levellist_entry &L = I.levels[lev];
        lev_eps[il] = L.epsilon;
        lev_g[il] = L.stat_weight;
        lev_nd[il] = L.ndowntrans;
        lev_nu[il] = L.nuptrans;
        lev_do[il] = id;
        lev_uo[il] = iu;
        for (int i = 0; i < L.ndowntrans; i++) down[id++] = L.downtrans[i].lineindex;
        for (int i = 0; i < L.nuptrans; i++) up[iu++] = L.uptrans[i].lineindex;
        lev_nt[il] = L.nphixstargets;
        lev_to[il] = it;
        for (int i = 0; i < L.nphixstargets; i++) {
          targ_prob[it] = L.phixstargets[i].probability;
          targ_level[it] = L.phixstargets[i].levelindex;
          it++;
        }
        lev_thr[il] = L.phixs_threshold;
        lev_ci[il] = L.cont_index;
#if (!NO_LUT_PHOTOION || !NO_LUT_BFHEATING)
        lev_cg[il] = L.closestgroundlevelcont;
#else
        lev_cg[il] = -1;
#endif
        if (L.photoion_xs != NULL) {
          tablerow[L.photoion_xs] = itab;
          memcpy(tables + (size_t)itab * t.nphixspoints, L.photoion_xs, sizeof(float) * t.nphixspoints);
          lev_tab[il] = itab++;
        } else {
          lev_tab[il] = -1;
        }
      }
    }
  }
  t.el_anumber = el_anumber;
  t.el_nions = el_nions;
  t.el_ionoffset = el_ionoffset;
  t.ion_ionstage = ion_ionstage;
  t.ion_nlevels = ion_nlevels;
  t.ion_ionisinglevels = ion_ionisinglevels;
  t.ion_maxrecombininglevel = ion_maxrec;
  t.ion_leveloffset = ion_leveloffset;
  t.ion_coolingoffset = ion_cooloff;
  t.ion_ncoolingterms = ion_ncool;
  t.ion_first_nlte = ion_firstnlte;
  t.ion_nlevels_nlte = ion_nlevnlte;
  t.ion_ionpot = ion_ionpot;
  t.lev_epsilon = lev_eps;
  t.lev_statweight = lev_g;
  t.lev_ndowntrans = lev_nd;
  t.lev_nuptrans = lev_nu;
  t.lev_downoffset = lev_do;
  t.lev_upoffset = lev_uo;
  t.lev_nphixstargets = lev_nt;
  t.lev_phixstargetoffset = lev_to;
  t.lev_phixs_threshold = lev_thr;
  t.lev_phixstable = lev_tab;
  t.lev_cont_index = lev_ci;
  t.lev_closestgroundlevelcont = lev_cg;
  t.downtrans_lineindex = down;
  t.uptrans_lineindex = up;
  t.phixstarget_prob = targ_prob;
  t.phixstarget_level = targ_level;
  t.phixs_tables = tables;

  // ---- line list
  t.nlines = globals::nlines;
  {
    const int N = t.nlines;
    double *nu = st->D(N);
    float *A = st->F(N), *f = st->F(N), *cs = st->F(N);
    int *e = st->I(N), *i = st->I(N), *u = st->I(N), *l = st->I(N), *fb = st->I(N);
    for (int n = 0; n < N; n++) {
      const linelist_entry &L = globals::linelist[n];
      nu[n] = L.nu;
      A[n] = L.einstein_A;
      f[n] = L.osc_strength;
      cs[n] = L.coll_str;
      e[n] = L.elementindex;
      i[n] = L.ionindex;
      u[n] = L.upperlevelindex;
      l[n] = L.lowerlevelindex;
      fb[n] = L.forbidden ? 1 : 0;
    }
    t.line_nu = nu;
    t.line_einstein_A = A;
    t.line_osc_strength = f;
    t.line_coll_str = cs;
    t.line_element = e;
    t.line_ion = i;
    t.line_upper = u;
    t.line_lower = l;
    t.line_forbidden = fb;
  }
  // ---- continua
  t.nbfcontinua = globals::nbfcontinua;
  {
    const int N = t.nbfcontinua;
    double *nue = st->D(N), *pr = st->D(N);
    int *e = st->I(N), *i = st->I(N), *l = st->I(N), *ti = st->I(N), *ul = st->I(N), *gi = st->I(N), *tb = st->I(N);
    for (int n = 0; n < N; n++) {
      const fullphixslist &C = globals::allcont[n];
      nue[n] = C.nu_edge;
      e[n] = C.element;
      i[n] = C.ion;
      l[n] = C.level;
      ti[n] = C.phixstargetindex;
      ul[n] = C.upperlevel;
      pr[n] = C.probability;
      gi[n] = C.index_in_groundphixslist;
      tb[n] = tablerow.at(C.photoion_xs);
    }
    t.cont_nu_edge = nue;
    t.cont_element = e;
    t.cont_ion = i;
    t.cont_level = l;
    t.cont_phixstargetindex = ti;
    t.cont_upperlevel = ul;
    t.cont_probability = pr;
    t.cont_index_in_groundphixslist = gi;
    t.cont_phixstable = tb;
  }
#if (!NO_LUT_PHOTOION || !NO_LUT_BFHEATING)
  t.nbfcontinua_ground = globals::nbfcontinua_ground;
  {
    const int N = t.nbfcontinua_ground;
    double *nue = st->D(N);
    int *e = st->I(N), *i = st->I(N), *l = st->I(N);
    for (int n = 0; n < N; n++) {
      nue[n] = globals::groundcont[n].nu_edge;
      e[n] = globals::groundcont[n].element;
      i[n] = globals::groundcont[n].ion;
      l[n] = globals::groundcont[n].level;
    }
    t.gcont_nu_edge = nue;
    t.gcont_element = e;
    t.gcont_ion = i;
    t.gcont_level = l;
  }
#else
  t.nbfcontinua_ground = 0;
  t.gcont_nu_edge = st->D(0);
  t.gcont_element = t.gcont_ion = t.gcont_level = st->I(0);
#endif
  // ---- LUTs
  {
    const size_t N = (size_t)TABLESIZE * t.nbfcontinua;
    double *a = st->D(N), *c = st->D(N);
    memcpy(a, globals::spontrecombcoeff, N * sizeof(double));
    memcpy(c, globals::bfcooling_coeff, N * sizeof(double));
    t.spontrecombcoeff = a;
    t.bfcooling_coeff = c;
#if (!NO_LUT_PHOTOION)
    double *b = st->D(N);
    memcpy(b, globals::corrphotoioncoeff, N * sizeof(double));
    t.corrphotoioncoeff = b;
#else
    t.corrphotoioncoeff = st->D(0);
#endif
  }
  // ---- cooling list (kpkt.cc:27-35 via wrap_kpkt.cc)
  t.ncoolingterms = globals::ncoolingterms;
  {
    const int N = t.ncoolingterms;
    int *ty = st->I(N), *lv = st->I(N), *ul = st->I(N);
    for (int n = 0; n < N; n++) refglue::coolinglist_entry(n, &ty[n], &lv[n], &ul[n]);
    t.cool_type = ty;
    t.cool_level = lv;
    t.cool_upperlevel = ul;
  }
  // ---- gamma lines (gammapkt.cc:24-31 via wrap_gammapkt.cc)
  t.nnuclides = decay::get_num_nuclides();
  {
    int *off = st->I(t.nnuclides + 1);
    int tot = 0;
    for (int n = 0; n < t.nnuclides; n++) {
      off[n] = tot;
      tot += refglue::gamma_nlines(n);
    }
    off[t.nnuclides] = tot;
    t.ngammalines_total = tot;
    double *en = st->D(tot), *pr = st->D(tot), *eg = st->D(t.nnuclides);
    for (int n = 0; n < t.nnuclides; n++) {
      for (int k = 0; k < refglue::gamma_nlines(n); k++) refglue::gamma_line(n, k, &en[off[n] + k], &pr[off[n] + k]);
      eg[n] = decay::nucdecayenergygamma(decay::get_nuc_z(n), decay::get_nuc_a(n));
    }
    t.gam_offset = off;
    t.gam_energy = en;
    t.gam_prob = pr;
    t.nuc_decayenergygamma = eg;
  }
  // ---- virtual packets (vpkt.cc:15-74, vpkt.h:27-47)
  t.nprocs = globals::nprocs;
#ifdef VPKT_ON
  {
    VpktParams vp;
    vpkt_params(&vp);
    t.vpkt_nobs = vp.nobs;
    t.vpkt_nspectra = vp.nspectra;
    t.vpkt_nrange = vp.nrange;
    t.vpkt_nrange_grid = vp.vgrid_flag == 1 ? vp.nrange_grid : 0;
    t.vpkt_vgrid_flag = vp.vgrid_flag;
    t.vpkt_vmtbins = VMTBINS;
    t.vpkt_vmnubins = VMNUBINS;
    t.vpkt_ny_vgrid = NY_VGRID;
    t.vpkt_nz_vgrid = NZ_VGRID;
    t.vpkt_tmin_input = vp.tmin_input;
    t.vpkt_tmax_input = vp.tmax_input;
    t.vpkt_tau_max = vp.tau_max;
    t.vpkt_tmin_grid = vp.tmin_grid;
    t.vpkt_tmax_grid = vp.tmax_grid;
    t.vpkt_tmin_vspec = tmin_vspec;
    t.vpkt_tmax_vspec = tmax_vspec;
    t.vpkt_numin_vspec = numin_vspec;
    t.vpkt_numax_vspec = numax_vspec;
    t.vpkt_dlogt = vp.dlogt;
    t.vpkt_dlognu = vp.dlognu;
    auto copyd = [&](const double *src, int n) {
      double *d = st->D(n);
      for (int i = 0; i < n; i++) d[i] = src[i];
      return d;
    };
    t.vpkt_nz_obs = copyd(vp.nz_obs, vp.nobs);
    t.vpkt_phiobs = copyd(vp.phiobs, vp.nobs);
    t.vpkt_exclude = copyd(vp.exclude, vp.nspectra);
    t.vpkt_numin_input = copyd(vp.numin_input, vp.nrange);
    t.vpkt_numax_input = copyd(vp.numax_input, vp.nrange);
    t.vpkt_nu_grid_min = copyd(vp.nu_grid_min, t.vpkt_nrange_grid);
    t.vpkt_nu_grid_max = copyd(vp.nu_grid_max, t.vpkt_nrange_grid);
    float *dt = st->F(VMTBINS), *df = st->F(VMNUBINS);
    for (int n = 0; n < VMTBINS; n++) dt[n] = vpkt_delta_t(n);
    for (int m = 0; m < VMNUBINS; m++) df[m] = vp.delta_freq[m];
    t.vpkt_delta_t = dt;
    t.vpkt_delta_freq = df;
  }
#else
  t.vpkt_nz_obs = t.vpkt_phiobs = t.vpkt_exclude = t.vpkt_numin_input = t.vpkt_numax_input = t.vpkt_nu_grid_min =
      t.vpkt_nu_grid_max = st->D(0);
  t.vpkt_delta_t = t.vpkt_delta_freq = st->F(0);
#endif
  // ---- multi-bin radiation field model (radfield.cc:43, artisoptions nu_lower_first_initial)
  t.radfield_nu_lower_first = nu_lower_first_initial;
  {
    const int nb = MULTIBIN_RADFIELD_MODEL_ON ? RADFIELDBINCOUNT : 0;
    double *up = st->D(nb);
    for (int b = 0; b < nb; b++) up[b] = refglue::radfieldbin_nu_upper(b);
    t.radfieldbin_nu_upper = up;
  }
}

inline void flatten_cellstate(ArtisCellState *C, Storage *st, const int nts) {
  ArtisCellState &c = *C;
  memset(&c, 0, sizeof(c));
  c.nts = nts;
  c.npts_model = grid::get_npts_model();
  c.nelements = get_nelements();
  c.max_nions = get_max_nions();
  c.includedions = get_includedions();
  c.max_path_step = globals::max_path_step;
  c.do_rlc_est = globals::do_rlc_est;
  c.do_comp_est = globals::do_comp_est;
  c.initial_iteration = globals::initial_iteration;
  const int NM = c.npts_model + 1, NE = c.nelements, NI = c.includedions;
  float *Te = st->F(NM), *TR = st->F(NM), *TJ = st->F(NM), *W = st->F(NM), *nne = st->F(NM), *nnetot = st->F(NM),
        *rho = st->F(NM), *kg = st->F(NM), *ffe = st->F(NM);
  int *thick = st->I(NM);
  double *totalcooling = st->D(NM);
  float *ab = st->F((size_t)NM * NE), *glp = st->F((size_t)NM * NI), *pf = st->F((size_t)NM * NI);
  double *cci = st->D((size_t)NM * NI);
  for (int m = 0; m < c.npts_model; m++) {
    const grid::modelgrid_t &M = grid::modelgrid[m];
    Te[m] = M.Te;
    TR[m] = M.TR;
    TJ[m] = M.TJ;
    W[m] = M.W;
    nne[m] = M.nne;
    nnetot[m] = M.nnetot;
    rho[m] = M.rho;
    kg[m] = M.kappagrey;
    ffe[m] = M.ffegrp;
    thick[m] = M.thick;
    totalcooling[m] = M.totalcooling;
    if (grid::get_numassociatedcells(m) <= 0 || M.composition == nullptr) continue;
    for (int el = 0; el < NE; el++) {
      ab[(size_t)m * NE + el] = M.composition[el].abundance;
      for (int ion = 0; ion < get_nions(el); ion++) {
        const int ui = get_uniqueionindex(el, ion);
        glp[(size_t)m * NI + ui] = M.composition[el].groundlevelpop[ion];
        pf[(size_t)m * NI + ui] = M.composition[el].partfunct[ion];
        cci[(size_t)m * NI + ui] = M.cooling_contrib_ion[el][ion];
      }
    }
  }
  c.Te = Te;
  c.TR = TR;
  c.TJ = TJ;
  c.W = W;
  c.nne = nne;
  c.nnetot = nnetot;
  c.rho = rho;
  c.kappagrey = kg;
  c.ffegrp = ffe;
  c.thick = thick;
  c.totalcooling = totalcooling;
  c.elem_abundance = ab;
  c.groundlevelpop = glp;
  c.partfunct = pf;
  c.cooling_contrib_ion = cci;
#if (!NO_LUT_PHOTOION)
  c.ncorrphotoionrenorm = c.npts_model * NE * c.max_nions;
  {
    double *r = st->D(c.ncorrphotoionrenorm);
    memcpy(r, globals::corrphotoionrenorm, sizeof(double) * c.ncorrphotoionrenorm);
    c.corrphotoionrenorm = r;
  }
#else
  c.ncorrphotoionrenorm = 0;
  c.corrphotoionrenorm = st->D(0);
#endif
  c.total_nlte_levels = NLTE_POPS_ON ? globals::total_nlte_levels : 0;
  {
    double *p = st->D(c.total_nlte_levels > 0 ? (size_t)NM * c.total_nlte_levels : 0);
    if (c.total_nlte_levels > 0)
      for (int m = 0; m < c.npts_model; m++)
        if (grid::modelgrid[m].nlte_pops != nullptr)
          memcpy(p + (size_t)m * c.total_nlte_levels, grid::modelgrid[m].nlte_pops,
                 sizeof(double) * c.total_nlte_levels);
    c.nlte_pops = p;
  }
  // ---- radiation-field bin solutions and bf-rate estimators of the previous timestep (radfield.cc:30-45,73)
  const int nonempty = grid::get_nonempty_npts_model();
  c.nradfieldbins = MULTIBIN_RADFIELD_MODEL_ON ? nonempty * RADFIELDBINCOUNT : 0;
  {
    float *bw = st->F(c.nradfieldbins), *bt = st->F(c.nradfieldbins);
    for (int i = 0; i < c.nradfieldbins; i++) refglue::radfieldbin_solution(i, &bw[i], &bt[i]);
    c.radfieldbin_W = bw;
    c.radfieldbin_TR = bt;
  }
  c.nprevbfrate = DETAILED_BF_ESTIMATORS_ON ? nonempty * globals::nbfcontinua : 0;
  {
    float *pb = st->F(c.nprevbfrate);
    for (int i = 0; i < c.nprevbfrate; i++) pb[i] = refglue::prev_bfrate_normed(i);
    c.prev_bfrate_normed = pb;
  }
  // ---- Spencer-Fano solution (nonthermal.cc:95-151); only the numbers the transport path reads
  const bool sf = NT_ON && NT_SOLVE_SPENCERFANO;
  c.nt_nions = NT_ON ? NM * NI : 0;
  c.nt_nauger = NT_MAX_AUGER_ELECTRONS + 1;
  {
    float *ep = st->F(c.nt_nions), *pa = st->F((size_t)c.nt_nions * c.nt_nauger),
          *ea = st->F((size_t)c.nt_nions * c.nt_nauger);
    double *wf = st->D(c.nt_nions), *dep = st->D(NT_ON ? NM : 0), *ntot = st->D(NT_ON ? NM : 0);
    if (NT_ON) {
      for (int m = 0; m < c.npts_model; m++) {
        if (grid::get_numassociatedcells(m) <= 0) continue;
        dep[m] = refglue::nt_deposition_rate_density_raw(m);
        ntot[m] = get_nntot(m);
        for (int el = 0; el < NE; el++)
          for (int ion = 0; ion < get_nions(el); ion++) {
            const int ui = get_uniqueionindex(el, ion);
            const size_t k = (size_t)m * NI + ui;
            if (ion < get_nions(el) - 1) wf[k] = refglue::nt_ratecoeff_wfapprox(m, el, ion);
            if (!sf) continue;
            ep[k] = refglue::nt_eff_ionpot(m, ui);
            for (int a = 0; a < c.nt_nauger; a++) {
              pa[k * c.nt_nauger + a] = refglue::nt_prob_num_auger(m, ui * c.nt_nauger + a, false);
              ea[k * c.nt_nauger + a] = refglue::nt_prob_num_auger(m, ui * c.nt_nauger + a, true);
            }
          }
      }
    }
    c.nt_eff_ionpot = ep;
    c.nt_ratecoeff_wfapprox = wf;
    c.nt_prob_num_auger = pa;
    c.nt_ionenfrac_num_auger = ea;
    c.nt_deposition_rate_density = dep;
    c.nntot = ntot;
  }
}

// Copies the reference's estimator globals (after the reference's own update_packets) into the E-list struct.
inline void flatten_estimators(ArtisEstimators *E, Storage *st, const int nts) {
  ArtisEstimators &e = *E;
  memset(&e, 0, sizeof(e));
  e.npts_model = grid::get_npts_model();
  e.nelements = get_nelements();
  e.max_nions = get_max_nions();
  e.nlines = globals::nlines;
  const int NM = e.npts_model + 1;
  double *J = st->D(NM), *nuJ = st->D(NM), *ff = st->D(NM), *col = st->D(NM), *re = st->D(NM);
  const size_t NG = (size_t)NM * e.nelements * e.max_nions;
  double *ga = st->D(NG), *bf = st->D(NG);
  for (int m = 0; m < e.npts_model; m++) {
    if (grid::get_numassociatedcells(m) <= 0) continue;  // the reference never zeroes the slots of empty cells
    J[m] = refglue::radfield_J()[m];
    nuJ[m] = refglue::radfield_nuJ()[m];
    ff[m] = globals::ffheatingestimator[m];
    col[m] = globals::colheatingestimator[m];
    re[m] = globals::rpkt_emiss[m];
  }
#if (!NO_LUT_PHOTOION)
  memcpy(ga, globals::gammaestimator, sizeof(double) * (size_t)e.npts_model * e.nelements * e.max_nions);
#endif
#if (!NO_LUT_BFHEATING)
  memcpy(bf, globals::bfheatingestimator, sizeof(double) * (size_t)e.npts_model * e.nelements * e.max_nions);
#endif
  double *ts = st->D(16);
  const struct time &T = globals::time_step[nts];
  ts[ARTIS_TS_GAMMA_DEP] = T.gamma_dep;
  ts[ARTIS_TS_POSITRON_DEP] = T.positron_dep;
  ts[ARTIS_TS_ELECTRON_DEP] = T.electron_dep;
  ts[ARTIS_TS_ELECTRON_EMISSION] = T.electron_emission;
  ts[ARTIS_TS_ALPHA_DEP] = T.alpha_dep;
  ts[ARTIS_TS_ALPHA_EMISSION] = T.alpha_emission;
  ts[ARTIS_TS_GAMMA_EMISSION] = T.gamma_emission;
  ts[ARTIS_TS_CMF_LUM] = T.cmf_lum;
  ts[ARTIS_TS_NT_ENERGY_DEPOSITED] = nonthermal::nt_energy_deposited;
  long long *ctr = st->L(ARTIS_CTR_COUNT);
  for (int i = 0; i < stats::COUNTER_COUNT; i++) ctr[i] = stats::get_counter((enum stats::eventcounters)i);
  ctr[ARTIS_CTR_NESC] = globals::nesc;
  ctr[ARTIS_CTR_PELLET_DECAYS] = T.pellet_decays;
  int *ec = st->I(e.nlines), *ac = st->I(e.nlines);
#ifdef RECORD_LINESTAT
  memcpy(ec, globals::ecounter, sizeof(int) * e.nlines);
  memcpy(ac, globals::acounter, sizeof(int) * e.nlines);
#endif
  e.nradfieldbins = MULTIBIN_RADFIELD_MODEL_ON ? grid::get_nonempty_npts_model() * RADFIELDBINCOUNT : 0;
  e.nbfrate = DETAILED_BF_ESTIMATORS_ON ? grid::get_nonempty_npts_model() * globals::nbfcontinua : 0;
  {
    double *bj = st->D(e.nradfieldbins), *bn = st->D(e.nradfieldbins), *br = st->D(e.nbfrate);
    int *bc = st->I(e.nradfieldbins);
    for (int i = 0; i < e.nradfieldbins; i++) refglue::radfieldbin_raw(i, &bj[i], &bn[i], &bc[i]);
    for (int i = 0; i < e.nbfrate; i++) br[i] = refglue::bfrate_raw(i);
    e.radfieldbin_J_raw = bj;
    e.radfieldbin_nuJ_raw = bn;
    e.radfieldbin_contribcount = bc;
    e.bfrate_raw = br;
  }
#ifdef VPKT_ON
  {
    VpktParams vp;
    vpkt_params(&vp);
    const int ncomb = vp.nspectra * vp.nobs;
    e.nvstokes = 3 * VMTBINS * ncomb * VMNUBINS;
    e.nvgrid = vp.vgrid_flag == 1 ? 3 * NY_VGRID * NZ_VGRID * vp.nrange_grid * vp.nobs : 0;
    double *vs = st->D(e.nvstokes), *vg = st->D(e.nvgrid);
    size_t at = 0;
    for (int s3 = 0; s3 < 3; s3++)
      for (int nt = 0; nt < VMTBINS; nt++)
        for (int ic = 0; ic < ncomb; ic++)
          for (int m = 0; m < VMNUBINS; m++) vs[at++] = vpkt_vstokes(s3, nt, ic, m);
    at = 0;
    if (e.nvgrid > 0)
      for (int s3 = 0; s3 < 3; s3++)
        for (int n = 0; n < NY_VGRID; n++)
          for (int m = 0; m < NZ_VGRID; m++)
            for (int r = 0; r < vp.nrange_grid; r++)
              for (int b = 0; b < vp.nobs; b++) vg[at++] = vpkt_vgrid(s3, n, m, r, b);
    e.vstokes = vs;
    e.vgrid = vg;
    int cnt[4];
    vpkt_counts(cnt);
    ts[ARTIS_TS_NVPKT] = cnt[0];
    ts[ARTIS_TS_NVPKT_ESC1] = cnt[1];
    ts[ARTIS_TS_NVPKT_ESC2] = cnt[2];
    ts[ARTIS_TS_NVPKT_ESC3] = cnt[3];
  }
#else
  e.vstokes = st->D(0);
  e.vgrid = st->D(0);
#endif
  e.J = J;
  e.nuJ = nuJ;
  e.ffheatingestimator = ff;
  e.colheatingestimator = col;
  e.rpkt_emiss = re;
  e.gammaestimator = ga;
  e.bfheatingestimator = bf;
  e.ts_scalars = ts;
  e.counters = ctr;
  e.ecounter = ec;
  e.acounter = ac;
}

}  // namespace refglue
#endif
