// vpkt.cuh -- virtual packets (VPKT_ON builds): vpkt.cc:76-406, 581-627, 837-896 of the reference.
//
// At every electron scattering (realtype 1), k-packet -> r-packet conversion (2) and macro-atom de-activation (3) in a
// cell that is not flagged thick for virtual packets, a copy of the packet is sent towards each observer: the optical
// depth to the edge of the grid is integrated cell by cell (continuum opacity of each cell times the path, plus the
// Sobolev depth of every line the ray redshifts through), once per requested spectrum (all opacities; one element's
// lines removed; no lines; no bf / ff / es), and the Stokes vector weighted by exp(-tau) is binned into the observer's
// time- and frequency-resolved spectra and, optionally, a velocity map.
//
// Included by rpkt_events.cuh after the frame transforms (vpkt.cc:898-1069, which it uses). The event handlers leave a
// request in the thread's Tally; vpkt_flush_request runs it after the event. The ray trace takes its packet BY VALUE and
// is not inlined into the transport kernels: the caller's packet stays in registers, and each kernel has one call site.
#pragma once

namespace artis {

constexpr int VPKT_MAX_SPECTRA = 8;  // Nspectra per observer (checked at init)

// vpkt.cc:374-385 check_tau: 0 when every spectrum's optical depth exceeds tau_max
ARTIS_DEV int vpkt_check_tau(const Ctx &c, const double *tau) {
  int count = 0;
  for (int i = 0; i < c.st.vpkt_nspectra; i++)
    if (tau[i] > c.st.vpkt_tau_max) count += 1;
  return (count == c.st.vpkt_nspectra) ? 0 : 1;
}

// vpkt.cc:388-406 add_to_vspecpol; estimator layout [stokes][nt][ind_comb][nnu] (artis_gpu_fields.def)
ARTIS_DEV void add_to_vspecpol(const Ctx &c, const double nu_rf, const double e_rf, const double stokes[3], const int bin,
                               const int ind, const double t_arrive) {
  const int ncomb = c.st.vpkt_nspectra * c.st.vpkt_nobs;
  const int ind_comb = c.st.vpkt_nspectra * bin + ind;
  if (t_arrive > c.st.vpkt_tmin_vspec && t_arrive < c.st.vpkt_tmax_vspec) {
    const int nt = (int)((log(t_arrive) - log(c.st.vpkt_tmin_vspec)) / c.st.vpkt_dlogt);
    if (nu_rf > c.st.vpkt_numin_vspec && nu_rf < c.st.vpkt_numax_vspec) {
      const int nnu = (int)((log(nu_rf) - log(c.st.vpkt_numin_vspec)) / c.st.vpkt_dlognu);
      const double PARSEC = 3.0857e+18;
      const double pktcontrib = e_rf / c.st.vpkt_delta_t[nt] / c.st.vpkt_delta_freq[nnu] / 4.e12 / PI / PARSEC / PARSEC /
                                c.st.nprocs * 4 * PI;
      const size_t plane = (size_t)c.st.vpkt_vmtbins * ncomb * c.st.vpkt_vmnubins;
      const size_t at = ((size_t)nt * ncomb + ind_comb) * c.st.vpkt_vmnubins + nnu;
      est_add(&c.es.vstokes[at], stokes[0] * pktcontrib);
      est_add(&c.es.vstokes[plane + at], stokes[1] * pktcontrib);
      est_add(&c.es.vstokes[2 * plane + at], stokes[2] * pktcontrib);
    }
  }
}

// vpkt.cc:581-627 add_to_vpkt_grid; estimator layout [stokes][nt][mt][bin_range][bin]
ARTIS_DEV void add_to_vpkt_grid(const Ctx &c, const double nu_rf, const double e_rf, const double stokes[3], const double *vel,
                                const int bin_range, const int bin, const double *obs) {
  double vref1, vref2;
  const double nx = obs[0], ny = obs[1], nz = obs[2];
  if (nx == 1) {
    vref1 = vel[1];
    vref2 = vel[2];
  } else if (nx == -1) {
    vref1 = -vel[1];
    vref2 = -vel[2];
  } else {
    vref1 = -ny * vel[0] + (nx + nz * nz / (1 + nx)) * vel[1] - ny * nz * (1 - nx) / sqrt(1 - nx * nx) * vel[2];
    vref2 = -nz * vel[0] - ny * nz * (1 - nx) / sqrt(1 - nx * nx) * vel[1] + (nx + ny * ny / (1 + nx)) * vel[2];
  }
  if (fabs(vref1) >= c.st.vmax || fabs(vref2) >= c.st.vmax) return;
  const double ybin = 2 * c.st.vmax / c.st.vpkt_ny_vgrid;
  const double zbin = 2 * c.st.vmax / c.st.vpkt_nz_vgrid;
  const int nt = (int)((c.st.vmax - vref1) / ybin);
  const int mt = (int)((c.st.vmax - vref2) / zbin);
  if (nu_rf > c.st.vpkt_nu_grid_min[bin_range] && nu_rf < c.st.vpkt_nu_grid_max[bin_range]) {
    const size_t plane = (size_t)c.st.vpkt_ny_vgrid * c.st.vpkt_nz_vgrid * c.st.vpkt_nrange_grid * c.st.vpkt_nobs;
    const size_t at = (((size_t)nt * c.st.vpkt_nz_vgrid + mt) * c.st.vpkt_nrange_grid + bin_range) * c.st.vpkt_nobs + bin;
    est_add(&c.es.vgrid[at], stokes[0] * e_rf);
    est_add(&c.es.vgrid[plane + at], stokes[1] * e_rf);
    est_add(&c.es.vgrid[2 * plane + at], stokes[2] * e_rf);
  }
}

// Continuum opacity as a virtual packet sees it in the reference. get_levelpop reads the per-thread cellhistory, and
// that cache belongs to the cell of the REAL packet for the whole ray: change_cell (boundary.cc:332) does not reset it
// and the cellhistory_reset of vpkt.cc:887 returns at once because the cell number still matches
// (update_grid.cc:665). Every level population along the ray -- in the Sobolev depths and in the bound-free sum -- and
// the cached departure ratios are therefore those of the cell where the ray STARTS, while n_e, T_e, the ion
// populations of the free-free term and the test whether a continuum's element / ion is present are taken from the cell
// the ray is in. The reference's own results are the parity target, so this is reproduced: `tc0` = table row of the
// start cell. A departure ratio that the start cell never needed is evaluated from the start cell's populations with the
// current cell's n_e and T_e, as the reference does on first use (rpkt.cc:1136-1145; later uses of that cached value by
// other cells are not tracked -- it only arises for continua absent from the start cell).
// One term of that bound-free sum (0 when the continuum is skipped): `tc` = table row of the current cell, `tc0` of the
// start cell.
ARTIS_DEV double vpkt_bf_term(const Ctx &c, const int tc, const int tc0, const int mgi, const int i, const double nu, const double T_e,
                              const double expfac) {
  if (c.tb.bf_cell[2 * ((size_t)tc * c.st.nbfcontinua + i)] < 0.) return 0.;  // absent from the current cell
  const int gl = c.cont_level_g[i];
  const double nnlevel = tab_levelpop(c, tc0, gl);
  const ContRec r = c.cont_rec[i];
  if (!(nnlevel > 0.) || !(nu <= r.nu_edge * c.st.last_phixs_nuovernuedge)) return 0.;
  const double *start = c.tb.bf_cell + 2 * ((size_t)tc0 * c.st.nbfcontinua + i);
  double departure_ratio;
  if (start[0] > 0.) {
    departure_ratio = start[1];
  } else {
    const int glu = c.cont_upper_g[i];
    const double sf = calculate_sahafact_g(c, gl, glu, T_e, H * r.nu_edge);
    departure_ratio = tab_levelpop(c, tc0, glu) / nnlevel * (double)c.cs.nne[mgi] * sf;
  }
  const double sigma_bf = phixs_fromtable(c, r.tablerow, r.nu_edge, nu, r.inv_nu_edge);
  double corrfactor = 1 - departure_ratio * expfac;
  if (corrfactor < 0) corrfactor = 0.;
  return nnlevel * sigma_bf * r.probability * corrfactor;
}
// ---- the ray trace, written once for two executions ---------------------------------------------------------------------------
//   COOP = false  one thread traces the ray (the host build, the kernel-per-kind scheduler, and a lane of the persistent kernels
//                 when most lanes of its warp have rays of their own to trace: 32 rays side by side);
//   COOP = true   all 32 lanes of a warp trace ONE ray (device only; the persistent kernels when few lanes have a request).
//                 Every lane carries the same copy of the ray -- the geometry (boundary_cross, move_pkt, the frame transforms)
//                 is computed redundantly and costs what it costs one lane -- and the two loops that are a ray's work are
//                 shared out: the bound-free sum of a cell's continuum opacity, a tile of 32 continua at a time, and the line
//                 sweep of a cell, 32 consecutive list entries at a time (a lane per line evaluates the distance to its
//                 resonance point and its Sobolev depth: five dependent table reads, 32 chains in flight instead of one).
// Both give the same bits: the bound-free sum runs in the tile order of rpkt.cuh (tiles of BF_TILE consecutive continua
// reduced as a binary tree -- PairTree for one thread, shuffle-down 1, 2, 4, 8, 16 for a warp -- tile sums added in order), and
// the Sobolev depths are added IN LIST ORDER with the tau_max test after each one, as vpkt.cc:283-292 does (a warp adds them
// from shuffled operands, every lane the same sums).
//
// Two things the reference does not do, neither of which changes a result:
//  (1) The continuum opacity of the START cell at the packet's frequency is the same for every observer of a request; it is
//      evaluated once (VpktKappa `first`) instead of once per ray. The reference's per-ray cache reset (vpkt.cc:887) would
//      recompute the identical number from identical inputs.
//  (2) Lines first. A ray is abandoned as soon as every spectrum's optical depth exceeds tau_max, and most rays of a
//      supernova model die on the lines of their second or third cell. Before paying for a cell's ~900-term continuum sum the
//      cell's line sweep is run on a COPY of the optical depths without the continuum term: if that alone passes tau_max for
//      every spectrum, so does the reference's own sequence -- floating-point addition is monotone, so adding the
//      (non-negative) continuum term first can only leave every later partial sum at least as large -- and the ray is dead
//      either way: it is dropped without the sum. A ray that survives the probe evaluates the sum and repeats the sweep in
//      the reference's order. The work counters count the real sweep only.
// -DARTIS_VPKT_PROFILE: where a ray's cycles go (measurement builds only): counters[54..56] = cycles in boundary_cross + continuum sum, both line sweeps, set-up + move;
// counters[59] = rays, [60] = cells
#if defined(ARTIS_VPKT_PROFILE) && defined(__CUDA_ARCH__)
#define VP_T0() long long vp_t = clock64()
#define VP_ADD(slot)                                                                       \
  do {                                                                                     \
    const long long vp_n = clock64();                                                      \
    if (lead) atomicAdd((unsigned long long *)&c.es.counters[slot], (unsigned long long)(vp_n - vp_t)); \
    vp_t = vp_n;                                                                           \
  } while (0)
#define VP_COUNT(slot) \
  if (lead) atomicAdd((unsigned long long *)&c.es.counters[slot], 1ull)
#else
#define VP_T0()
#define VP_ADD(slot)
#define VP_COUNT(slot)
#endif
struct VpktKappa {  // the opacity scratch of a ray: KappaCache without the ground-continuum array nobody reads here
  bool valid;
  int mgi;
  double nu, total, es, ff, bf;
};
#ifdef __CUDACC__
__device__ __forceinline__ int vp_popc(const unsigned m) { return __popc(m); }
__device__ __forceinline__ int vp_ctz(const unsigned m) { return __ffs(m) - 1; }
template <bool COOP>
__device__ __forceinline__ int vp_lane() {
  return COOP ? (int)(threadIdx.x & 31) : 0;
}
template <bool COOP>
__device__ __forceinline__ unsigned vp_ballot(const bool pred) {
  return COOP ? __ballot_sync(0xffffffffu, pred) : (pred ? 1u : 0u);
}
template <bool COOP>
__device__ __forceinline__ double vp_shfl(const double v, const int src) {
  return COOP ? __shfl_sync(0xffffffffu, v, src) : v;
}
template <bool COOP>
__device__ __forceinline__ int vp_shfl_i(const int v, const int src) {
  return COOP ? __shfl_sync(0xffffffffu, v, src) : v;
}
#else
inline int vp_popc(const unsigned m) { return __builtin_popcount(m); }
inline int vp_ctz(const unsigned m) { return __builtin_ctz(m); }
template <bool COOP>
inline int vp_lane() {
  return 0;
}
template <bool COOP>
inline unsigned vp_ballot(const bool pred) {
  return pred ? 1u : 0u;
}
template <bool COOP>
inline double vp_shfl(const double v, const int) {
  return v;
}
template <bool COOP>
inline int vp_shfl_i(const int v, const int) {
  return v;
}
#endif

template <bool COOP>
ARTIS_DEV void vpkt_kappa_cont(const Ctx &c, const double nu, const int mgi, const int tc0, VpktKappa &k, Tally &tl) {
  const bool lead = vp_lane<COOP>() == 0;
  if (k.valid && (mgi == k.mgi) && (fabs(k.nu / nu - 1.0) < 1e-4)) return;  // rpkt.cuh kappa_cache_hit
  if (lead) tl.kappa_evals++;
  double sigma = 0., kappa_ff = 0., kappa_bf = 0.;
  if (c.st.do_r_lc) {
    if (c.st.opacity_case == 4) {
      sigma = SIGMA_T * c.cs.nne[mgi];
      kappa_ff = calculate_kappa_ff(c, mgi, nu);
      const int tc = tabcell(c, mgi);
      const double T_e = (double)c.cs.Te[mgi];
      const double expfac = exp(-HOVERKB * nu / T_e);
      const int i0 = bf_stop_index(c, nu);
      const int ilo = bf_start_index(c, nu, i0);
      if (lead) tl.bf_evaluated += i0 - ilo;
#ifdef __CUDACC__
      if constexpr (COOP) {
        const int lane = vp_lane<COOP>();
        for (int base = ilo; base < i0; base += BF_TILE) {
          const int i = base + lane;
          double t = (i < i0) ? vpkt_bf_term(c, tc, tc0, mgi, i, nu, T_e, expfac) : 0.;
#pragma unroll
          for (int off = 1; off < BF_TILE; off <<= 1) t += __shfl_down_sync(0xffffffffu, t, off);
          kappa_bf += t;  // lane 0 holds the tile sums in tree order
        }
        kappa_bf = __shfl_sync(0xffffffffu, kappa_bf, 0);
      } else
#endif
      {
        const int iend = ilo + ((i0 - ilo + BF_TILE - 1) / BF_TILE) * BF_TILE;  // the last tile is padded with zero terms
        PairTree tree;
        for (int i = ilo; i < iend; i++) {
          const double t = (i < i0) ? vpkt_bf_term(c, tc, tc0, mgi, i, nu, T_e, expfac) : 0.;
          double tile_sum;
          if (tree.push((i - ilo) & (BF_TILE - 1), t, &tile_sum)) kappa_bf += tile_sum;
        }
      }
    } else {
      kappa_ff = 1e5 * calculate_kappa_ff(c, mgi, nu);
    }
  }
  k.nu = nu;  // rpkt.cuh kappa_store
  k.mgi = mgi;
  k.valid = true;
  k.total = sigma + kappa_bf + kappa_ff;
  k.es = sigma;
  k.ff = kappa_ff;
  k.bf = kappa_bf;
  if (!isfinite(k.total)) {  // rpkt.cc:1280-1294 soft fallback
    if (isfinite(k.es)) {
      k.ff = 0.;
      k.bf = 0.;
      k.total = k.es;
      if (lead) tl.soft_repairs++;
    } else {
      fatal(c, 101);
    }
  }
}

// vpkt.cc:254-303, the `while (ldist < sdist)` loop of one cell: adds the Sobolev depth of every line the ray redshifts through
// on its way across the cell (path sdist, entered at co-moving frequency nu_cmf) to the spectra's optical depths, in list
// order from list entry next_trans on; returns true as soon as every spectrum has passed tau_max (the ray is dead).
// W list entries per round: 1 for one thread, 32 for a warp.
template <bool COOP>
ARTIS_DEV bool vpkt_line_sweep(const Ctx &c, const double nu_cmf, const double t_current, const double sdist, const int tc0,
                               int &next_trans, double *tau_vpkt, Tally &tl, const bool count) {
  if (!(0. < sdist)) return false;
  const int lane = vp_lane<COOP>();
  const int nspectra = c.st.vpkt_nspectra;
  const int right = c.st.nlines - 1;
  while (true) {
    // closest_transition (rpkt.cuh): no line left at this frequency, or the list is exhausted. (The reference stores
    // nlines + 1 and keeps looping with ldist unchanged, vpkt.cc:297-300, which does not terminate; it is only reached by a
    // packet redder than the reddest line. No line is left in this or any later cell: the sweep ends.)
    if (nu_cmf < c.st.line_nu[right] || next_trans > right) {
      next_trans = c.st.nlines + 1;
      return false;
    }
    int first = next_trans;
    if (first <= 0) first = (nu_cmf >= c.st.line_nu[0]) ? 0 : linelist_lower_bound(c, first, nu_cmf);
    const int lineindex = first + lane;
    const bool inlist = lineindex <= right;
    double nutrans = 1., ldist = 0.;
    if (inlist) {
      nutrans = c.st.line_nu[lineindex];
      ldist = (nu_cmf < nutrans) ? 0. : CLIGHT * t_current * (nu_cmf / nutrans - 1);
    }
    const int nlist = vp_popc(vp_ballot<COOP>(inlist));
    // the entry that ends the sweep: the first one beyond the cell (not consumed) or exactly on its edge (consumed)
    const unsigned stop = vp_ballot<COOP>(inlist && !(ldist < sdist));
    int consumed = nlist;
    bool ends = false, beyond = false;
    if (stop != 0u) {
      const int s = vp_ctz(stop);
      beyond = vp_shfl<COOP>(ldist, s) > sdist;
      consumed = beyond ? s : s + 1;
      ends = true;
    }
    double dtau = 0.;
    int anumber = 0;
    if (lane < consumed) {
      const double t_line = t_current + ldist / CLIGHT;
      const int glu = c.line_upper_g[lineindex], gll = c.line_lower_g[lineindex];
      const double A_ul = (double)c.st.line_einstein_A[lineindex];
      const double B_ul = CLIGHTSQUAREDOVERTWOH / pow(nutrans, 3) * A_ul;
      const double B_lu = (double)c.st.lev_statweight[glu] / (double)c.st.lev_statweight[gll] * B_ul;
      const double n_u = tab_levelpop(c, tc0, glu);
      const double n_l = tab_levelpop(c, tc0, gll);
      anumber = c.st.el_anumber[c.st.line_element[lineindex]];
      dtau = (B_lu * n_l - B_ul * n_u) * HCLIGHTOVERFOURPI * t_line;
    }
    if (count && lane == 0) {
      tl.lines_scanned += consumed + (beyond ? 1 : 0);
      tl.lines_tau += consumed;
    }
    for (int j = 0; j < consumed; j++) {  // in list order, with the test after every line
      const double v = vp_shfl<COOP>(dtau, j);
      const int a = vp_shfl_i<COOP>(anumber, j);
      for (int ind = 0; ind < nspectra; ind++) {
        const double ex = c.st.vpkt_exclude[ind];
        if (ex != -1 && ((double)a != ex)) tau_vpkt[ind] += v;
      }
      if (vpkt_check_tau(c, tau_vpkt) == 0) return true;
    }
    next_trans = first + consumed;
    if (ends) return false;
  }
}

// ---- vpkt.cc:76-364 rlc_emiss_vpkt in three pieces (the one-thread / one-warp ray below runs them in a row, the lock-step
// tracer of the request list runs each piece for all its lanes at once) ----------------------------------------------------------
// The Stokes vector a ray starts with and its probability density towards the observer (vpkt.cc:103-196)
struct VpktStokes {
  double I, Q, U, pn;
};
// set-up: `dummy` becomes the packet sent towards `obs` (which is rewritten with the rest-frame direction for a scattering,
// vpkt.cc:178), `vel_vec` the velocity of the start position
ARTIS_DEV void vpkt_ray_setup(const Ctx &c, const Pkt &pk, const double stokes_q, const double stokes_u, const double t_current,
                              double *obs, const int realtype, Pkt &dummy, VpktStokes &sv, double *vel_vec) {
  dummy = pk;
  dummy.dir[0] = obs[0];
  dummy.dir[1] = obs[1];
  dummy.dir[2] = obs[2];
  double old_dir_cmf[3], obs_cmf[3], vel_rev[3], ref1[3], ref2[3];
  get_velocity(pk.pos, vel_vec, t_current);
  dummy.nu_rf = dummy.nu_cmf / doppler_nucmf_on_nurf(c, dummy.dir, vel_vec);
  dummy.e_rf = dummy.e_cmf * dummy.nu_rf / dummy.nu_cmf;
  double Qi = stokes_q, Ui = stokes_u;  // Stokes Q, U of the real packet (used by the scattering phase function only)
  sv.I = 0.;
  sv.Q = 0.;
  sv.U = 0.;
  sv.pn = 0.;
  if (realtype == 1) {  // electron scattering: dipole phase function and Stokes rotation (Bulla et al. 2015)
    double dir_in[3] = {pk.dir[0], pk.dir[1], pk.dir[2]};
    frame_transform(dir_in, &Qi, &Ui, vel_vec, old_dir_cmf);
    angle_ab(dummy.dir, vel_vec, obs_cmf);
    meridian(old_dir_cmf, ref1, ref2);
    const double i1 = rot_angle(old_dir_cmf, obs_cmf, ref1, ref2);
    const double cos2i1 = cos(2 * i1), sin2i1 = sin(2 * i1);
    const double Qold = Qi * cos2i1 - Ui * sin2i1;
    const double Uold = Qi * sin2i1 + Ui * cos2i1;
    const double mu = dot3(old_dir_cmf, obs_cmf);
    sv.pn = 3. / (16. * PI) * (1 + mu * mu + (mu * mu - 1) * Qold);  // pow(mu, 2.) in the reference
    const double Inew = 0.75 * ((mu * mu + 1.0) + Qold * (mu * mu - 1.0));
    double Qnew = 0.75 * ((mu * mu - 1.0) + Qold * (mu * mu + 1.0));
    double Unew = 1.5 * mu * Uold;
    Qnew = Qnew / Inew;
    Unew = Unew / Inew;
    sv.I = Inew / Inew;
    meridian(obs_cmf, ref1, ref2);
    const double i2 = PI + rot_angle(obs_cmf, old_dir_cmf, ref1, ref2);
    const double cos2i2 = cos(2 * i2), sin2i2 = sin(2 * i2);
    sv.Q = Qnew * cos2i2 + Unew * sin2i2;
    sv.U = -Qnew * sin2i2 + Unew * cos2i2;
    vel_rev[0] = -vel_vec[0];
    vel_rev[1] = -vel_vec[1];
    vel_rev[2] = -vel_vec[2];
    frame_transform(obs_cmf, &sv.Q, &sv.U, vel_rev, obs);  // writes the rest-frame direction back into obs (vpkt.cc:178)
  }
  if (realtype == 2 || realtype == 3) {  // isotropic emission
    sv.I = 1;
    sv.Q = 0;
    sv.U = 0;
    sv.pn = 1 / (4 * PI);
  }
}
// the continuum term of a cell on every spectrum's optical depth (vpkt.cc:222-250); returns true when the ray is dead
ARTIS_DEV bool vpkt_add_continuum(const Ctx &c, const VpktKappa &k, const double s_cont, double *tau_vpkt) {
  const double kap_cont = k.total;
  const double kap_cont_nobf = kap_cont - k.bf, kap_cont_noff = kap_cont - k.ff, kap_cont_noes = kap_cont - k.es;
  for (int ind = 0; ind < c.st.vpkt_nspectra; ind++) {
    const double ex = c.st.vpkt_exclude[ind];
    if (ex == -2)
      tau_vpkt[ind] += kap_cont_nobf * s_cont;
    else if (ex == -3)
      tau_vpkt[ind] += kap_cont_noff * s_cont;
    else if (ex == -4)
      tau_vpkt[ind] += kap_cont_noes * s_cont;
    else
      tau_vpkt[ind] += kap_cont * s_cont;
  }
  return vpkt_check_tau(c, tau_vpkt) == 0;
}
// the end of a ray that has left the model: its Stokes vector, binned on the fly (vpkt.cc:336-362)
ARTIS_DEV void vpkt_ray_bin(const Ctx &c, const Pkt &pk, const Pkt &dummy, const double *tau_vpkt, const VpktStokes &sv,
                            const double *vel_vec, const int bin, const double *obs, const double t_current) {
  const double t_arrive = t_current - (dot3(pk.pos, dummy.dir) / CLIGHT_PROP);
  for (int ind = 0; ind < c.st.vpkt_nspectra; ind++) {
    const double prob = sv.pn * exp(-tau_vpkt[ind]);
    const double stokes[3] = {sv.I * prob, sv.Q * prob, sv.U * prob};
    add_to_vspecpol(c, dummy.nu_rf, dummy.e_rf, stokes, bin, ind, t_arrive);
  }
  if (c.st.vpkt_vgrid_flag == 1) {
    const double prob = sv.pn * exp(-tau_vpkt[0]);
    const double stokes[3] = {sv.I * prob, sv.Q * prob, sv.U * prob};
    for (int bin_range = 0; bin_range < c.st.vpkt_nrange_grid; bin_range++) {
      if (dummy.nu_rf > c.st.vpkt_nu_grid_min[bin_range] && dummy.nu_rf < c.st.vpkt_nu_grid_max[bin_range]) {
        if (t_arrive > c.st.vpkt_tmin_grid && t_arrive < c.st.vpkt_tmax_grid)
          add_to_vpkt_grid(c, dummy.nu_rf, dummy.e_rf, stokes, vel_vec, bin_range, bin, obs);
      }
    }
  }
}

// One ray from set-up to binning. `pk` is the real packet (read only), `k` the opacity scratch of the request, `first` the
// start cell's opacity shared by the request's rays. Returns the change of the virtual-packet counters: bit 0 = a virtual
// packet was created, bit 1 = it reached the edge of the grid. COOP: warp-uniform arguments and result.
template <bool COOP>
ARTIS_DEV_NOINLINE int rlc_emiss_vpkt(const Ctx &c, const Pkt &pk, const double stokes_q, const double stokes_u,
                                      const double t_current, const int bin, double *obs, const int realtype, VpktKappa &k,
                                      VpktKappa &first, Tally &tl) {
  const bool lead = vp_lane<COOP>() == 0;
  VP_T0();
  VP_COUNT(59);
  double tau_vpkt[VPKT_MAX_SPECTRA];
  for (int ind = 0; ind < VPKT_MAX_SPECTRA; ind++) tau_vpkt[ind] = 0;
  Pkt dummy;
  VpktStokes sv;
  double vel_vec[3];
  vpkt_ray_setup(c, pk, stokes_q, stokes_u, t_current, obs, realtype, dummy, sv, vel_vec);
  bool end_packet = false;
  double t_future = t_current;
  int mgi = cell_mgi(c, dummy.where);
  const int tc0 = tabcell(c, mgi);  // level populations of the whole ray come from the start cell (see vpkt_bf_term)
  bool at_start = true;
  k.valid = false;  // cellhistory_reset(mgi, false) at vpkt.cc:887: a ray starts without a cached opacity ...
  if (first.valid) k = first;  // ... and would evaluate exactly this one first
  VP_ADD(56);  // (set-up shares the slot of the move: 51 is ARTIS_CTR_BF_LISTED)
  while (!end_packet) {
    VP_COUNT(60);
    int snext = 0;
    const double sdist = boundary_cross(c, dummy, &snext);
    const double s_cont = sdist * t_current * t_current * t_current / (t_future * t_future * t_future);
    VP_ADD(54);  // (boundary_cross shares the slot of the continuum sum: 52 / 53 are the thin role's cycle counters)
    {  // lines first (see above): the cell's sweep on a copy, without the continuum term
      double tau_probe[VPKT_MAX_SPECTRA];
      for (int ind = 0; ind < VPKT_MAX_SPECTRA; ind++) tau_probe[ind] = tau_vpkt[ind];
      int next_probe = dummy.next_trans;
      const bool dead = vpkt_line_sweep<COOP>(c, dummy.nu_cmf, t_current, sdist, tc0, next_probe, tau_probe, tl, false);
      VP_ADD(55);  // (probe sweep shares the slot of the real sweep)
      if (dead) return 1;
    }
    vpkt_kappa_cont<COOP>(c, dummy.nu_cmf, mgi, tc0, k, tl);
    VP_ADD(54);
    if (at_start) {
      first = k;
      at_start = false;
    }
    if (vpkt_add_continuum(c, k, s_cont, tau_vpkt)) return 1;
    {
      const bool dead = vpkt_line_sweep<COOP>(c, dummy.nu_cmf, t_current, sdist, tc0, dummy.next_trans, tau_vpkt, tl, true);
      VP_ADD(55);
      if (dead) return 1;
    }
    t_future += (sdist / CLIGHT_PROP);
    dummy.prop_time = t_future;
    move_pkt(c, dummy, sdist);
    if (snext == -99) {  // change_cell (boundary.cc:332-357). The reference also bumps globals::nesc for a virtual
      end_packet = true; // packet leaving the grid; here COUNTER_NESC counts real packets only
    } else {
      dummy.where = snext;
      if (lead) stat_inc(tl, COUNTER_CELLCROSSINGS);
    }
    mgi = cell_mgi(c, dummy.where);
    VP_ADD(56);
    if (mgi == c.st.npts_model) break;        // reached an empty cell
    if (c.cs.thick[mgi] == 1) return 1;       // a virtual packet dies in a grey cell
  }
  if (lead) vpkt_ray_bin(c, pk, dummy, tau_vpkt, sv, vel_vec, bin, obs, t_current);
  return 3;
}

// vpkt.cc:837-896 vpkt_call_estimators for one interaction of the real packet `pk` (read only -- the one field the reference
// modifies, next_trans, is rewritten with the value it already has). Returns 1 when at least one virtual packet was sent
// (the caller then drops its own cached opacity, as cellhistory_reset + calculate_kappa_rpkt_cont at vpkt.cc:887-893
// re-evaluate it at the packet's exact frequency). COOP: warp-uniform arguments and result; lane 0 counts.
template <bool COOP>
ARTIS_DEV int vpkt_call_estimators(const Ctx &c, const Pkt &pk, const double stokes_q, const double stokes_u, const double t_current,
                                   const int realtype, Tally &tl) {
  const bool lead = vp_lane<COOP>() == 0;
  const int mgi = cell_mgi(c, pk.where);
  if (c.cs.thick[mgi] != 0) return 0;
  VpktKappa k, first;
  k.valid = false;
  first.valid = false;
  double vel_vec[3];
  get_velocity(pk.pos, vel_vec, t_current);
  int vflag = 0, created = 0, escaped = 0;
  for (int bin = 0; bin < c.st.vpkt_nobs; bin++) {
    double obs[3];
    const double nz = c.st.vpkt_nz_obs[bin];
    obs[0] = sqrt(1 - nz * nz) * cos(c.st.vpkt_phiobs[bin]);
    obs[1] = sqrt(1 - nz * nz) * sin(c.st.vpkt_phiobs[bin]);
    obs[2] = nz;
    const double t_arrive = t_current - (dot3(pk.pos, obs) / CLIGHT_PROP);
    if (t_arrive >= c.st.vpkt_tmin_input && t_arrive <= c.st.vpkt_tmax_input) {
      for (int i = 0; i < c.st.vpkt_nrange; i++) {
        const double nu_rf = pk.nu_cmf / doppler_nucmf_on_nurf(c, obs, vel_vec);
        if (nu_rf > c.st.vpkt_numin_input[i] && nu_rf < c.st.vpkt_numax_input[i]) {
          const int r = rlc_emiss_vpkt<COOP>(c, pk, stokes_q, stokes_u, t_current, bin, obs, realtype, k, first, tl);
          created += (r & 1);
          escaped += (r >> 1) & 1;
          vflag = 1;
        }
      }
    }
  }
  if (lead) {
    if (created) est_add(&c.es.ts_scalars[ARTIS_TS_NVPKT], (double)created);
    if (escaped) est_add(&c.es.ts_scalars[ARTIS_TS_NVPKT_ESC1 + realtype - 1], (double)escaped);
  }
  return vflag;
}

// The events only leave a request (Tally.vpkt_req): the reference calls vpkt_call_estimators at the end of the emitting
// routine (macroatom.cc:292,376, kpkt.cc:631,691 -- nothing the ray depends on changes afterwards) or just before the
// electron scattering changes the direction and the Stokes vector (rpkt.cc:359 -- those five numbers travel with the
// request). Each scheduler calls this once, right after an event returns, so that the transport kernels hold a single
// call site of the out-of-line ray trace.
ARTIS_DEV void vpkt_flush_request(const Ctx &c, const Pkt &p, KappaCache &k, Tally &tl) {
  const int realtype = tl.vpkt_req;
  if (realtype == 0) return;
  tl.vpkt_req = 0;
  Pkt q = p;
  double stokes_q = 0., stokes_u = 0.;  // only an electron scattering (realtype 1) looks at the incoming polarisation
  if (realtype == 1) {
    q.dir[0] = tl.vpkt_dir[0];
    q.dir[1] = tl.vpkt_dir[1];
    q.dir[2] = tl.vpkt_dir[2];
    stokes_q = tl.vpkt_q;
    stokes_u = tl.vpkt_u;
  }
  if (vpkt_call_estimators<false>(c, q, stokes_q, stokes_u, q.prop_time, realtype, tl)) k.valid = false;
}

// vpkt_call_estimators without the rays: would this interaction send at least one virtual packet? (the thick-cell test and
// the observers' time and frequency windows, vpkt.cc:850-880)
ARTIS_DEV bool vpkt_wanted(const Ctx &c, const Pkt &pk, const double t_current) {
  const int mgi = cell_mgi(c, pk.where);
  if (c.cs.thick[mgi] != 0) return false;
  double vel_vec[3];
  get_velocity(pk.pos, vel_vec, t_current);
  for (int bin = 0; bin < c.st.vpkt_nobs; bin++) {
    double obs[3];
    const double nz = c.st.vpkt_nz_obs[bin];
    obs[0] = sqrt(1 - nz * nz) * cos(c.st.vpkt_phiobs[bin]);
    obs[1] = sqrt(1 - nz * nz) * sin(c.st.vpkt_phiobs[bin]);
    obs[2] = nz;
    const double t_arrive = t_current - (dot3(pk.pos, obs) / CLIGHT_PROP);
    if (t_arrive >= c.st.vpkt_tmin_input && t_arrive <= c.st.vpkt_tmax_input) {
      for (int i = 0; i < c.st.vpkt_nrange; i++) {
        const double nu_rf = pk.nu_cmf / doppler_nucmf_on_nurf(c, obs, vel_vec);
        if (nu_rf > c.st.vpkt_numin_input[i] && nu_rf < c.st.vpkt_numax_input[i]) return true;
      }
    }
  }
  return false;
}

#ifdef __CUDACC__
// ---- deferred ray trace (the barrier-free scheduler's default) ------------------------------------------------------------------
// Nothing a real packet does afterwards depends on its virtual packets, and nothing a virtual packet reads changes within a
// timestep (cell state, level populations, the packet as it was at the event). So the event only RECORDS the request -- the
// packet's position, frequency, energy, arrival direction and Stokes Q / U, 128 bytes -- and goes on; when the scheduler has
// finished, k_vpkt_trace (transport.cu) traces the whole list with every warp of the device running the same few KB of
// ray-trace code. In the transport kernel the same rays cost 4-5x as much: its warps are few (168 registers), they interleave
// the ray trace with five other kinds of event code, and they wait for instructions more than for data (profiles/r2_notes.md).
// Warp-uniform call; returns true on the lanes whose event sent virtual packets (they drop their cached opacity, as the
// reference re-evaluates it at vpkt.cc:887-893).
__device__ __forceinline__ bool vpkt_enqueue(const Ctx &c, const bool have, const Pkt &p, Tally &tl) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int realtype = have ? tl.vpkt_req : 0;
  tl.vpkt_req = 0;
  if (__ballot_sync(full, realtype != 0) == 0u) return false;
  const bool send = realtype != 0 && vpkt_wanted(c, p, p.prop_time);
  const unsigned m = __ballot_sync(full, send);
  if (m == 0u) return false;
  const int leader = __ffs(m) - 1;
  unsigned long long base = 0ull;
  if (lane == leader) {
    base = atomicAdd(c.vreq_n, (unsigned long long)__popc(m));
    if (base + (unsigned long long)__popc(m) > c.vreq_soft) *reinterpret_cast<volatile int *>(c.vreq_stop) = 1;
  }
  base = __shfl_sync(full, base, leader);
  if (send) {
    const unsigned long long at = base + (unsigned long long)__popc(m & ((1u << lane) - 1u));
    if (at < c.vreq_cap) {  // always, by the margin between vreq_soft and vreq_cap (engine.inc)
      VpktReq r;
      r.where = p.where;
      r.last_cross = p.last_cross;
      r.next_trans = p.next_trans;
      r.realtype = realtype;
      for (int d = 0; d < 3; d++) {
        r.pos[d] = p.pos[d];
        r.dir[d] = (realtype == 1) ? tl.vpkt_dir[d] : p.dir[d];
      }
      r.e_cmf = p.e_cmf;
      r.nu_cmf = p.nu_cmf;
      r.prop_time = p.prop_time;
      r.stokes_q = (realtype == 1) ? tl.vpkt_q : 0.;
      r.stokes_u = (realtype == 1) ? tl.vpkt_u : 0.;
      r.pad_[0] = r.pad_[1] = r.pad_[2] = 0.;
      const uint4 *src = reinterpret_cast<const uint4 *>(&r);
      uint4 *dst = reinterpret_cast<uint4 *>(c.vreq + at);
#pragma unroll
      for (int q = 0; q < 8; q++) __stcg(dst + q, src[q]);
    } else {
      fatal(c, 950);
    }
  }
  return send;
}
// one request of the list, by one thread (COOP = false) or by the 32 lanes of a warp together (COOP = true)
template <bool COOP>
__device__ __forceinline__ void vpkt_trace_request(const Ctx &c, const VpktReq &r, Tally &tl) {
  Pkt q;
  q.where = r.where;
  q.type = ARTIS_TYPE_RPKT;
  q.last_cross = r.last_cross;
  q.interactions = 0;
  q.last_event = 0;
  q.next_trans = r.next_trans;
  q.number = 0;
  q.ma_element = q.ma_ion = q.ma_level = q.ma_activatingline = 0;
  for (int d = 0; d < 3; d++) {
    q.pos[d] = r.pos[d];
    q.dir[d] = r.dir[d];
  }
  q.e_cmf = r.e_cmf;
  q.e_rf = 0.;  // a ray sets its own rest-frame energy and frequency (rlc_emiss_vpkt)
  q.nu_cmf = r.nu_cmf;
  q.nu_rf = 0.;
  q.prop_time = r.prop_time;
  q.cold = nullptr;
  vpkt_call_estimators<COOP>(c, q, r.stokes_q, r.stokes_u, r.prop_time, r.realtype, tl);
}

// ---- the request list traced in lock step (k_vpkt_trace, transport.cu) -----------------------------------------------------------
// A thread per request that simply calls vpkt_call_estimators runs at 3.7 active lanes per instruction
// (profiles/r2_c5_vpkt_trace_v1_ncu_summary.txt): the 32 rays of a warp are each somewhere else in their cell loop -- one in a
// continuum sum, one in a line sweep, one crossing a boundary -- and the hardware runs the groups one after the other. Here a
// lane still owns a request and traces its rays one after the other, but the WARP moves through the phases of a cell
// together: (A) lanes without a request take the next one of the list, (B) lanes without a ray set up the next observer's,
// (C) every ray finds its cell boundary, (D) runs its lines-first probe, (E) the rays that need a continuum opacity get
// their bound-free sums from the whole warp, an owner at a time, 32 continua per round (coalesced rows, tile-tree order),
// (F) continuum term and the real line sweep, (G) the move into the next cell or the end of the ray. Same arithmetic per ray
// as rlc_emiss_vpkt<false>, in the same order: the same bits.
__device__ __forceinline__ void vpkt_trace_lockstep(const Ctx &c, const VpktReq *req, const long long n, unsigned long long *cursor,
                                                    Tally &tl) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int nobs = c.st.vpkt_nobs, nrange = c.st.vpkt_nrange;
  // the lane's request
  bool has_req = false, exhausted = false;
  Pkt pk;
  double req_q = 0., req_u = 0.;
  int realtype = 0, next_pair = 0, created = 0, escaped = 0;
  VpktKappa first;
  first.valid = false;
  // the lane's ray
  bool ray = false;
  Pkt d;
  VpktStokes sv;
  VpktKappa k;
  k.valid = false;
  double tau_vpkt[VPKT_MAX_SPECTRA], vel_vec[3], obs[3], t_future = 0.;
  int mgi = 0, tc0 = 0, bin = 0;
  bool at_start = false;
  while (true) {
    // (A) requests
    const unsigned want = __ballot_sync(full, !has_req && !exhausted);
    if (want != 0u) {
      unsigned long long base = 0ull;
      if (lane == __ffs(want) - 1) base = atomicAdd(cursor, (unsigned long long)__popc(want));
      base = __shfl_sync(full, base, __ffs(want) - 1);
      if (!has_req && !exhausted) {
        const long long idx = (long long)base + __popc(want & lt_mask);
        if (idx < n) {
          VpktReq r;
          const uint4 *src = reinterpret_cast<const uint4 *>(req + idx);
          uint4 *dst = reinterpret_cast<uint4 *>(&r);
#pragma unroll
          for (int q = 0; q < 8; q++) dst[q] = __ldcg(src + q);
          pk.where = r.where;
          pk.type = ARTIS_TYPE_RPKT;
          pk.last_cross = r.last_cross;
          pk.interactions = 0;
          pk.last_event = 0;
          pk.next_trans = r.next_trans;
          pk.number = 0;
          pk.ma_element = pk.ma_ion = pk.ma_level = pk.ma_activatingline = 0;
          for (int q = 0; q < 3; q++) {
            pk.pos[q] = r.pos[q];
            pk.dir[q] = r.dir[q];
          }
          pk.e_cmf = r.e_cmf;
          pk.e_rf = 0.;
          pk.nu_cmf = r.nu_cmf;
          pk.nu_rf = 0.;
          pk.prop_time = r.prop_time;
          pk.cold = nullptr;
          req_q = r.stokes_q;
          req_u = r.stokes_u;
          realtype = r.realtype;
          has_req = true;
          next_pair = (c.cs.thick[cell_mgi(c, pk.where)] != 0) ? nobs * nrange : 0;  // vpkt.cc:846
          created = escaped = 0;
          first.valid = false;
        } else {
          exhausted = true;
        }
      }
    }
    if (__ballot_sync(full, has_req) == 0u) break;  // the list is done and so is every ray of this warp
    // (B) the next observer (and frequency range) whose windows the request passes: vpkt.cc:850-880
    bool starting = false;
    if (has_req && !ray) {
      const double t_current = pk.prop_time;
      double vel0[3];
      get_velocity(pk.pos, vel0, t_current);
      while (next_pair < nobs * nrange) {
        const int b = next_pair / nrange, i = next_pair - b * nrange;
        next_pair++;
        const double nz = c.st.vpkt_nz_obs[b];
        obs[0] = sqrt(1 - nz * nz) * cos(c.st.vpkt_phiobs[b]);
        obs[1] = sqrt(1 - nz * nz) * sin(c.st.vpkt_phiobs[b]);
        obs[2] = nz;
        const double t_arrive = t_current - (dot3(pk.pos, obs) / CLIGHT_PROP);
        if (t_arrive >= c.st.vpkt_tmin_input && t_arrive <= c.st.vpkt_tmax_input) {
          const double nu_rf = pk.nu_cmf / doppler_nucmf_on_nurf(c, obs, vel0);
          if (nu_rf > c.st.vpkt_numin_input[i] && nu_rf < c.st.vpkt_numax_input[i]) {
            bin = b;
            starting = true;
            break;
          }
        }
      }
      if (!starting) {  // the request is finished
        if (created) est_add(&c.es.ts_scalars[ARTIS_TS_NVPKT], (double)created);
        if (escaped) est_add(&c.es.ts_scalars[ARTIS_TS_NVPKT_ESC1 + realtype - 1], (double)escaped);
        has_req = false;
      }
    }
    __syncwarp(full);
    if (starting) {
      for (int ind = 0; ind < VPKT_MAX_SPECTRA; ind++) tau_vpkt[ind] = 0;
      vpkt_ray_setup(c, pk, req_q, req_u, pk.prop_time, obs, realtype, d, sv, vel_vec);
      t_future = pk.prop_time;
      mgi = cell_mgi(c, d.where);
      tc0 = tabcell(c, mgi);
      at_start = true;
      k.valid = false;
      if (first.valid) k = first;
      created++;
      ray = true;
    }
    __syncwarp(full);
    if (__ballot_sync(full, ray) == 0u) continue;
    const double t_current = pk.prop_time;
    // (C) the cell boundary
    int snext = 0;
    double sdist = 0., s_cont = 0.;
    if (ray) {
      sdist = boundary_cross(c, d, &snext);
      s_cont = sdist * t_current * t_current * t_current / (t_future * t_future * t_future);
    }
    __syncwarp(full);
    // (D) lines first
    bool dead = false;
    if (ray) {
      double tau_probe[VPKT_MAX_SPECTRA];
      for (int ind = 0; ind < VPKT_MAX_SPECTRA; ind++) tau_probe[ind] = tau_vpkt[ind];
      int next_probe = d.next_trans;
      dead = vpkt_line_sweep<false>(c, d.nu_cmf, t_current, sdist, tc0, next_probe, tau_probe, tl, false);
    }
    __syncwarp(full);
    // (E) the continuum opacity of the cell: vpkt_kappa_cont<false> with the bound-free sum shared out
    {
      const double nu = d.nu_cmf;
      const bool needk = ray && !dead && !(k.valid && (mgi == k.mgi) && (fabs(k.nu / nu - 1.0) < 1e-4));
      double sigma = 0., kappa_ff = 0., kappa_bf = 0., T_e = 1., expfac = 0.;
      int tc = 0, ilo = 0, i0 = 0;
      bool needbf = false;
      if (needk) {
        tl.kappa_evals++;
        if (c.st.do_r_lc) {
          if (c.st.opacity_case == 4) {
            sigma = SIGMA_T * c.cs.nne[mgi];
            kappa_ff = calculate_kappa_ff(c, mgi, nu);
            tc = tabcell(c, mgi);
            T_e = (double)c.cs.Te[mgi];
            expfac = exp(-HOVERKB * nu / T_e);
            i0 = bf_stop_index(c, nu);
            ilo = bf_start_index(c, nu, i0);
            tl.bf_evaluated += i0 - ilo;
            needbf = true;
          } else {
            kappa_ff = 1e5 * calculate_kappa_ff(c, mgi, nu);
          }
        }
      }
      unsigned owners = __ballot_sync(full, needbf);
      while (owners != 0u) {
        const int o = __ffs(owners) - 1;
        owners &= owners - 1u;
        const double o_nu = __shfl_sync(full, nu, o), o_Te = __shfl_sync(full, T_e, o), o_exp = __shfl_sync(full, expfac, o);
        const int o_tc = __shfl_sync(full, tc, o), o_tc0 = __shfl_sync(full, tc0, o), o_mgi = __shfl_sync(full, mgi, o);
        const int o_ilo = __shfl_sync(full, ilo, o), o_i0 = __shfl_sync(full, i0, o);
        double sum = 0.;
        for (int base = o_ilo; base < o_i0; base += BF_TILE) {
          const int i = base + lane;
          double t = (i < o_i0) ? vpkt_bf_term(c, o_tc, o_tc0, o_mgi, i, o_nu, o_Te, o_exp) : 0.;
#pragma unroll
          for (int off = 1; off < BF_TILE; off <<= 1) t += __shfl_down_sync(full, t, off);
          sum += t;  // lane 0 holds the tile sums in tree order
        }
        sum = __shfl_sync(full, sum, 0);
        if (lane == o) kappa_bf = sum;
      }
      if (needk) {  // rpkt.cuh kappa_store
        k.nu = nu;
        k.mgi = mgi;
        k.valid = true;
        k.total = sigma + kappa_bf + kappa_ff;
        k.es = sigma;
        k.ff = kappa_ff;
        k.bf = kappa_bf;
        if (!isfinite(k.total)) {
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
    }
    __syncwarp(full);
    // (F) the continuum term, then the cell's lines in the reference's order
    if (ray && !dead) {
      if (at_start) {
        first = k;
        at_start = false;
      }
      dead = vpkt_add_continuum(c, k, s_cont, tau_vpkt);
      if (!dead) dead = vpkt_line_sweep<false>(c, d.nu_cmf, t_current, sdist, tc0, d.next_trans, tau_vpkt, tl, true);
    }
    __syncwarp(full);
    // (G) into the next cell, or out of the model
    if (ray && !dead) {
      t_future += (sdist / CLIGHT_PROP);
      d.prop_time = t_future;
      move_pkt(c, d, sdist);
      bool out = false;
      if (snext == -99) {
        out = true;
      } else {
        d.where = snext;
        stat_inc(tl, COUNTER_CELLCROSSINGS);
      }
      mgi = cell_mgi(c, d.where);
      if (out || mgi == c.st.npts_model) {  // left the grid / reached an empty cell: the ray arrives
        vpkt_ray_bin(c, pk, d, tau_vpkt, sv, vel_vec, bin, obs, t_current);
        escaped++;
        ray = false;
      } else if (c.cs.thick[mgi] == 1) {  // a virtual packet dies in a grey cell
        ray = false;
      }
    }
    if (ray && dead) ray = false;
  }
}

// The persistent kernels' one call site when the rays are NOT deferred (A/B builds), reached by all 32 lanes together. The requests the lanes' events have left are
// served by the whole warp one after the other when there are few of them (the other lanes would only wait), and by their own
// lanes side by side when there are many (32 rays in flight per warp, nobody waits and no geometry is computed 32 times).
#ifndef ARTIS_VPKT_COOP_MAXREQ
#define ARTIS_VPKT_COOP_MAXREQ 6
#endif
__device__ __forceinline__ void coop_vpkt_flush(const Ctx &c, const bool have, const Pkt &p, KappaCache &k, Tally &tl) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int myreq = have ? tl.vpkt_req : 0;
  unsigned req = __ballot_sync(full, myreq != 0);
  if (req == 0u) return;
  if (__popc(req) > ARTIS_VPKT_COOP_MAXREQ) {
    if (myreq != 0) vpkt_flush_request(c, p, k, tl);
    __syncwarp(full);
    return;
  }
  tl.vpkt_req = 0;
  Pkt mine = p;
  double my_q = 0., my_u = 0.;
  if (myreq == 1) {  // an electron scattering: the packet as it arrived (vpkt_flush_request)
    mine.dir[0] = tl.vpkt_dir[0];
    mine.dir[1] = tl.vpkt_dir[1];
    mine.dir[2] = tl.vpkt_dir[2];
    my_q = tl.vpkt_q;
    my_u = tl.vpkt_u;
  }
  while (req != 0u) {
    const int owner = __ffs(req) - 1;
    req &= req - 1u;
    Pkt q;
    q.where = __shfl_sync(full, mine.where, owner);
    q.type = __shfl_sync(full, mine.type, owner);
    q.last_cross = __shfl_sync(full, mine.last_cross, owner);
    q.interactions = 0;
    q.last_event = 0;
    q.next_trans = __shfl_sync(full, mine.next_trans, owner);
    q.number = 0;
    q.ma_element = q.ma_ion = q.ma_level = q.ma_activatingline = 0;
#pragma unroll
    for (int d = 0; d < 3; d++) {
      q.pos[d] = __shfl_sync(full, mine.pos[d], owner);
      q.dir[d] = __shfl_sync(full, mine.dir[d], owner);
    }
    q.e_cmf = __shfl_sync(full, mine.e_cmf, owner);
    q.e_rf = __shfl_sync(full, mine.e_rf, owner);
    q.nu_cmf = __shfl_sync(full, mine.nu_cmf, owner);
    q.nu_rf = __shfl_sync(full, mine.nu_rf, owner);
    q.prop_time = __shfl_sync(full, mine.prop_time, owner);
    q.cold = nullptr;  // the ray trace reads nothing of the cold record
    const double sq = __shfl_sync(full, my_q, owner), su = __shfl_sync(full, my_u, owner);
    const int realtype = __shfl_sync(full, myreq, owner);
    const int sent = vpkt_call_estimators<true>(c, q, sq, su, q.prop_time, realtype, tl);
    if (sent && lane == owner) k.valid = false;
  }
}
#endif  // __CUDACC__

}  // namespace artis
