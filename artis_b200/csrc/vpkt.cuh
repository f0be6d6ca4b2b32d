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
ARTIS_DEV void vpkt_kappa_cont(const Ctx &c, const Pkt &d, const int mgi, const int tc0, KappaCache &k, Tally &tl) {
  const double nu = d.nu_cmf;
  if (kappa_cache_hit(k, mgi, nu)) return;
  tl.kappa_evals++;
  double sigma = 0., kappa_ff = 0., kappa_bf = 0., kappa_ffheating = 0.;
  if (c.st.do_r_lc) {
    if (c.st.opacity_case == 4) {
      sigma = SIGMA_T * c.cs.nne[mgi];
      kappa_ff = calculate_kappa_ff(c, mgi, nu);
      kappa_ffheating = kappa_ff;
      const int tc = tabcell(c, mgi);
      const double T_e = (double)c.cs.Te[mgi];
      const double expfac = exp(-HOVERKB * nu / T_e);
      const int i0 = bf_stop_index(c, nu);
      const int ilo = bf_start_index(c, nu, i0);
      tl.bf_evaluated += i0 - ilo;
      for (int i = ilo; i < i0; i++) {
        if (c.tb.bf_cell[2 * ((size_t)tc * c.st.nbfcontinua + i)] < 0.) continue;  // absent from the current cell
        const int gl = c.cont_level_g[i];
        const double nnlevel = tab_levelpop(c, tc0, gl);
        const ContRec r = c.cont_rec[i];
        if (!(nnlevel > 0.) || !(nu <= r.nu_edge * c.st.last_phixs_nuovernuedge)) continue;
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
        kappa_bf += nnlevel * sigma_bf * r.probability * corrfactor;
      }
    } else {
      kappa_ff = 1e5 * calculate_kappa_ff(c, mgi, nu);
    }
  }
  for (int g = 0; g < c.st.nbfcontinua_ground; g++) k.gamma_contr[g] = 0.;  // no estimator reads it for a virtual packet
  kappa_store(c, k, mgi, nu, sigma, kappa_ff, kappa_bf, kappa_ffheating, tl);
}

// vpkt.cc:76-364 rlc_emiss_vpkt. `pk` is the real packet (read only), `k` the opacity scratch of this thread (the
// reference's globals::kappa_rpkt_cont[tid], shared between the real packet and its virtual packets). Returns the
// change of the virtual-packet counters: bit 0 = a virtual packet was created, bit 1 = it reached the edge of the grid.
ARTIS_DEV int rlc_emiss_vpkt(const Ctx &c, const Pkt &pk, const double stokes_q, const double stokes_u, const double t_current,
                             const int bin, double *obs, const int realtype, KappaCache &k, Tally &tl) {
  const int nspectra = c.st.vpkt_nspectra;
  double tau_vpkt[VPKT_MAX_SPECTRA];
  for (int ind = 0; ind < VPKT_MAX_SPECTRA; ind++) tau_vpkt[ind] = 0;
  Pkt dummy = pk;
  bool end_packet = false;
  double t_future = t_current;
  dummy.dir[0] = obs[0];
  dummy.dir[1] = obs[1];
  dummy.dir[2] = obs[2];
  double vel_vec[3], old_dir_cmf[3], obs_cmf[3], vel_rev[3], ref1[3], ref2[3];
  get_velocity(pk.pos, vel_vec, t_current);
  dummy.nu_rf = dummy.nu_cmf / doppler_nucmf_on_nurf(c, dummy.dir, vel_vec);
  dummy.e_rf = dummy.e_cmf * dummy.nu_rf / dummy.nu_cmf;
  double Qi = stokes_q, Ui = stokes_u;  // Stokes Q, U of the real packet (used by the scattering phase function only)
  double I = 0., Q = 0., U = 0., pn = 0.;
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
    pn = 3. / (16. * PI) * (1 + mu * mu + (mu * mu - 1) * Qold);  // pow(mu, 2.) in the reference
    const double Inew = 0.75 * ((mu * mu + 1.0) + Qold * (mu * mu - 1.0));
    double Qnew = 0.75 * ((mu * mu - 1.0) + Qold * (mu * mu + 1.0));
    double Unew = 1.5 * mu * Uold;
    Qnew = Qnew / Inew;
    Unew = Unew / Inew;
    I = Inew / Inew;
    meridian(obs_cmf, ref1, ref2);
    const double i2 = PI + rot_angle(obs_cmf, old_dir_cmf, ref1, ref2);
    const double cos2i2 = cos(2 * i2), sin2i2 = sin(2 * i2);
    Q = Qnew * cos2i2 + Unew * sin2i2;
    U = -Qnew * sin2i2 + Unew * cos2i2;
    vel_rev[0] = -vel_vec[0];
    vel_rev[1] = -vel_vec[1];
    vel_rev[2] = -vel_vec[2];
    frame_transform(obs_cmf, &Q, &U, vel_rev, obs);  // writes the rest-frame direction back into obs (vpkt.cc:178)
  }
  if (realtype == 2 || realtype == 3) {  // isotropic emission
    I = 1;
    Q = 0;
    U = 0;
    pn = 1 / (4 * PI);
  }
  int mgi = cell_mgi(c, dummy.where);
  const int tc0 = tabcell(c, mgi);  // level populations of the whole ray come from the start cell (see vpkt_kappa_cont)
  while (!end_packet) {
    double ldist = 0;
    int snext = 0;
    const double sdist = boundary_cross(c, dummy, &snext);
    const double s_cont = sdist * t_current * t_current * t_current / (t_future * t_future * t_future);
    vpkt_kappa_cont(c, dummy, mgi, tc0, k, tl);
    const double kap_cont = k.total;
    const double kap_cont_nobf = kap_cont - k.bf, kap_cont_noff = kap_cont - k.ff, kap_cont_noes = kap_cont - k.es;
    for (int ind = 0; ind < nspectra; ind++) {
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
    if (vpkt_check_tau(c, tau_vpkt) == 0) return 1;
    while (ldist < sdist) {
      const int lineindex = closest_transition(c, dummy.nu_cmf, dummy.next_trans);
      if (lineindex >= 0) {
        tl.lines_scanned++;
        const double nutrans = c.st.line_nu[lineindex];
        dummy.next_trans = lineindex + 1;
        if (dummy.nu_cmf < nutrans) {
          ldist = 0;
        } else {
          ldist = CLIGHT * t_current * (dummy.nu_cmf / nutrans - 1);
        }
        if (ldist > sdist) {  // the line lies in the next cell: start there with this line
          dummy.next_trans -= 1;
          break;
        }
        tl.lines_tau++;
        const double t_line = t_current + ldist / CLIGHT;
        const int glu = c.line_upper_g[lineindex], gll = c.line_lower_g[lineindex];
        const double A_ul = (double)c.st.line_einstein_A[lineindex];
        const double B_ul = CLIGHTSQUAREDOVERTWOH / pow(nutrans, 3) * A_ul;
        const double B_lu = (double)c.st.lev_statweight[glu] / (double)c.st.lev_statweight[gll] * B_ul;
        const double n_u = tab_levelpop(c, tc0, glu);
        const double n_l = tab_levelpop(c, tc0, gll);
        const int anumber = c.st.el_anumber[c.st.line_element[lineindex]];
        for (int ind = 0; ind < nspectra; ind++) {
          const double ex = c.st.vpkt_exclude[ind];
          if (ex != -1 && ((double)anumber != ex)) tau_vpkt[ind] += (B_lu * n_l - B_ul * n_u) * HCLIGHTOVERFOURPI * t_line;
        }
        if (vpkt_check_tau(c, tau_vpkt) == 0) return 1;
      } else {
        // The reference sets next_trans = nlines + 1 and keeps looping with ldist unchanged (vpkt.cc:297-300), which
        // does not terminate; it is only reached by a packet redder than the reddest line of the list. No line is left
        // in this or any later cell, so the sweep of the cell ends here.
        dummy.next_trans = c.st.nlines + 1;
        break;
      }
    }
    t_future += (sdist / CLIGHT_PROP);
    dummy.prop_time = t_future;
    move_pkt(c, dummy, sdist);
    if (snext == -99) {  // change_cell (boundary.cc:332-357). The reference also bumps globals::nesc for a virtual
      end_packet = true; // packet leaving the grid; here COUNTER_NESC counts real packets only
    } else {
      dummy.where = snext;
      stat_inc(tl, COUNTER_CELLCROSSINGS);
    }
    mgi = cell_mgi(c, dummy.where);
    if (mgi == c.st.npts_model) break;        // reached an empty cell
    if (c.cs.thick[mgi] == 1) return 1;       // a virtual packet dies in a grey cell
  }
  // final Stokes vector, binned on the fly
  const double t_arrive = t_current - (dot3(pk.pos, dummy.dir) / CLIGHT_PROP);
  for (int ind = 0; ind < nspectra; ind++) {
    const double prob = pn * exp(-tau_vpkt[ind]);
    const double stokes[3] = {I * prob, Q * prob, U * prob};
    add_to_vspecpol(c, dummy.nu_rf, dummy.e_rf, stokes, bin, ind, t_arrive);
  }
  if (c.st.vpkt_vgrid_flag == 1) {
    const double prob = pn * exp(-tau_vpkt[0]);
    const double stokes[3] = {I * prob, Q * prob, U * prob};
    for (int bin_range = 0; bin_range < c.st.vpkt_nrange_grid; bin_range++) {
      if (dummy.nu_rf > c.st.vpkt_nu_grid_min[bin_range] && dummy.nu_rf < c.st.vpkt_nu_grid_max[bin_range]) {
        if (t_arrive > c.st.vpkt_tmin_grid && t_arrive < c.st.vpkt_tmax_grid)
          add_to_vpkt_grid(c, dummy.nu_rf, dummy.e_rf, stokes, vel_vec, bin_range, bin, obs);
      }
    }
  }
  return 3;
}

// vpkt.cc:837-896 vpkt_call_estimators for one interaction of the real packet `pk` (taken by value: the caller's copy is
// not touched -- the only field the reference modifies, next_trans, is rewritten with the value it already has).
// Owns its opacity scratch and work tally, like a host thread of the reference owns kappa_rpkt_cont[tid]; returns 1 when
// at least one virtual packet was sent (the caller then drops its own cached opacity, as cellhistory_reset +
// calculate_kappa_rpkt_cont at vpkt.cc:887-893 re-evaluate it at the packet's exact frequency).
ARTIS_DEV_NOINLINE int vpkt_call_estimators(const Ctx &c, const Pkt pk, const double stokes_q, const double stokes_u,
                                            const double t_current, const int realtype) {
  const int mgi = cell_mgi(c, pk.where);
  if (c.cs.thick[mgi] != 0) return 0;
  Tally tl = {};
  KappaCache k;
  k.valid = false;
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
          const int r = rlc_emiss_vpkt(c, pk, stokes_q, stokes_u, t_current, bin, obs, realtype, k, tl);
          created += (r & 1);
          escaped += (r >> 1) & 1;
          vflag = 1;
          k.valid = false;  // cellhistory_reset(mgi, false) at vpkt.cc:887
        }
      }
    }
  }
  if (created) est_add(&c.es.ts_scalars[ARTIS_TS_NVPKT], (double)created);
  if (escaped) est_add(&c.es.ts_scalars[ARTIS_TS_NVPKT_ESC1 + realtype - 1], (double)escaped);
  tally_flush(c, tl);
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
  if (vpkt_call_estimators(c, q, stokes_q, stokes_u, q.prop_time, realtype)) k.valid = false;
}

}  // namespace artis
