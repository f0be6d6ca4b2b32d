// rpkt_events.cuh -- estimator update, emission, electron scattering with Stokes parameters, event handlers and
// do_rpkt_step (rpkt.cc:330-509, 557-819, 975-1025; polarization.cc:6-140; vpkt.cc:898-1069 frame transforms).
#pragma once
#include "rpkt.cuh"

namespace artis {

// ---- rpkt.cc:557-621 update_estimators + radfield.cc:831-860 + grey_emissivities.cc:79-106 ---------------------
// The sweep of radfield.cc:764-829 for one r-packet step: continua first, first + stride, ... of those whose table
// covers nu (a single thread passes first = 0, stride = 1; a warp passes lane, 32). gamma_contr[i] is zero from the first
// continuum with knu < nu_edge on, and for a continuum whose table does not reach knu.
ARTIS_DEV void bfest_sweep(const Ctx &c, const int tc, const int mgi, const double knu, const double nu, const double factor,
                           const int first, const int stride, Tally &tl) {
  const int i0k = bf_stop_index(c, knu);
  int i0 = bf_stop_index(c, nu);
  if (i0k < i0) i0 = i0k;
  const int ilo = bf_start_index(c, nu, i0);
  const double expfac = exp(-HOVERKB * knu / (double)c.cs.Te[mgi]);
  double *row = c.es.bfrate_raw + (size_t)tc * c.st.nbfcontinua;
  for (int i = ilo + first; i < i0; i += stride) {
    double gcontr;
    kappa_bf_term(c, tc, i, knu, expfac, &gcontr);
    if (gcontr != 0.) {
      tl.estim_atomics += est_add(&row[i], gcontr * factor);
    }
  }
}
ARTIS_DEV void update_estimators(const Ctx &c, const Pkt &p, const double distance, KappaCache &k, Tally &tl) {
  const int mgi = cell_mgi(c, p.where);
  if (mgi == c.st.npts_model) return;
  const double distance_e_cmf = distance * p.e_cmf;
  const double nu = p.nu_cmf;
  tl.estim_atomics += est_add(&c.es.J[mgi], distance_e_cmf);
  tl.estim_atomics += est_add(&c.es.nuJ[mgi], distance_e_cmf * nu);
  // The reference adds globals::kappa_rpkt_cont[tid].ffheating and phixslist[tid].groundcont_gamma_contr, i.e.
  // whatever the last opacity evaluation of that host thread left behind -- in grey cells that is the opacity of
  // some other cell/packet. Here the scratch only counts when it was evaluated for this cell.
  const bool kv = k.valid && (k.mgi == mgi);
  // grey cells (no valid opacity scratch) would add +0.0: skipped
  if (kv) tl.estim_atomics += est_add(&c.es.ffheatingestimator[mgi], distance_e_cmf * k.ffheating);
  // radfield.cc:764-829 update_bfestimators (DETAILED_BF_ESTIMATORS_ON): every continuum whose cross-section table
  // covers nu gets gamma_contr[i] * distance_e_cmf / nu * doppler, where gamma_contr is the per-continuum scratch of the
  // last opacity evaluation (at k.nu). The reference keeps that scratch as an nbfcontinua-long array per host
  // thread; here the term is re-evaluated at k.nu from the per-cell table -- same function, same inputs, same bits.
  if (ARTIS_OPT_DETAILED_BF(c) && kv && distance_e_cmf != 0) {
    const int tc = tabcell(c, mgi);
    const double distance_e_cmf_over_nu = distance_e_cmf / nu * doppler_packet(c, p);
    if (k.defer_bfest) {  // the warp runs the sweep right after the step (same terms, same factor)
      k.bfest_tc = tc;
      k.bfest_nu = nu;
      k.bfest_factor = distance_e_cmf_over_nu;
    } else {
      bfest_sweep(c, tc, mgi, k.nu, nu, distance_e_cmf_over_nu, 0, 1, tl);
    }
  }
  // radfield.cc:843-874: frequency-binned J and nuJ of the multi-bin radiation field model
  if (ARTIS_OPT_MULTIBIN(c)) {
    const int binindex = radfield_select_bin(c, nu);
    if (binindex >= 0) {
      const size_t mgibinindex = (size_t)tabcell(c, mgi) * c.st.opt_radfieldbincount + binindex;
      tl.estim_atomics += est_add(&c.es.radfieldbin_J_raw[mgibinindex], distance_e_cmf);
      tl.estim_atomics += est_add(&c.es.radfieldbin_nuJ_raw[mgibinindex], distance_e_cmf * nu);
      int_add(&c.es.radfieldbin_contribcount[mgibinindex], 1);
      tl.estim_atomics += 1;
    }
  }
  // without a valid opacity scratch for this cell every contribution below is +0.0: nothing to add (grey cells)
  if (kv && (!ARTIS_OPT_NOLUT_PHOTOION(c) || !ARTIS_OPT_NOLUT_BFHEATING(c))) {
    const int nelements = c.st.nelements, max_nions = c.st.max_nions;
    const double distance_e_cmf_over_nu = distance_e_cmf / nu;
    const int ng = c.st.nbfcontinua_ground;
    for (int i = 0; i < ng; i++) {
      const double nu_edge = c.st.gcont_nu_edge[i];
      if (nu > nu_edge) {
        const int element = c.st.gcont_element[i];
        if (elem_abundance(c, mgi, element) > 0) {
          const int ion = c.st.gcont_ion[i];
          const int ionestimindex = mgi * nelements * max_nions + element * max_nions + ion;
          const double contr = k.gamma_contr[i];
          if (contr == 0.) continue;  // adds +0.0 to both estimators
          if (!ARTIS_OPT_NOLUT_PHOTOION(c)) {
            tl.estim_atomics += est_add(&c.es.gammaestimator[ionestimindex], contr * distance_e_cmf_over_nu);
          }
          if (!ARTIS_OPT_NOLUT_BFHEATING(c)) {
            tl.estim_atomics += est_add(&c.es.bfheatingestimator[ionestimindex], contr * distance_e_cmf * (1. - nu_edge / nu));
          }
        }
      } else {
        break;
      }
    }
  }
}
ARTIS_DEV void rlc_emiss_rpkt(const Ctx &c, const Pkt &p, const double dist, Tally &tl) {
  const int mgi = cell_mgi(c, p.where);
  if (dist > 0.0) {
    double vel_vec[3];
    get_velocity(p.pos, vel_vec, p.prop_time);
    double cont = (double)(c.cs.kappagrey[mgi] * c.cs.rho[mgi]);  // float product (grey_emissivities.cc:97)
    cont = cont * p.e_rf * dist * (1. - (2. * dot3(vel_vec, p.dir) / CLIGHT));
    tl.estim_atomics += est_add(&c.es.rpkt_emiss[mgi], 1.e-20 * cont);
  }
}

// ---- rpkt.cc:975-1025 emitt_rpkt ---------------------------------------------------------------------------------------
ARTIS_DEV void set_pol_dir(Pkt &p) {
  double dummy_dir[3] = {0., 0., 1.};
  double pol_dir[3];
  cross_prod(p.dir, dummy_dir, pol_dir);
  if (dot3(pol_dir, pol_dir) < 1.e-8) {
    dummy_dir[0] = dummy_dir[2] = 0.0;
    dummy_dir[1] = 1.0;
    cross_prod(p.dir, dummy_dir, pol_dir);
  }
  vec_norm(pol_dir, pol_dir);
  p.cold->pol_dir[0] = pol_dir[0];
  p.cold->pol_dir[1] = pol_dir[1];
  p.cold->pol_dir[2] = pol_dir[2];
}
ARTIS_DEV void emitt_rpkt_body(const Ctx &c, Pkt &p, Rng &rng) {
  p.type = ARTIS_TYPE_RPKT;
  p.last_cross = ARTIS_CROSS_NONE;
  double dir_cmf[3];
  get_rand_isotropic_unitvec(rng, dir_cmf);
  double vel_vec[3];
  get_velocity(p.pos, vel_vec, -1. * p.prop_time);
  angle_ab(dir_cmf, vel_vec, p.dir);
  const double dopplerfactor = doppler_packet(c, p);
  p.nu_rf = p.nu_cmf / dopplerfactor;
  p.e_rf = p.e_cmf / dopplerfactor;
  p.cold->stokes[0] = 1.;
  p.cold->stokes[1] = 0.;
  p.cold->stokes[2] = 0.;
  set_pol_dir(p);
}
ARTIS_DEV_NOINLINE void emitt_rpkt_ool(const Ctx &c, EvState *s) { emitt_rpkt_body(c, s->p, s->rng); }
ARTIS_DEV void emitt_rpkt(const Ctx &c, Pkt &p, Rng &rng) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  emitt_rpkt_ool(c, &s);
  p = s.p;
  rng = s.rng;
#else
  emitt_rpkt_body(c, p, rng);
#endif
}

// ---- vpkt.cc:898-1069 rot_angle / meridian / lorentz / frame_transform ----------------------------------------------
ARTIS_DEV double rot_angle(const double *n1, const double *n2, const double *ref1, const double *ref2) {
  double i = 0;
  double ref1_sc[3];
  const double n1n2 = dot3(n1, n2);
  ref1_sc[0] = n1[0] * n1n2 - n2[0];
  ref1_sc[1] = n1[1] * n1n2 - n2[1];
  ref1_sc[2] = n1[2] * n1n2 - n2[2];
  vec_norm(ref1_sc, ref1_sc);
  double cos1 = dot3(ref1_sc, ref1);
  const double cos2 = dot3(ref1_sc, ref2);
  if (cos1 < -1) cos1 = -1;
  if (cos1 > 1) cos1 = 1;
  const double pi_ = acos(-1.);
  if ((cos1 > 0) && (cos2 > 0)) i = acos(cos1);
  if ((cos1 > 0) && (cos2 < 0)) i = 2 * pi_ - acos(cos1);
  if ((cos1 < 0) && (cos2 < 0)) i = pi_ + acos(fabs(cos1));
  if ((cos1 < 0) && (cos2 > 0)) i = pi_ - acos(fabs(cos1));
  if (cos1 == 0) i = pi_ / 2.;
  if (cos2 == 0) i = 0.0;
  return i;
}
ARTIS_DEV void meridian(const double *n, double *ref1, double *ref2) {
  const double s = sqrt(n[0] * n[0] + n[1] * n[1]);
  ref1[0] = -1. * n[0] * n[2] / s;
  ref1[1] = -1. * n[1] * n[2] / s;
  ref1[2] = (1 - (n[2] * n[2])) / s;
  ref2[0] = n[2] * ref1[1] - n[1] * ref1[2];
  ref2[1] = n[0] * ref1[2] - n[2] * ref1[0];
  ref2[2] = n[1] * ref1[0] - n[0] * ref1[1];
}
ARTIS_DEV void lorentz(const double *e_rf, const double *n_rf, const double *v, double *e_cmf) {
  double beta[3], e_par[3], e_perp[3], b_rf[3], v_cr_b[3];
  beta[0] = v[0] / CLIGHT;
  beta[1] = v[1] / CLIGHT;
  beta[2] = v[2] / CLIGHT;
  const double vsqr = dot3(beta, beta);
  const double gamma_rel = 1. / (sqrt(1 - vsqr));
  const double edotb = (e_rf[0] * beta[0] + e_rf[1] * beta[1] + e_rf[2] * beta[2]);
  for (int d = 0; d < 3; d++) {
    e_par[d] = edotb * beta[d] / (vsqr);
    e_perp[d] = e_rf[d] - e_par[d];
  }
  b_rf[0] = n_rf[1] * e_rf[2] - n_rf[2] * e_rf[1];
  b_rf[1] = n_rf[2] * e_rf[0] - n_rf[0] * e_rf[2];
  b_rf[2] = n_rf[0] * e_rf[1] - n_rf[1] * e_rf[0];
  v_cr_b[0] = beta[1] * b_rf[2] - beta[2] * b_rf[1];
  v_cr_b[1] = beta[2] * b_rf[0] - beta[0] * b_rf[2];
  v_cr_b[2] = beta[0] * b_rf[1] - beta[1] * b_rf[0];
  for (int d = 0; d < 3; d++) e_cmf[d] = e_par[d] + gamma_rel * (e_perp[d] + v_cr_b[d]);
  vec_norm(e_cmf, e_cmf);  // the magnetic-field part of the reference routine does not feed any output
}
ARTIS_DEV void frame_transform(const double *n_rf, double *Q, double *U, const double *v, double *n_cmf) {
  double e_rf[3], e_cmf[3], ref1[3], ref2[3];
  double theta_rot = 0.;
  const double pi_ = acos(-1.);
  meridian(n_rf, ref1, ref2);
  const double Q0 = *Q, U0 = *U;
  const double pp = sqrt(Q0 * Q0 + U0 * U0);
  double rot = 0;
  if (pp > 0) {
    const double cos2rot = Q0 / pp, sin2rot = U0 / pp;
    if ((cos2rot > 0) && (sin2rot > 0)) rot = acos(Q0 / pp) / 2.;
    if ((cos2rot < 0) && (sin2rot > 0)) rot = (pi_ - acos(fabs(Q0 / pp))) / 2.;
    if ((cos2rot < 0) && (sin2rot < 0)) rot = (pi_ + acos(fabs(Q0 / pp))) / 2.;
    if ((cos2rot > 0) && (sin2rot < 0)) rot = (2. * pi_ - acos(fabs(Q0 / pp))) / 2.;
    if (cos2rot == 0) {
      rot = 0.25 * pi_;
      if (U0 < 0) rot = 0.75 * pi_;
    }
    if (sin2rot == 0) {
      rot = 0.0;
      if (Q0 < 0) rot = 0.5 * pi_;
    }
  }
  const double cr = cos(rot), sr = sin(rot);
  for (int d = 0; d < 3; d++) e_rf[d] = cr * ref1[d] - sr * ref2[d];
  angle_ab(n_rf, v, n_cmf);
  lorentz(e_rf, n_rf, v, e_cmf);
  meridian(n_cmf, ref1, ref2);
  const double e1 = dot3(e_cmf, ref1);
  const double e2 = dot3(e_cmf, ref2);
  if ((e1 > 0) && (e2 < 0)) theta_rot = acos(e1);
  if ((e1 < 0) && (e2 < 0)) theta_rot = pi_ - acos(fabs(e1));
  if ((e1 < 0) && (e2 > 0)) theta_rot = pi_ + acos(fabs(e1));
  if ((e1 > 0) && (e2 > 0)) theta_rot = 2 * pi_ - acos(e1);
  if (e1 == 0) theta_rot = pi_ / 2.;
  if (e2 == 0) theta_rot = 0.0;
  if (e1 > 1) theta_rot = 0.0;
  if (e1 < -1) theta_rot = pi_;
  *Q = cos(2 * theta_rot) * pp;
  *U = sin(2 * theta_rot) * pp;
}

// ---- polarization.cc:6-140 escat_rpkt ----------------------------------------------------------------------------------
ARTIS_DEV void escat_rpkt_body(const Ctx &c, Pkt &p, Rng &rng) {
  p.type = ARTIS_TYPE_RPKT;
  p.last_cross = ARTIS_CROSS_NONE;
  double vel_vec[3];
  get_velocity(p.pos, vel_vec, p.prop_time);
  double Qi = ldcold(&p.cold->stokes[1]), Ui = ldcold(&p.cold->stokes[2]);
  double old_dir_cmf[3];
  frame_transform(p.dir, &Qi, &Ui, vel_vec, old_dir_cmf);
  double M = 0., mu = 0., phisc = 0.;
  if (c.st.opt_dipole) {
    double pr = 0., x = 0.;
    do {
      const double zrand = rng_uniform(rng);
      const double zrand2 = rng_uniform(rng);
      const double zrand3 = rng_uniform(rng);
      M = 2 * zrand - 1;
      mu = M * M;
      phisc = 2 * PI * zrand2;
      pr = (mu + 1) + (mu - 1) * (cos(2 * phisc) * Qi + sin(2 * phisc) * Ui);
      x = 2 * zrand3;
    } while (x > pr);
  } else {
    const double zrand = rng_uniform(rng);
    const double zrand2 = rng_uniform(rng);
    M = 2. * zrand - 1;
    mu = M * M;
    phisc = 2 * PI * zrand2;
  }
  const double tsc = acos(M);
  double new_dir_cmf[3];
  if (fabs(old_dir_cmf[2]) < 0.99999) {
    const double s = sqrt(1. - old_dir_cmf[2] * old_dir_cmf[2]);
    new_dir_cmf[0] = sin(tsc) / s * (old_dir_cmf[1] * sin(phisc) - old_dir_cmf[0] * old_dir_cmf[2] * cos(phisc)) +
                     old_dir_cmf[0] * cos(tsc);
    new_dir_cmf[1] = sin(tsc) / s * (-old_dir_cmf[0] * sin(phisc) - old_dir_cmf[1] * old_dir_cmf[2] * cos(phisc)) +
                     old_dir_cmf[1] * cos(tsc);
    new_dir_cmf[2] = sin(tsc) * cos(phisc) * s + old_dir_cmf[2] * cos(tsc);
  } else {
    new_dir_cmf[0] = sin(tsc) * cos(phisc);
    new_dir_cmf[1] = sin(tsc) * sin(phisc);
    new_dir_cmf[2] = (old_dir_cmf[2] > 0) ? cos(tsc) : -cos(tsc);
  }
  double ref1[3], ref2[3];
  meridian(old_dir_cmf, ref1, ref2);
  const double i1 = rot_angle(old_dir_cmf, new_dir_cmf, ref1, ref2);
  const double cos2i1 = cos(2 * i1), sin2i1 = sin(2 * i1);
  const double Qold = Qi * cos2i1 - Ui * sin2i1;
  const double Uold = Qi * sin2i1 + Ui * cos2i1;
  mu = dot3(old_dir_cmf, new_dir_cmf);
  const double Inew = 0.75 * ((mu * mu + 1.0) + Qold * (mu * mu - 1.0));
  double Qnew = 0.75 * ((mu * mu - 1.0) + Qold * (mu * mu + 1.0));
  double Unew = 1.5 * mu * Uold;
  Qnew = Qnew / Inew;
  Unew = Unew / Inew;
  meridian(new_dir_cmf, ref1, ref2);
  const double i2 = PI + rot_angle(new_dir_cmf, old_dir_cmf, ref1, ref2);
  const double cos2i2 = cos(2 * i2), sin2i2 = sin(2 * i2);
  double Q = Qnew * cos2i2 + Unew * sin2i2;
  double U = -Qnew * sin2i2 + Unew * cos2i2;
  double vel_rev[3] = {-vel_vec[0], -vel_vec[1], -vel_vec[2]};
  double dummy_dir[3];
  frame_transform(new_dir_cmf, &Q, &U, vel_rev, dummy_dir);
  p.cold->stokes[0] = 1.0;
  p.cold->stokes[1] = Q;
  p.cold->stokes[2] = U;
  p.dir[0] = dummy_dir[0];
  p.dir[1] = dummy_dir[1];
  p.dir[2] = dummy_dir[2];
  const double dopplerfactor = doppler_packet(c, p);
  p.nu_rf = p.nu_cmf / dopplerfactor;
  p.e_rf = p.e_cmf / dopplerfactor;
}
ARTIS_DEV_NOINLINE void escat_rpkt_ool(const Ctx &c, EvState *s) { escat_rpkt_body(c, s->p, s->rng); }
ARTIS_DEV void escat_rpkt(const Ctx &c, Pkt &p, Rng &rng) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  escat_rpkt_ool(c, &s);
  p = s.p;
  rng = s.rng;
#else
  escat_rpkt_body(c, p, rng);
#endif
}

// ---- rpkt.cc:330-447 rpkt_event_continuum --------------------------------------------------------------------------
}  // namespace artis
#include "vpkt.cuh"
namespace artis {

ARTIS_DEV void rpkt_event_continuum_body(const Ctx &c, Pkt &p, const KappaCache &k, const int mgi, Rng &rng, Tally &tl) {
  const double nu = p.nu_cmf;
  const double dopplerfactor = doppler_packet(c, p);
  const double kappa_cont = k.total * dopplerfactor;
  const double sigma = k.es * dopplerfactor;
  const double kappa_ff = k.ff * dopplerfactor;
  const double kappa_bf = k.bf * dopplerfactor;
  const double zrand = rng_uniform(rng);
  if (zrand * kappa_cont < sigma) {
    p.interactions += 1;
    tl.interactions++;
    p.cold->nscatterings = ldcold(&p.cold->nscatterings) + 1;
    p.last_event = 12;
    stat_inc(tl, COUNTER_ESCOUNTER);
    if (ARTIS_OPT_VPKT(c)) {  // rpkt.cc:359-363: virtual packets see the packet as it arrives at the scattering
      p.last_cross = ARTIS_CROSS_NONE;
      tl.vpkt_req = 1;
      tl.vpkt_dir[0] = p.dir[0];
      tl.vpkt_dir[1] = p.dir[1];
      tl.vpkt_dir[2] = p.dir[2];
      tl.vpkt_q = ldcold(&p.cold->stokes[1]);
      tl.vpkt_u = ldcold(&p.cold->stokes[2]);
    }
    escat_rpkt(c, p, rng);
    p.cold->em_pos[0] = p.pos[0];
    p.cold->em_pos[1] = p.pos[1];
    p.cold->em_pos[2] = p.pos[2];
    p.cold->em_time = (int)p.prop_time;
  } else if (zrand * kappa_cont < sigma + kappa_ff) {
    stat_inc(tl, COUNTER_K_STAT_FROM_FF);
    p.interactions += 1;
    tl.interactions++;
    p.last_event = 5;
    p.type = ARTIS_TYPE_KPKT;
    p.cold->absorptiontype = -1;
  } else if (zrand * kappa_cont < sigma + kappa_ff + kappa_bf) {
    p.cold->absorptiontype = -2;
    const double kappa_bf_inrest = k.bf;
    const double zrand2 = rng_uniform(rng);
    const double kappa_bf_rand = zrand2 * kappa_bf_inrest;
    // std::lower_bound over kappa_bf_sum[0 .. nbf-1): the prefix sum is re-accumulated in index order at
    // the frequency of the cached opacity, so the values are those the reference stored in its per-thread scratch
    const int tc = tabcell(c, mgi);
    const double expfac = exp(-HOVERKB * k.nu / (double)c.cs.Te[mgi]);
    const int nbf = c.st.nbfcontinua;
    int allcontindex = nbf - 1;
    double sum = 0.;
    // entries at and beyond the stop index all equal the last sum (rpkt.cc:1117-1125), so only i < i0 can be selected
    // before lower_bound runs off the end -- except that an all-zero list with a zero draw selects entry 0
    const int i0 = bf_stop_index(c, k.nu);
    const int lim = i0 < nbf - 1 ? i0 : nbf - 1;
    if (nbf > 1 && !(sum < kappa_bf_rand)) {
      allcontindex = 0;  // a zero draw selects the first entry whatever it holds
    } else {
      // entries below the table-reach window are exact zeros (bf_start_index): the running sum starts there
      for (int i = bf_start_index(c, k.nu, i0); i < lim; i++) {
        double gcontr;
        sum += kappa_bf_term(c, tc, i, k.nu, expfac, &gcontr);
        if (!(sum < kappa_bf_rand)) {
          allcontindex = i;
          break;
        }
      }
    }
    const double nu_edge = c.st.cont_nu_edge[allcontindex];
    const int element = c.st.cont_element[allcontindex];
    const int ion = c.st.cont_ion[allcontindex];
    const double zrand3 = rng_uniform(rng);
    if (zrand3 < nu_edge / nu) {
      stat_inc(tl, COUNTER_MA_STAT_ACTIVATION_BF);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 3;
      p.type = ARTIS_TYPE_MA;
      p.ma_element = element;
      p.ma_ion = ion + 1;
      p.ma_level = c.st.cont_upperlevel[allcontindex];  // == get_phixsupperlevel(element, ion, level, target)
      p.ma_activatingline = -99;
    } else {
      stat_inc(tl, COUNTER_K_STAT_FROM_BF);
      p.interactions += 1;
      tl.interactions++;
      p.last_event = 4;
      p.type = ARTIS_TYPE_KPKT;
    }
  } else {
    fatal(c, 103);
  }
}
ARTIS_DEV_NOINLINE void rpkt_event_continuum_ool(const Ctx &c, EvState *s, const KappaCache &k, const int mgi) { rpkt_event_continuum_body(c, s->p, k, mgi, s->rng, s->tl); }
ARTIS_DEV void rpkt_event_continuum(const Ctx &c, Pkt &p, const KappaCache &k, const int mgi, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  rpkt_event_continuum_ool(c, &s, k, mgi);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  rpkt_event_continuum_body(c, p, k, mgi, rng, tl);
#endif
}

// rpkt.cc:449-489
ARTIS_DEV void rpkt_event_boundbound(const Ctx &c, Pkt &p, Tally &tl) {
  stat_inc(tl, COUNTER_MA_STAT_ACTIVATION_BB);
  p.interactions += 1;
  tl.interactions++;
  p.last_event = 1;
  p.cold->absorptiontype = p.ma_activatingline;
  p.cold->absorptionfreq = p.nu_rf;
  p.cold->absorptiondir[0] = p.dir[0];
  p.cold->absorptiondir[1] = p.dir[1];
  p.cold->absorptiondir[2] = p.dir[2];
  p.type = ARTIS_TYPE_MA;
  if (c.st.opt_record_linestat) int_add(&c.es.acounter[p.next_trans - 1], 1);  // all threads
}
// rpkt.cc:491-509
ARTIS_DEV void rpkt_event_thickcell(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
  p.interactions += 1;
  tl.interactions++;
  p.cold->nscatterings = ldcold(&p.cold->nscatterings) + 1;
  p.last_event = 12;
  stat_inc(tl, COUNTER_ESCOUNTER);
  emitt_rpkt(c, p, rng);
  p.cold->em_pos[0] = p.pos[0];
  p.cold->em_pos[1] = p.pos[1];
  p.cold->em_pos[2] = p.pos[2];
  p.cold->em_time = (int)p.prop_time;
}

// ---- rpkt.cc:623-813 do_rpkt_step ----------------------------------------------------------------------------------------
// returns true while the packet is still an r-packet inside its time step (the reference additionally stops at a
// model-cell change because of its cellhistory cache; the wavefront scheduler decides that)
// The step is written in two halves so that a scheduler can evaluate the continuum opacity between them with a whole
// warp (transport.cu coop_kappa_warm): rpkt_step_begin runs up to the point where get_event needs kappa and returns
// false when the step is already over (cell change at zero distance, fatal geometry); rpkt_step_finish is the rest.
struct RpktStepState {
  int mgi, snext;
  int mgi_next;  // model cell of snext, requested as soon as snext is known so that the gather flies during the move
  double tau_next, sdist;
};
ARTIS_DEV int rpkt_step_mgi_next(const Ctx &c, const Pkt &p, const RpktStepState &st) {
  return (st.snext >= 0 && st.snext != p.where) ? cell_mgi(c, st.snext) : st.mgi;
}
template <int MODE = 0>
ARTIS_DEV bool rpkt_step_begin(const Ctx &c, Pkt &p, Rng &rng, Tally &tl, RpktStepState &st) {
  tl.rpkt_steps++;
  st.mgi = cell_mgi(c, p.where);
  const double zrand = rng_uniform_pos(rng);
  st.tau_next = -1. * log(zrand);
  int snext = 0;
  double sdist = boundary_cross(c, p, &snext);
  if (sdist == 0) {
    change_cell(c, p, snext, tl);
    return false;
  }
  const double maxsdist = max_sdist(c, p, sdist);
  if (sdist > maxsdist) {
    fatal(c, 104);
    p.type = ARTIS_TYPE_ESCAPE;  // park the packet; the run is reported as failed
    return false;
  }
  if (((snext != -99) && (snext < 0)) || (snext >= c.st.ngrid)) {
    fatal(c, 105);
    p.type = ARTIS_TYPE_ESCAPE;
    return false;
  }
  if (sdist > c.cs.max_path_step) {
    sdist = c.cs.max_path_step;
    snext = p.where;
  }
  st.sdist = sdist;
  st.snext = snext;
  st.mgi_next = rpkt_step_mgi_next(c, p, st);
  return true;
}
template <int MODE = 0>
ARTIS_DEV bool rpkt_step_finish(const Ctx &c, Pkt &p, const double t2, KappaCache &k, Rng &rng, Tally &tl,
                                const RpktStepState &st) {
  int mgi = st.mgi;
  const double tau_next = st.tau_next;
  const double sdist = st.sdist;
  const int snext = st.snext;
  const double tdist = (t2 - p.prop_time) * CLIGHT_PROP;
  double edist;
  int rpkt_eventtype = RPKT_EVENTTYPE_NONE;
  bool find_nextline = false;
  const bool thick = (MODE != 1) && (mgi != c.st.npts_model) && (c.cs.thick[mgi] == 1);
  if (MODE != 1 && mgi == c.st.npts_model) {
    edist = DBL_MAX_;
    find_nextline = true;
  } else if (thick) {
    const double kappa = (double)(c.cs.kappagrey[mgi] * c.cs.rho[mgi]) * doppler_packet(c, p);
    edist = (tau_next - 0.0) / kappa;
    find_nextline = true;
  } else {
    if constexpr (MODE == 2) {
      fatal(c, 107);
      p.type = ARTIS_TYPE_ESCAPE;
      return false;
    } else {
      edist = get_event(c, mgi, p, &rpkt_eventtype, tau_next, fmin(tdist, sdist), k, tl);
    }
  }
  const bool do_rlc = (c.cs.do_rlc_est != 0 && c.cs.do_rlc_est != 3);
  // rpkt.cc:700-810: the three outcomes (cell boundary, event, end of the time step) all begin with the same half move,
  // estimator update and second half move over the shortest of the three distances; it is written once here so that the
  // lanes of a warp stay converged through it whatever their outcome is.
  int outcome;  // 0 boundary, 1 event, 2 end of time step
  double dist;
  if ((sdist < tdist) && (sdist < edist)) {
    outcome = 0;
    dist = sdist;
  } else if ((edist < sdist) && (edist < tdist)) {
    outcome = 1;
    dist = edist;
  } else if ((tdist < sdist) && (tdist < edist)) {
    outcome = 2;
    dist = tdist;
  } else {
    fatal(c, 107);  // exact tie: the reference aborts
    p.type = ARTIS_TYPE_ESCAPE;
    return false;
  }
  move_pkt_withtime(c, p, dist / 2.);
  update_estimators(c, p, dist, k, tl);
  if (do_rlc) rlc_emiss_rpkt(c, p, dist, tl);
  if (outcome == 2) {
    p.prop_time = t2;
    move_pkt(c, p, dist / 2.);
  } else {
    move_pkt_withtime(c, p, dist / 2.);
  }
  if (outcome == 0) {
    if (snext != p.where) {
      change_cell(c, p, snext, tl);
      mgi = st.mgi_next;  // = cell_mgi(c, p.where), also when the packet escaped and stays where it is
    }
    // scat_count = 0 (rpkt.cc:729): it is already 0, update_packets cleared it for every packet at the start of the timestep
    p.last_event = p.last_event + 100;
    if (find_nextline) {
      if (p.type == ARTIS_TYPE_RPKT && mgi != c.st.npts_model && c.cs.thick[mgi] != 1) closest_transition_empty(c, p);
    }
    return p.type == ARTIS_TYPE_RPKT;
  }
  if (outcome == 1) {
    if (thick) {
      rpkt_event_thickcell(c, p, rng, tl);
    } else if (rpkt_eventtype == RPKT_EVENTTYPE_BB) {
      rpkt_event_boundbound(c, p, tl);
    } else if (rpkt_eventtype == RPKT_EVENTTYPE_CONT) {
      rpkt_event_continuum(c, p, k, mgi, rng, tl);
    } else {
      fatal(c, 106);
    }
    return p.type == ARTIS_TYPE_RPKT;
  }
  p.last_event = p.last_event + 1000;
  if (find_nextline) closest_transition_empty(c, p);
  return false;
}
// MODE 0: any cell; 1: the caller guarantees a non-grey, non-empty cell; 2: the caller guarantees a grey or empty cell
// (the scheduler's two r-packet kinds, kernels.cuh) -- only prunes code, the arithmetic is the same.
template <int MODE = 0>
ARTIS_DEV bool do_rpkt_step(const Ctx &c, Pkt &p, const double t2, KappaCache &k, Rng &rng, Tally &tl) {
  RpktStepState st;
  if (!rpkt_step_begin<MODE>(c, p, rng, tl, st)) return p.type == ARTIS_TYPE_RPKT;
  return rpkt_step_finish<MODE>(c, p, t2, k, rng, tl, st);
}

}  // namespace artis
