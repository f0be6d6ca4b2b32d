// gamma.cuh -- gamma-ray packets and pellets: gammapkt.cc:227-700, photo_electric.cc,
// grey_emissivities.cc:11-77, update_packets.cc:16-135.
#pragma once
#include "kpkt.cuh"

namespace artis {

// gammapkt.cc:320-333
ARTIS_DEV double sigma_compton_partial(const double x, const double f) {
  const double term1 = ((x * x) - (2 * x) - 2) * log(f) / x / x;
  const double term2 = (((f * f) - 1) / (f * f)) / 2;
  const double term3 = ((f - 1) / x) * ((1 / x) + (2 / f) + (1 / (x * f)));
  return (3 * SIGMA_T * (term1 + term2 + term3) / (8 * x));
}
// gammapkt.cc:335-361
ARTIS_DEV double sig_comp(const Ctx &c, const Pkt &p, const int mgi) {
  const double xx = H * p.nu_cmf / ME / CLIGHT / CLIGHT;
  double sigma_cmf;
  if (xx < THOMSON_LIMIT) {
    sigma_cmf = SIGMA_T;
  } else {
    sigma_cmf = sigma_compton_partial(xx, 1 + (2 * xx));
  }
  sigma_cmf *= c.cs.nnetot[mgi];
  return sigma_cmf * doppler_packet(c, p);
}
// photo_electric.cc:10-49
ARTIS_DEV double sig_photo_electric(const Ctx &c, const Pkt &p, const int mgi) {
  double sigma_cmf;
  const double rho = (double)c.cs.rho[mgi];
  if (c.st.gamma_grey < 0) {
    double sigma_cmf_si = 1.16e-24 * pow(p.nu_cmf / 2.41326e19, -3.13);
    double sigma_cmf_fe = 25.7e-24 * pow(p.nu_cmf / 2.41326e19, -3.0);
    sigma_cmf_si *= rho / MH / 28;
    sigma_cmf_fe *= rho / MH / 56;
    const double f_fe = (double)c.cs.ffegrp[mgi];
    sigma_cmf = (sigma_cmf_fe * f_fe) + (sigma_cmf_si * (1. - f_fe));
  } else {
    sigma_cmf = c.st.gamma_grey * rho;
  }
  return sigma_cmf * doppler_packet(c, p);
}
// photo_electric.cc:51-114
ARTIS_DEV double sig_pair_prod(const Ctx &c, const Pkt &p, const int mgi) {
  double sigma_cmf = 0.0;
  const double rho = (double)c.cs.rho[mgi];
  if (c.st.gamma_grey < 0 && p.nu_cmf > 2.46636e+20) {
    double sigma_cmf_si, sigma_cmf_fe;
    const double f_fe = (double)c.cs.ffegrp[mgi];
    if (p.nu_cmf > 3.61990e+20) {
      sigma_cmf_si = (0.0481 + (0.301 * ((p.nu_cmf / 2.41326e+20) - 1.5))) * 196.e-27;
      sigma_cmf_fe = (0.0481 + (0.301 * ((p.nu_cmf / 2.41326e+20) - 1.5))) * 784.e-27;
    } else {
      sigma_cmf_si = 1.0063 * ((p.nu_cmf / 2.41326e+20) - 1.022) * 196.e-27;
      sigma_cmf_fe = 1.0063 * ((p.nu_cmf / 2.41326e+20) - 1.022) * 784.e-27;
    }
    sigma_cmf_si *= rho / MH / 28;
    sigma_cmf_fe *= rho / MH / 56;
    sigma_cmf = (sigma_cmf_fe * f_fe) + (sigma_cmf_si * (1. - f_fe));
  }
  double sigma_rf = sigma_cmf * doppler_packet(c, p);
  if (sigma_rf < 0) sigma_rf = 0.0;
  return sigma_rf;
}
// grey_emissivities.cc:11-26
ARTIS_DEV double meanf_sigma(const double x) {
  const double f = 1 + (2 * x);
  const double term0 = 2 / x;
  const double term1 = (1 - (2 / x) - (3 / (x * x))) * log(f);
  const double term2 = ((4 / x) + (3 / (x * x)) - 1) * 2 * x / f;
  const double term3 = (1 - (2 / x) - (1 / (x * x))) * 2 * x * (1 + x) / f / f;
  const double term4 = -2. * x * ((4 * x * x) + (6 * x) + 3) / 3 / f / f / f;
  return 3 * SIGMA_T * (term0 + term1 + term2 + term3 + term4) / (8 * x);
}
// grey_emissivities.cc:28-77
ARTIS_DEV void rlc_emiss_gamma(const Ctx &c, const Pkt &p, const int mgi, const double dist, Tally &tl) {
  if (dist > 0) {
    double vel_vec[3];
    get_velocity(p.pos, vel_vec, p.prop_time);
    const double xx = H * p.nu_cmf / ME / CLIGHT / CLIGHT;
    double heating_cont = ((meanf_sigma(xx) * c.cs.nnetot[mgi]) + sig_photo_electric(c, p, mgi) +
                           (sig_pair_prod(c, p, mgi) * (1. - (2.46636e+20 / p.nu_cmf))));
    heating_cont = heating_cont * p.e_rf * dist * (1. - (2. * dot3(vel_vec, p.dir) / CLIGHT));
    tl.estim_atomics += est_add(&c.es.rpkt_emiss[mgi], 1.e-20 * heating_cont);
  }
}
// gammapkt.cc:363-397: bisection on the partial Compton cross-section
ARTIS_DEV double choose_f(const double xx, const double zrand) {
  double f_max = 1 + (2 * xx);
  double f_min = 1;
  const double norm = zrand * sigma_compton_partial(xx, f_max);
  int count = 0;
  double err = 1e20;
  double ftry = (f_max + f_min) / 2;
  while ((err > 1.e-4) && (count < 1000)) {
    ftry = (f_max + f_min) / 2;
    const double sigma_try = sigma_compton_partial(xx, ftry);
    if (sigma_try > norm) {
      f_max = ftry;
      err = (sigma_try - norm) / norm;
    } else {
      f_min = ftry;
      err = (norm - sigma_try) / norm;
    }
    count++;
  }
  return ftry;
}
// gammapkt.cc:399-419
ARTIS_DEV double thomson_angle(const Ctx &c, Rng &rng) {
  const double zrand = rng_uniform(rng);
  const double B_coeff = (8. * zrand) - 4.;
  double t_coeff = sqrt((B_coeff * B_coeff) + 4);
  t_coeff = t_coeff - B_coeff;
  t_coeff = t_coeff / 2;
  t_coeff = cbrt(t_coeff);
  const double mu = (1 / t_coeff) - t_coeff;
  if (fabs(mu) > 1) fatal(c, 401);
  return mu;
}
// gammapkt.cc:421-531
ARTIS_DEV void compton_scatter_body(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
  double f;
  const double xx = H * p.nu_cmf / ME / CLIGHT / CLIGHT;
  bool stay_gamma;
  if (xx < THOMSON_LIMIT) {
    f = 1.0;
    stay_gamma = true;
  } else {
    const double zrand = rng_uniform(rng);
    f = choose_f(xx, zrand);
    if ((f < 1) || (f > (2 * xx + 1))) fatal(c, 402);
    const double prob_gamma = 1. / f;
    const double zrand2 = rng_uniform(rng);
    stay_gamma = (zrand2 < prob_gamma);
  }
  if (stay_gamma) {
    p.nu_cmf = p.nu_cmf / f;
    double vel_vec[3];
    get_velocity(p.pos, vel_vec, p.prop_time);
    double cmf_dir[3];
    angle_ab(p.dir, vel_vec, cmf_dir);
    double cos_theta;
    if (xx < THOMSON_LIMIT) {
      cos_theta = thomson_angle(c, rng);
    } else {
      cos_theta = 1. - ((f - 1) / xx);
    }
    double new_dir[3];
    scatter_dir(rng, cmf_dir, cos_theta, new_dir);
    if (fabs(1. - dot3(new_dir, new_dir)) > 1.e-8) fatal(c, 403);
    if (fabs(dot3(new_dir, cmf_dir) - cos_theta) > 1.e-8) fatal(c, 404);
    vec_scale(vel_vec, -1.);
    double final_dir[3];
    angle_ab(new_dir, vel_vec, final_dir);
    p.dir[0] = final_dir[0];
    p.dir[1] = final_dir[1];
    p.dir[2] = final_dir[2];
    const double dopplerfactor = doppler_packet(c, p);
    p.nu_rf = p.nu_cmf / dopplerfactor;
    p.e_rf = p.e_cmf / dopplerfactor;
    p.last_cross = ARTIS_CROSS_NONE;
  } else {
    p.type = ARTIS_TYPE_NTLEPTON;
    p.cold->absorptiontype = -3;
    stat_inc(tl, COUNTER_NT_STAT_FROM_GAMMA);
  }
}
ARTIS_DEV_NOINLINE void compton_scatter_ool(const Ctx &c, EvState *s) { compton_scatter_body(c, s->p, s->rng, s->tl); }
ARTIS_DEV void compton_scatter(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  compton_scatter_ool(c, &s);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  compton_scatter_body(c, p, rng, tl);
#endif
}
// photo_electric.cc:116-164
ARTIS_DEV void pair_prod_body(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
  const double prob_gamma = 1.022 * MEV / (H * p.nu_cmf);
  if (prob_gamma < 0) fatal(c, 405);
  const double zrand = rng_uniform(rng);
  if (zrand > prob_gamma) {
    p.type = ARTIS_TYPE_NTLEPTON;
    p.cold->absorptiontype = -5;
    stat_inc(tl, COUNTER_NT_STAT_FROM_GAMMA);
  } else {
    p.nu_cmf = 0.511 * MEV / H;
    double dir_cmf[3];
    get_rand_isotropic_unitvec(rng, dir_cmf);
    double vel_vec[3];
    get_velocity(p.pos, vel_vec, -1. * p.prop_time);
    angle_ab(dir_cmf, vel_vec, p.dir);
    const double dopplerfactor = doppler_packet(c, p);
    p.nu_rf = p.nu_cmf / dopplerfactor;
    p.e_rf = p.e_cmf / dopplerfactor;
    p.type = ARTIS_TYPE_GAMMA;
    p.last_cross = ARTIS_CROSS_NONE;
  }
}
ARTIS_DEV_NOINLINE void pair_prod_ool(const Ctx &c, EvState *s) { pair_prod_body(c, s->p, s->rng, s->tl); }
ARTIS_DEV void pair_prod(const Ctx &c, Pkt &p, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  pair_prod_ool(c, &s);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  pair_prod_body(c, p, rng, tl);
#endif
}

// gammapkt.cc:533-700 do_gamma (GRID_UNIFORM)
ARTIS_DEV void do_gamma(const Ctx &c, Pkt &p, const double t2, Rng &rng, Tally &tl) {
  tl.gamma_steps++;
  const double zrand0 = rng_uniform_pos(rng);
  const double tau_next = -1. * log(zrand0);
  int snext = 0;
  double sdist = boundary_cross(c, p, &snext);
  const double maxsdist = max_sdist(c, p, sdist);
  if (sdist > maxsdist) {
    fatal(c, 410);
    p.type = ARTIS_TYPE_ESCAPE;
    return;
  }
  if (sdist < 0) sdist = 0;
  if (((snext < 0) && (snext != -99)) || (snext >= c.st.ngrid)) {
    fatal(c, 411);
    p.type = ARTIS_TYPE_ESCAPE;
    return;
  }
  if (sdist > c.cs.max_path_step) {
    sdist = c.cs.max_path_step;
    snext = p.where;
  }
  const int mgi = cell_mgi(c, p.where);
  double kap_compton = 0.0;
  if (c.st.gamma_grey < 0) kap_compton = sig_comp(c, p, mgi);
  const double kap_photo_electric = sig_photo_electric(c, p, mgi);
  const double kap_pair_prod = sig_pair_prod(c, p, mgi);
  const double kap_tot = kap_compton + kap_photo_electric + kap_pair_prod;
  const double edist = (tau_next - 0.0) / kap_tot;
  const double tdist = (t2 - p.prop_time) * CLIGHT_PROP;
  if (edist < 0 || tdist < 0) {
    fatal(c, 412);
    p.type = ARTIS_TYPE_ESCAPE;
    return;
  }
  const bool est = (kap_tot > 0) && (c.cs.do_rlc_est != 0);  // do_comp_est is false whenever do_r_lc is set (sn3d.cc:539)
  if ((sdist < tdist) && (sdist < edist)) {
    p.prop_time += sdist / 2. / CLIGHT_PROP;
    move_pkt(c, p, sdist / 2.);
    if (est) rlc_emiss_gamma(c, p, mgi, sdist, tl);
    p.prop_time += sdist / 2. / CLIGHT_PROP;
    move_pkt(c, p, sdist / 2.);
    if (snext != p.where) change_cell(c, p, snext, tl);
  } else if ((tdist < sdist) && (tdist < edist)) {
    p.prop_time += tdist / 2. / CLIGHT_PROP;
    move_pkt(c, p, tdist / 2.);
    if (est) rlc_emiss_gamma(c, p, mgi, tdist, tl);
    p.prop_time = t2;
    move_pkt(c, p, tdist / 2.);
  } else if ((edist < sdist) && (edist < tdist)) {
    p.prop_time += edist / 2. / CLIGHT_PROP;
    move_pkt(c, p, edist / 2.);
    if (est) rlc_emiss_gamma(c, p, mgi, edist, tl);
    p.prop_time += edist / 2. / CLIGHT_PROP;
    move_pkt(c, p, edist / 2.);
    const double zrand = rng_uniform(rng);
    if (kap_compton > (zrand * kap_tot)) {
      compton_scatter(c, p, rng, tl);
    } else if ((kap_compton + kap_photo_electric) > (zrand * kap_tot)) {
      p.type = ARTIS_TYPE_NTLEPTON;
      p.cold->absorptiontype = -4;
      stat_inc(tl, COUNTER_NT_STAT_FROM_GAMMA);
    } else if ((kap_compton + kap_photo_electric + kap_pair_prod) > (zrand * kap_tot)) {
      pair_prod(c, p, rng, tl);
    } else {
      fatal(c, 413);
      p.type = ARTIS_TYPE_ESCAPE;
    }
  } else {
    fatal(c, 414);
    p.type = ARTIS_TYPE_ESCAPE;
  }
}

// gammapkt.cc:227-318 choose_gamma_ray + pellet_gamma_decay
ARTIS_DEV void pellet_gamma_decay(const Ctx &c, Pkt &p, Rng &rng) {
  const int nuc = ldcold(&p.cold->pellet_nucindex);
  const int off = c.st.gam_offset[nuc];
  const int nlines = c.st.gam_offset[nuc + 1] - off;
  if (nlines == 0) {
    p.type = ARTIS_TYPE_KPKT;
    p.cold->absorptiontype = -6;
    return;
  }
  double dir_cmf[3];
  get_rand_isotropic_unitvec(rng, dir_cmf);
  double vel_vec[3];
  const double tdecay = ldcold(&p.cold->tdecay);
  get_velocity(p.pos, vel_vec, -1. * tdecay);
  angle_ab(dir_cmf, vel_vec, p.dir);
  const double E_gamma = c.st.nuc_decayenergygamma[nuc];
  const double zrand = rng_uniform(rng);
  int nselected = -1;
  double runtot = 0.;
  for (int n = 0; n < nlines; n++) {
    runtot += c.st.gam_prob[off + n] * c.st.gam_energy[off + n] / E_gamma;
    if (zrand <= runtot) {
      nselected = n;
      break;
    }
  }
  if (nselected < 0) {
    fatal(c, 420);
    nselected = nlines - 1;
  }
  p.nu_cmf = c.st.gam_energy[off + nselected] / H;
  p.prop_time = tdecay;
  const double dopplerfactor = doppler_packet(c, p);
  p.nu_rf = p.nu_cmf / dopplerfactor;
  p.e_rf = p.e_cmf / dopplerfactor;
  p.type = ARTIS_TYPE_GAMMA;
  p.last_cross = ARTIS_CROSS_NONE;
  p.cold->stokes[0] = 1.0;
  p.cold->stokes[1] = p.cold->stokes[2] = 0.0;
  set_pol_dir(p);
}

// update_packets.cc:71-135 update_pellet
ARTIS_DEV void update_pellet_body(const Ctx &c, Pkt &p, const double t2, Rng &rng, Tally &tl) {
  const double ts = p.prop_time;
  const double tdecay = ldcold(&p.cold->tdecay);
  double *tsc = c.es.ts_scalars;
  if (tdecay > t2) {
    vec_scale(p.pos, t2 / ts);
    p.prop_time = t2;
  } else if (tdecay > ts) {
    ctr_add(&c.es.counters[ARTIS_CTR_PELLET_DECAYS], 1);
    p.prop_time = tdecay;
    vec_scale(p.pos, tdecay / ts);
    const int decaytype = ldcold(&p.cold->pellet_decaytype);
    if ((ldcold(&p.cold->origin_bytes) & 0xffu)) {
      if (decaytype == ARTIS_DECAYTYPE_BETAPLUS) {
        est_add(&tsc[ARTIS_TS_POSITRON_DEP], p.e_cmf);
        p.type = ARTIS_TYPE_NTLEPTON;
        p.cold->absorptiontype = -10;
      } else if (decaytype == ARTIS_DECAYTYPE_BETAMINUS) {
        est_add(&tsc[ARTIS_TS_ELECTRON_EMISSION], p.e_cmf);
        p.cold->em_time = (int)p.prop_time;
        p.type = ARTIS_TYPE_NONTHERMAL_PREDEPOSIT;
        p.cold->absorptiontype = -10;
      } else if (decaytype == ARTIS_DECAYTYPE_ALPHA) {
        est_add(&tsc[ARTIS_TS_ALPHA_EMISSION], p.e_cmf);
        p.cold->em_time = (int)p.prop_time;
        p.type = ARTIS_TYPE_NONTHERMAL_PREDEPOSIT;
        p.cold->absorptiontype = -10;
      }
    } else {
      est_add(&tsc[ARTIS_TS_GAMMA_EMISSION], p.e_cmf);
      pellet_gamma_decay(c, p, rng);
    }
  } else if ((tdecay > 0) && (c.nts == 0)) {
    p.e_cmf *= tdecay / c.st.tmin;
    p.type = ARTIS_TYPE_PRE_KPKT;
    p.cold->absorptiontype = -7;
    stat_inc(tl, COUNTER_K_STAT_FROM_EARLIERDECAY);
    p.prop_time = c.st.tmin;
  } else {
    fatal(c, 430);
    p.prop_time = t2;
  }
}
ARTIS_DEV_NOINLINE void update_pellet_ool(const Ctx &c, EvState *s, const double t2) { update_pellet_body(c, s->p, t2, s->rng, s->tl); }
ARTIS_DEV void update_pellet(const Ctx &c, Pkt &p, const double t2, Rng &rng, Tally &tl) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  s.tl = tl;
  update_pellet_ool(c, &s, t2);
  p = s.p;
  rng = s.rng;
  tl = s.tl;
#else
  update_pellet_body(c, p, t2, rng, tl);
#endif
}

// update_packets.cc:16-69 do_nonthermal_predeposit
ARTIS_DEV void do_nonthermal_predeposit_body(const Ctx &c, Pkt &p, const double t2, Rng &rng) {
  const double ts = p.prop_time;
  const int decaytype = ldcold(&p.cold->pellet_decaytype);
  const double particle_en = H * p.nu_cmf;
  double endot = 0.;
  double t_absorb = ts;
  if (!ARTIS_OPT_INSTANT_DEPOSITION(c)) {
    const double rho = (double)c.cs.rho[cell_mgi(c, p.where)];
    endot = (decaytype == ARTIS_DECAYTYPE_ALPHA) ? 5.e11 * MEV * rho : 4.e10 * MEV * rho;
    const double zrand = rng_uniform(rng);
    const double en_absorb = zrand * particle_en;
    t_absorb = ts + en_absorb / endot;
  }
  double *tsc = c.es.ts_scalars;
  if (t_absorb <= t2) {
    if (decaytype == ARTIS_DECAYTYPE_ALPHA) {
      est_add(&tsc[ARTIS_TS_ALPHA_DEP], p.e_cmf);
    } else if (decaytype == ARTIS_DECAYTYPE_BETAMINUS) {
      est_add(&tsc[ARTIS_TS_ELECTRON_DEP], p.e_cmf);
    } else if (decaytype == ARTIS_DECAYTYPE_BETAPLUS) {
      est_add(&tsc[ARTIS_TS_POSITRON_DEP], p.e_cmf);
    }
    vec_scale(p.pos, t_absorb / ts);
    p.prop_time = t_absorb;
    p.type = ARTIS_TYPE_NTLEPTON;
  } else {
    p.nu_cmf = (particle_en - endot * (t2 - ts)) / H;
    vec_scale(p.pos, t2 / ts);
    p.prop_time = t2;
  }
}
ARTIS_DEV_NOINLINE void do_nonthermal_predeposit_ool(const Ctx &c, EvState *s, const double t2) { do_nonthermal_predeposit_body(c, s->p, t2, s->rng); }
ARTIS_DEV void do_nonthermal_predeposit(const Ctx &c, Pkt &p, const double t2, Rng &rng) {
#if ARTIS_EVENT_OOL
  EvState s;
  s.p = p;
  s.rng = rng;
  do_nonthermal_predeposit_ool(c, &s, t2);
  p = s.p;
  rng = s.rng;
#else
  do_nonthermal_predeposit_body(c, p, t2, rng);
#endif
}

// update_packets.cc:137-202 do_packet. The r-packet case runs ONE do_rpkt_step instead of do_rpkt's whole loop, so
// that the scheduler can interleave packet kinds; everything the reference does after do_rpkt returns (cmf_lum on
// escape) happens as soon as the packet escapes.
ARTIS_DEV void do_packet(const Ctx &c, Pkt &p, const double t2, KappaCache &k, Rng &rng, Tally &tl) {
  tl.packet_updates++;
  switch (p.type) {
    case ARTIS_TYPE_RADIOACTIVE_PELLET:
      update_pellet(c, p, t2, rng, tl);
      break;
    case ARTIS_TYPE_GAMMA:
      do_gamma(c, p, t2, rng, tl);
      if (p.type != ARTIS_TYPE_GAMMA && p.type != ARTIS_TYPE_ESCAPE) est_add(&c.es.ts_scalars[ARTIS_TS_GAMMA_DEP], p.e_cmf);
      break;
    case ARTIS_TYPE_RPKT:
      do_rpkt_step(c, p, t2, k, rng, tl);
      if (p.type == ARTIS_TYPE_ESCAPE) est_add(&c.es.ts_scalars[ARTIS_TS_CMF_LUM], p.e_cmf);
      break;
    case ARTIS_TYPE_NONTHERMAL_PREDEPOSIT:
      do_nonthermal_predeposit(c, p, t2, rng);
      break;
    case ARTIS_TYPE_NTLEPTON:
      do_ntlepton(c, p, rng, tl);
      break;
    case ARTIS_TYPE_KPKT:
    case ARTIS_TYPE_PRE_KPKT:
    case ARTIS_TYPE_GAMMA_KPKT: {
      const int mgi = cell_mgi(c, p.where);
      if (p.type == ARTIS_TYPE_PRE_KPKT || c.cs.thick[mgi] == 1) {
        do_kpkt_bb(c, p, rng, tl);
      } else if (p.type == ARTIS_TYPE_KPKT) {
        do_kpkt(c, p, t2, rng, tl);
      } else {
        fatal(c, 440);
        p.prop_time = t2;
      }
      break;
    }
    case ARTIS_TYPE_MA:
      do_macroatom(c, p, rng, tl);
      break;
    default:
      fatal(c, 441);
      p.prop_time = t2;
  }
}

// do_packet restricted to the packet types of one scheduler kind (kernels.cuh kind_of): each role only carries the code
// and registers of the physics it can meet. KIND 5 (macro-atom jumps) and the macro-atom branch of KIND 0 run the whole
// iteration of do_macroatom here; the device scheduler splits it into ma_jump / ma_event itself (transport.cu).
template <int KIND>
ARTIS_DEV void do_packet_kind(const Ctx &c, Pkt &p, const double t2, KappaCache &k, Rng &rng, Tally &tl) {
  if constexpr (KIND == 1 || KIND == 4) {
    tl.packet_updates++;
    do_rpkt_step<(KIND == 1 ? 1 : 2)>(c, p, t2, k, rng, tl);
    if (p.type == ARTIS_TYPE_ESCAPE) est_add(&c.es.ts_scalars[ARTIS_TS_CMF_LUM], p.e_cmf);
  } else if constexpr (KIND == 2) {
    tl.packet_updates++;
    do_gamma(c, p, t2, rng, tl);
    if (p.type != ARTIS_TYPE_GAMMA && p.type != ARTIS_TYPE_ESCAPE) est_add(&c.es.ts_scalars[ARTIS_TS_GAMMA_DEP], p.e_cmf);
  } else if constexpr (KIND == 5) {
    tl.packet_updates++;
    do_macroatom(c, p, rng, tl);
  } else if constexpr (KIND == 0) {
    tl.packet_updates++;
    if (p.type == ARTIS_TYPE_MA) {
      do_macroatom(c, p, rng, tl);
    } else if (p.type == ARTIS_TYPE_NTLEPTON) {
      do_ntlepton(c, p, rng, tl);
    } else {
      const int mgi = cell_mgi(c, p.where);
      if (p.type == ARTIS_TYPE_PRE_KPKT || c.cs.thick[mgi] == 1) {
        do_kpkt_bb(c, p, rng, tl);
      } else if (p.type == ARTIS_TYPE_KPKT) {
        do_kpkt(c, p, t2, rng, tl);
      } else {
        fatal(c, 440);
        p.prop_time = t2;
      }
    }
  } else {
    // KIND_OTHER: pellets and non-thermal pre-deposition only (update_packets.cc:137-202 for these two types) -- calling
    // the whole do_packet here would put a second copy of every other kind's physics into the kernel
    tl.packet_updates++;
    if (p.type == ARTIS_TYPE_RADIOACTIVE_PELLET) {
      update_pellet(c, p, t2, rng, tl);
    } else if (p.type == ARTIS_TYPE_NONTHERMAL_PREDEPOSIT) {
      do_nonthermal_predeposit(c, p, t2, rng);
    } else {
      fatal(c, 441);
      p.prop_time = t2;
    }
  }
}

}  // namespace artis
