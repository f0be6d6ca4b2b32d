// spectra.cuh -- light curves and the angle-averaged spectrum of the packets resident on the device.
//
// Replaces the packet loop of write_partial_lightcurve_spectra (spectrum.cc:672-682) with its two callees
//     add_to_lc_res   (light_curve.cc:34-54, current_abin == -1)
//     add_to_spec_res (spectrum.cc:602-615 -> add_to_spec, spectrum.cc:339-372, current_abin == -1, no emission resolution:
//                      WRITE_PARTIAL_EMISSIONABSORPTIONSPEC is false in every artisoptions_*.h)
// so that a driver which keeps its packets resident between timesteps downloads ntstep x (4 + nnubins) doubles after a
// timestep instead of npkts x 304 bytes. One thread per packet: a packet that has not escaped costs one 32-byte sector of
// its hot record (the type), an escaped one its hot line and the sector of the cold record that holds the escape record.
#pragma once
#include "common.cuh"

namespace artis {

struct BinSpec {
  double tmin, tmax;       // globals::tmin, globals::tmax
  double nu_min, nu_max;   // globals::nu_min_r / nu_max_r (init_spectra's arguments at spectrum.cc:670)
  double dlognu;           // (log(nu_max) - log(nu_min)) / nnubins, spectrum.cc:354
  int nnubins;             // globals::nnubins (MNUBINS, exspec.h:5)
  int nprocs;              // globals::nprocs
  int ntstep;
  const double *ts_start;  // globals::time_step[].start / .width
  const double *ts_width;
  double vmax;             // globals::vmax
  const float *delta_freq; // spectra->delta_freq[nnubins]: a FLOAT table in the reference (spectrum.h:15, spectrum.cc:498-499)
  double *lc;              // [4][ntstep]: r-packet lum, r-packet lumcmf, gamma lum, gamma lumcmf
  double *flux;            // [ntstep][nnubins]
  double *stokes;          // [3][ntstep][nnubins] (I, Q, U) or nullptr
};

#ifdef __CUDACC__
ARTIS_DEV void bin_add(double *addr, double v) { atomicAdd(addr, v); }
#else
ARTIS_DEV void bin_add(double *addr, double v) { *addr += v; }
#endif

// get_timestep (sn3d.h:168-180) for tmin <= time < tmax: the reference scans the steps from 0, the starts are increasing,
// so the last step whose start is <= time is the same answer.
ARTIS_DEV int bin_timestep(const BinSpec &b, const double time) {
  int lo = 0, hi = b.ntstep - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (b.ts_start[mid] <= time) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// lc_local: the block's private copy of b.lc (shared memory) or b.lc itself
ARTIS_DEV void body_bin_escaped(const BinSpec &b, const PktHot *hot, const PktCold *cold, const long long n, double *lc_local) {
  const double tmin = b.tmin;
  const PktHot &p = hot[n];
  if (pkt_lo16(p.type_cross) != ARTIS_TYPE_ESCAPE) return;
  const PktCold &pc = cold[n];
  const int et = pc.escape_type;
  if (et != ARTIS_TYPE_RPKT && et != ARTIS_TYPE_GAMMA) return;
  double *const lum = lc_local + (et == ARTIS_TYPE_RPKT ? 0 : 2 * b.ntstep);
  double *const lumcmf = lum + b.ntstep;
  const double escape_time = (double)pc.escape_time;  // an int member (packet.h:60)
  // get_arrive_time (vectors.h:146-152)
  const double t_arrive = escape_time - (dot3(p.pos, p.dir) / CLIGHT_PROP);
  const double e_rf = p.e_rf;
  int nt = -1;
  if (t_arrive > tmin && t_arrive < b.tmax) {
    nt = bin_timestep(b, t_arrive);
    bin_add(&lum[nt], e_rf / b.ts_width[nt] / b.nprocs);
  }
  // get_arrive_time_cmf (vectors.h:154-156)
  const double cmf_factor = sqrt(1. - (b.vmax * b.vmax / CLIGHTSQUARED));
  const double t_arrive_cmf = escape_time * cmf_factor;
  if (t_arrive_cmf > tmin && t_arrive_cmf < b.tmax) {
    const int ntc = bin_timestep(b, t_arrive_cmf);
    bin_add(&lumcmf[ntc], p.e_cmf / b.ts_width[ntc] / b.nprocs / cmf_factor);
  }
  if (et != ARTIS_TYPE_RPKT || nt < 0) return;
  const double nu_rf = p.nu_rf;
  if (!(nu_rf > b.nu_min && nu_rf < b.nu_max)) return;
  const int nnu = (int)((log(nu_rf) - log(b.nu_min)) / b.dlognu);
  if (nnu >= b.nnubins) return;  // assert_always in the reference (spectrum.cc:357); cannot happen for nu_rf < nu_max
  constexpr double PARSEC = 3.0857e+18;  // constants.h:20
  const double deltaE = e_rf / b.ts_width[nt] / b.delta_freq[nnu] / 4.e12 / PI / PARSEC / PARSEC / b.nprocs * 1.;
  const size_t bin = (size_t)nt * b.nnubins + nnu;
  bin_add(&b.flux[bin], deltaE);
  if (b.stokes != nullptr) {
    const size_t plane = (size_t)b.ntstep * b.nnubins;
    bin_add(&b.stokes[bin], pc.stokes[0] * deltaE);
    bin_add(&b.stokes[plane + bin], pc.stokes[1] * deltaE);
    bin_add(&b.stokes[2 * plane + bin], pc.stokes[2] * deltaE);
  }
}


// ---- resolved outputs (spectrum.cc:339-452 with do_emission_res; :602-620 and light_curve.cc:55-62 with current_abin >= 0) ----
constexpr int NPHIBINS = 10, NCOSTHETABINS = 10, MABINS = NPHIBINS * NCOSTHETABINS;  // exspec.h:7-9
struct ResSpec {
  double syn_dir[3];           // globals::syn_dir
  int nelements, max_nions;
  const int *line_element, *line_ion;          // linelist[].elementindex / .ionindex
  const int *allcont_of_lut;                   // bflist index (-1 - emissiontype) -> allcont index
  const int *cont_element, *cont_ion;          // allcont[].element / .ion
  int nlines, nbfcontinua;
  double *emission, *trueemission;             // [ntstep][nnubins][proccount], proccount = 2 * nelements * max_nions + 1
  double *absorption;                          // [ntstep][nnubins][nelements * max_nions]
  double *lc_res;                              // [MABINS][ntstep]
  double *flux_res;                            // [MABINS][ntstep][nnubins]
};
// vectors.h:158-192
ARTIS_DEV int escapedirectionbin(const double dir_in[3], const double syn_dir[3]) {
  const double xhat[3] = {1.0, 0.0, 0.0};
  const double dirmag = vec_len(dir_in);
  const double dir[3] = {dir_in[0] / dirmag, dir_in[1] / dirmag, dir_in[2] / dirmag};
  const double costheta = dot3(dir, syn_dir);
  const int costhetabin = (int)((costheta + 1.0) * NPHIBINS / 2.0);
  double vec1[3], vec2[3], vec3[3];
  cross_prod(dir, syn_dir, vec1);
  cross_prod(xhat, syn_dir, vec2);
  const double cosphi = dot3(vec1, vec2) / vec_len(vec1) / vec_len(vec2);
  cross_prod(vec2, syn_dir, vec3);
  const double testphi = dot3(vec1, vec3);
  int phibin;
  if (testphi > 0) {
    phibin = (int)(acos(cosphi) / 2. / PI * NPHIBINS);
  } else {
    phibin = (int)((acos(cosphi) + PI) / 2. / PI * NPHIBINS);
  }
  return (costhetabin * NPHIBINS) + phibin;
}
// spectrum.cc:306-337 columnindex_from_emissiontype; -1 where the reference would assert
ARTIS_DEV int columnindex_from_emissiontype(const ResSpec &r, const int et) {
  if (et >= 0) return et < r.nlines ? r.line_element[et] * r.max_nions + r.line_ion[et] : -1;
  if (et == -9999999) return 2 * r.nelements * r.max_nions;
  const int et_new = -1 - et;
  if (r.nbfcontinua == 0) return 2 * r.nelements * r.max_nions;
  if (et_new >= r.nbfcontinua) return -1;
  const int i = r.allcont_of_lut[et_new];
  return r.nelements * r.max_nions + r.cont_element[i] * r.max_nions + r.cont_ion[i];
}
ARTIS_DEV void body_bin_resolved(const BinSpec &b, const ResSpec &r, const PktHot *hot, const PktCold *cold, const long long n,
                                 int *error_flag) {
  const PktHot &p = hot[n];
  if (pkt_lo16(p.type_cross) != ARTIS_TYPE_ESCAPE) return;
  const PktCold &pc = cold[n];
  if (pc.escape_type != ARTIS_TYPE_RPKT) return;
  const double t_arrive = (double)pc.escape_time - (dot3(p.pos, p.dir) / CLIGHT_PROP);
  if (!(t_arrive > b.tmin && t_arrive < b.tmax)) return;
  const int nt = bin_timestep(b, t_arrive);
  const int abin = escapedirectionbin(p.dir, r.syn_dir);
  if (abin < 0 || abin >= MABINS) {  // assert_always(na < MABINS)
    if (error_flag) *error_flag = 801;
    return;
  }
  bin_add(&r.lc_res[(size_t)abin * b.ntstep + nt], p.e_rf / b.ts_width[nt] * MABINS / b.nprocs);
  const double nu_rf = p.nu_rf;
  if (!(nu_rf > b.nu_min && nu_rf < b.nu_max)) return;
  const int nnu = (int)((log(nu_rf) - log(b.nu_min)) / b.dlognu);
  if (nnu >= b.nnubins) return;
  constexpr double PARSEC = 3.0857e+18;
  const double deltaE = p.e_rf / b.ts_width[nt] / b.delta_freq[nnu] / 4.e12 / PI / PARSEC / PARSEC / b.nprocs * 1.;
  const double deltaE_abin = p.e_rf / b.ts_width[nt] / b.delta_freq[nnu] / 4.e12 / PI / PARSEC / PARSEC / b.nprocs * (double)MABINS;
  bin_add(&r.flux_res[((size_t)abin * b.ntstep + nt) * b.nnubins + nnu], deltaE_abin);
  const int proccount = 2 * r.nelements * r.max_nions + 1, ioncount = r.nelements * r.max_nions;
  const size_t cell = (size_t)nt * b.nnubins;
  const int nproc = columnindex_from_emissiontype(r, pc.emissiontype);
  const int truenproc = columnindex_from_emissiontype(r, pc.trueemissiontype);
  if (nproc < 0 || nproc >= proccount || truenproc < 0 || truenproc >= proccount) {  // assert_always in the reference
    if (error_flag) *error_flag = 802;
    return;
  }
  bin_add(&r.emission[(cell + nnu) * proccount + nproc], deltaE);
  bin_add(&r.trueemission[(cell + nnu) * proccount + truenproc], deltaE);
  const int nnu_abs = (int)((log(pc.absorptionfreq) - log(b.nu_min)) / b.dlognu);
  if (pc.absorptionfreq > 0. && nnu_abs >= 0 && nnu_abs < b.nnubins) {
    const int at = pc.absorptiontype;
    if (at >= 0 && at < r.nlines) {
      const double deltaE_absorption = p.e_rf / b.ts_width[nt] / b.delta_freq[nnu_abs] / 4.e12 / PI / PARSEC / PARSEC / b.nprocs * 1.;
      bin_add(&r.absorption[(cell + nnu_abs) * ioncount + r.line_element[at] * r.max_nions + r.line_ion[at]], deltaE_absorption);
    }
  }
}

// ---- estimator normalisation: update_grid.cc:1041-1134 (estimator_normfactor, normalise_J / normalise_nuJ,
// ffheating / colheating scaling), radfield.cc:1458-1475 get_T_J_from_J, grey_emissivities.cc:124-139 normalise_grey -- per
// model cell, so that a driver whose packets and estimators stay on the device downloads six numbers per cell
constexpr double STEBO = 5.670400e-5;            // constants.h:22
constexpr double ONEOVER4PI = 7.957747153555701e-02;  // constants.h:35
struct NormSpec {
  int npts_model, nprocs;
  double t_mid, tmin, deltat, mintemp, maxtemp;
  const int *nonemptymgi_of_mgi;   // < 0: no propagation cell maps to this model cell (get_numassociatedcells == 0)
  const double *vol_init;          // grid::vol_init_modelcell(mgi)
  const float *TR;                 // the cell's current T_R: get_T_J_from_J falls back to it when J is not finite
  const double *J, *nuJ, *ff, *col, *rpkt_emiss;  // raw estimators
  double *out;                     // [npts_model][6]: J, nuJ, ffheating, colheating, rpkt_emiss (all normalised), T_J
  double *gamma_dep_pathint;       // time_step[nts].gamma_dep_pathint (sum over cells)
};
ARTIS_DEV void body_normalise(const NormSpec &n, const int mgi) {
  double *o = n.out + (size_t)mgi * 6;
  if (n.nonemptymgi_of_mgi[mgi] < 0) {
    for (int k = 0; k < 6; k++) o[k] = 0.;
    return;
  }
  const double deltaV = n.vol_init[mgi] * pow(n.t_mid / n.tmin, 3);
  const double estimator_normfactor = 1 / deltaV / n.deltat / n.nprocs;
  const double estimator_normfactor_over4pi = ONEOVER4PI * estimator_normfactor;
  const double J = n.J[mgi] * estimator_normfactor_over4pi;
  o[0] = J;
  o[1] = n.nuJ[mgi] * estimator_normfactor_over4pi;
  o[2] = n.ff[mgi] * estimator_normfactor;
  o[3] = n.col[mgi] * estimator_normfactor;
  double T_J = pow(J * PI / STEBO, 1. / 4.);
  if (!isfinite(T_J)) {
    T_J = (double)n.TR[mgi];
  } else if (T_J > n.maxtemp) {
    T_J = n.maxtemp;
  } else if (T_J < n.mintemp) {
    T_J = n.mintemp;
  }
  o[5] = T_J;
  // normalise_grey
  bin_add(n.gamma_dep_pathint, n.rpkt_emiss[mgi] * 1.e20 / n.nprocs);
  o[4] = n.rpkt_emiss[mgi] * ONEOVER4PI / deltaV / n.deltat / n.nprocs;
}

// ---- escaped-packet compaction: the escaped packets, in array order, packed for one small D2H -----------
constexpr int ESC_CHUNK = 256;  // packets per block of the count / scatter kernels
// `reported` marks packets that an earlier only_new call has already delivered
ARTIS_DEV bool escaped_selected(const PktHot *hot, const unsigned char *reported, const int only_new, const long long i) {
  return pkt_lo16(hot[i].type_cross) == ARTIS_TYPE_ESCAPE && !(only_new && reported[i]);
}

}  // namespace artis
