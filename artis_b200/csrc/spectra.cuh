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


// ---- escaped-packet compaction: the escaped packets, in array order, packed for one small D2H -----------
constexpr int ESC_CHUNK = 256;  // packets per block of the count / scatter kernels
// `reported` marks packets that an earlier only_new call has already delivered
ARTIS_DEV bool escaped_selected(const PktHot *hot, const unsigned char *reported, const int only_new, const long long i) {
  return pkt_lo16(hot[i].type_cross) == ARTIS_TYPE_ESCAPE && !(only_new && reported[i]);
}

}  // namespace artis
