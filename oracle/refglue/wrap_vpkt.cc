// wrapper TU: reference vpkt.cc compiled unmodified + accessors to its file-scope parameters and accumulators
#include "vpkt.cc"
#include "refglue.h"
namespace refglue {
void vpkt_params(VpktParams *o) {
  o->nobs = Nobs;
  o->nspectra = Nspectra;
  o->nrange = Nrange;
  o->nrange_grid = Nrange_grid;
  o->vgrid_flag = vgrid_flag;
  o->tmin_input = tmin_vspec_input;
  o->tmax_input = tmax_vspec_input;
  o->tau_max = tau_max_vpkt;
  o->tmin_grid = tmin_grid;
  o->tmax_grid = tmax_grid;
  o->dlogt = dlogt_vspec;
  o->dlognu = dlognu_vspec;
  o->nz_obs = nz_obs_vpkt;
  o->phiobs = phiobs;
  o->exclude = exclude;
  o->numin_input = numin_vspec_input;
  o->numax_input = numax_vspec_input;
  o->nu_grid_min = nu_grid_min;
  o->nu_grid_max = nu_grid_max;
  o->delta_freq = delta_freq_vspec;
}
float vpkt_delta_t(int nt) { return vstokes_i[nt][0].delta_t; }
double vpkt_vstokes(int stokes, int nt, int ind_comb, int nnu) {
  struct vspecpol **v = stokes == 0 ? vstokes_i : (stokes == 1 ? vstokes_q : vstokes_u);
  return v[nt][ind_comb].flux[nnu];
}
double vpkt_vgrid(int stokes, int nt, int mt, int bin_range, int bin) {
  struct vgrid *v = stokes == 0 ? &vgrid_i[nt][mt] : (stokes == 1 ? &vgrid_q[nt][mt] : &vgrid_u[nt][mt]);
  return v->flux[bin_range][bin];
}
void vpkt_zero_accumulators(void) {
  if (vstokes_i == nullptr) return;
  for (int nt = 0; nt < VMTBINS; nt++)
    for (int ic = 0; ic < Nspectra * Nobs; ic++)
      for (int m = 0; m < VMNUBINS; m++) vstokes_i[nt][ic].flux[m] = vstokes_q[nt][ic].flux[m] = vstokes_u[nt][ic].flux[m] = 0.;
  if (vgrid_flag == 1)
    for (int n = 0; n < NY_VGRID; n++)
      for (int m = 0; m < NZ_VGRID; m++)
        for (int r = 0; r < Nrange_grid; r++)
          for (int b = 0; b < Nobs; b++) vgrid_i[n][m].flux[r][b] = vgrid_q[n][m].flux[r][b] = vgrid_u[n][m].flux[r][b] = 0.;
  nvpkt = nvpkt_esc1 = nvpkt_esc2 = nvpkt_esc3 = 0;
}
void vpkt_counts(int out[4]) {
  out[0] = nvpkt;
  out[1] = nvpkt_esc1;
  out[2] = nvpkt_esc2;
  out[3] = nvpkt_esc3;
}
}  // namespace refglue
