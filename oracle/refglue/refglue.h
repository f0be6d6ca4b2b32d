// refglue.h -- accessors for the reference's file-static state (TEST INFRASTRUCTURE / integration example).
// Each accessor is defined in a wrapper TU (wrap_<file>.cc) that #includes the reference's <file>.cc unmodified
// and appends the accessor, so no reference source is edited or copied.
#ifndef REFGLUE_H
#define REFGLUE_H
namespace refglue {
double *radfield_J(void);    // radfield.cc:100
double *radfield_nuJ(void);  // radfield.cc:108
// multi-bin radiation field and detailed bf estimators (radfield.cc:30-45,73-74), flat index nonemptymgi * N + i
double radfieldbin_nu_upper(int binindex);
void radfieldbin_solution(int mgibinindex, float *W, float *T_R);
void radfieldbin_raw(int mgibinindex, double *J_raw, double *nuJ_raw, int *contribcount);
float prev_bfrate_normed(int index);
double bfrate_raw(int index);
// Spencer-Fano solution of the host solver (nonthermal.cc:95-151) and the work-function fallback (:1362)
float nt_eff_ionpot(int modelgridindex, int uniqueionindex);
float nt_prob_num_auger(int modelgridindex, int index, bool energyweighted);
double nt_ratecoeff_wfapprox(int modelgridindex, int element, int ion);
double nt_deposition_rate_density_raw(int modelgridindex);
// virtual packets (vpkt.cc:15-74): parameters of vpkt.txt, the compile-time grids and the accumulated spectra / maps
struct VpktParams {
  int nobs, nspectra, nrange, nrange_grid, vgrid_flag;
  double tmin_input, tmax_input, tau_max, tmin_grid, tmax_grid, dlogt, dlognu;
  const double *nz_obs, *phiobs, *exclude, *numin_input, *numax_input, *nu_grid_min, *nu_grid_max;
  const float *delta_freq;
};
void vpkt_params(VpktParams *out);
float vpkt_delta_t(int nt);
double vpkt_vstokes(int stokes, int nt, int ind_comb, int nnu);
double vpkt_vgrid(int stokes, int nt, int mt, int bin_range, int bin);
void vpkt_zero_accumulators(void);  // the hook zeroes them before a snapshot step so that the dump is that step's sum
void vpkt_counts(int out[4]);
void coolinglist_entry(int i, int *type, int *level, int *upperlevel);  // kpkt.cc:27-35
int gamma_nlines(int nucindex);                                         // gammapkt.cc:24-31
void gamma_line(int nucindex, int k, double *energy, double *probability);
// probes of file-static functions on the path (wrap_rpkt.cc, wrap_macroatom.cc, wrap_kpkt.cc, wrap_gammapkt.cc)
double probe_get_event(int modelgridindex, struct packet *pkt_ptr, int *rpkt_eventtype, double tau_rnd, double abort_dist);
void probe_macroatom_rates(int modelgridindex, int element, int ion, int level, double t_mid, double *rates9);
double probe_cooling_ion(int modelgridindex, int element, int ion);
double probe_sig_comp(const struct packet *pkt_ptr);
}  // namespace refglue
// the reference's own transport entry point, renamed at compile time in wrap_update_packets.cc
void ref_update_packets(const int my_rank, const int nts, struct packet *packets);
#endif
