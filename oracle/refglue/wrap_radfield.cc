// wrapper TU: reference radfield.cc compiled unmodified + accessors to its file-static estimator arrays
#include "radfield.cc"
#include "refglue.h"
namespace refglue {
double *radfield_J(void) { return radfield::J; }
double *radfield_nuJ(void) { return radfield::nuJ; }
double radfieldbin_nu_upper(int binindex) { return radfield::radfieldbin_nu_upper[binindex]; }
void radfieldbin_solution(int i, float *W, float *T_R) {
  *W = radfield::radfieldbin_solutions[i].W;
  *T_R = radfield::radfieldbin_solutions[i].T_R;
}
void radfieldbin_raw(int i, double *J_raw, double *nuJ_raw, int *contribcount) {
  *J_raw = radfield::radfieldbins[i].J_raw;
  *nuJ_raw = radfield::radfieldbins[i].nuJ_raw;
  *contribcount = radfield::radfieldbins[i].contribcount;
}
#if (DETAILED_BF_ESTIMATORS_ON)
float prev_bfrate_normed(int index) { return radfield::prev_bfrate_normed[index]; }
double bfrate_raw(int index) { return radfield::bfrate_raw[index]; }
#else
float prev_bfrate_normed(int) { return -1.f; }
double bfrate_raw(int) { return 0.; }
#endif
}  // namespace refglue
