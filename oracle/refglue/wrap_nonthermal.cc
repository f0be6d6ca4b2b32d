// wrapper TU: reference nonthermal.cc compiled unmodified + accessors to the file-static Spencer-Fano solution
#include "nonthermal.cc"
#include "refglue.h"
namespace refglue {
float nt_eff_ionpot(int modelgridindex, int uniqueionindex) {
  if (nonthermal::nt_solution == nullptr || nonthermal::nt_solution[modelgridindex].eff_ionpot == nullptr) return 0.f;
  return nonthermal::nt_solution[modelgridindex].eff_ionpot[uniqueionindex];
}
float nt_prob_num_auger(int modelgridindex, int index, bool energyweighted) {
  if (nonthermal::nt_solution == nullptr || nonthermal::nt_solution[modelgridindex].prob_num_auger == nullptr) return 0.f;
  return energyweighted ? nonthermal::nt_solution[modelgridindex].ionenfrac_num_auger[index]
                        : nonthermal::nt_solution[modelgridindex].prob_num_auger[index];
}
double nt_ratecoeff_wfapprox(int modelgridindex, int element, int ion) {
  return nonthermal::nt_ionization_ratecoeff_wfapprox(modelgridindex, element, ion);
}
double nt_deposition_rate_density_raw(int modelgridindex) {
  return nonthermal::deposition_rate_density == nullptr ? 0. : nonthermal::deposition_rate_density[modelgridindex];
}
}  // namespace refglue
