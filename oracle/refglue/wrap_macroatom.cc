// wrapper TU: reference macroatom.cc compiled unmodified + a probe entry to its file-static get_transitionrates
// (macroatom.cc:161), which fills the cellhistory row of one level via calculate_macroatom_transitionrates (:57)
#include "macroatom.cc"
#include "refglue.h"
namespace refglue {
void probe_macroatom_rates(int modelgridindex, int element, int ion, int level, double t_mid, double *rates9) {
  const double *pr = get_transitionrates(modelgridindex, element, ion, level, t_mid, tid);
  for (int a = 0; a < MA_ACTION_COUNT; a++) rates9[a] = pr[a];
}
}  // namespace refglue
