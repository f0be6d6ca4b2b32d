// wrapper TU: reference kpkt.cc compiled unmodified + accessor to its file-static cooling list
#include "kpkt.cc"
#include "refglue.h"
namespace refglue {
void coolinglist_entry(int i, int *type, int *level, int *upperlevel) {
  *type = kpkt::coolinglist[i].type;
  *level = kpkt::coolinglist[i].level;
  *upperlevel = kpkt::coolinglist[i].upperlevel;
}
}  // namespace refglue
namespace refglue {
// cumulative cooling at the end of one ion's section of the cooling list, computed by the reference's own
// calculate_kpkt_rates_ion (kpkt.cc:167) exactly as do_kpkt calls it (kpkt.cc:527-566)
double probe_cooling_ion(int modelgridindex, int element, int ion) {
  double oldcoolingsum = 0.;
  for (int el = 0; el < get_nelements(); el++)
    for (int i = 0; i < get_nions(el); i++) {
      if (el == element && i == ion) goto done;
      oldcoolingsum += grid::modelgrid[modelgridindex].cooling_contrib_ion[el][i];
    }
done:
  const int ilow = kpkt::get_coolinglistoffset(element, ion);
  kpkt::calculate_kpkt_rates_ion(modelgridindex, element, ion, ilow, oldcoolingsum, tid);
  return globals::cellhistory[tid].cooling_contrib[ilow + kpkt::get_ncoolingterms(element, ion) - 1];
}
}  // namespace refglue
