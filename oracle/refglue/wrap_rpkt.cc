// wrapper TU: reference rpkt.cc compiled unmodified + a probe entry to its file-static get_event (rpkt.cc:67)
#include "rpkt.cc"
#include "refglue.h"
namespace refglue {
double probe_get_event(int modelgridindex, struct packet *pkt_ptr, int *rpkt_eventtype, double tau_rnd, double abort_dist) {
  globals::kappa_rpkt_cont[tid].recalculate_required = true;  // every probe evaluates the continuum opacity afresh
  return get_event(modelgridindex, pkt_ptr, rpkt_eventtype, tau_rnd, abort_dist);
}
}  // namespace refglue
