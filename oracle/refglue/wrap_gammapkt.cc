// wrapper TU: reference gammapkt.cc compiled unmodified + accessors to its file-static gamma line lists
#include "gammapkt.cc"
#include "refglue.h"
namespace refglue {
int gamma_nlines(int nucindex) { return gammapkt::gamma_spectra[nucindex].nlines; }
void gamma_line(int nucindex, int k, double *energy, double *probability) {
  *energy = gammapkt::gamma_spectra[nucindex].energy[k];
  *probability = gammapkt::gamma_spectra[nucindex].probability[k];
}
}  // namespace refglue
namespace refglue {
double probe_sig_comp(const struct packet *pkt_ptr) { return gammapkt::sig_comp(pkt_ptr); }  // gammapkt.cc:335
}  // namespace refglue
