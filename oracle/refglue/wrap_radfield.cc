// wrapper TU: reference radfield.cc compiled unmodified + accessors to its file-static estimator arrays
#include "radfield.cc"
#include "refglue.h"
namespace refglue {
double *radfield_J(void) { return radfield::J; }
double *radfield_nuJ(void) { return radfield::nuJ; }
}  // namespace refglue
