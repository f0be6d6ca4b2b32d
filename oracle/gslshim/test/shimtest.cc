// shimtest.cc -- TEST INFRASTRUCTURE: C entry points for tests/test_gslshim.py, which checks the local GSL stand-in
// (oracle/gslshim) against things that are NOT this repository's own code: integrals with analytic answers (the family
// QUADPACK's own tests use) for gsl_integration_qag, and the recurrence of Knuth's subtractive generator for gsl_rng_ran3.
#include <cmath>

#include "gsl/gsl_integration.h"
#include "gsl/gsl_rng.h"

namespace {
struct P {
  double alpha;
};
double f_xalog(double x, void *p) {  // x^alpha ln(1/x): integral over (0,1) = 1/(alpha+1)^2 (QUADPACK test family 1)
  const double a = static_cast<P *>(p)->alpha;
  return x > 0. ? std::pow(x, a) * std::log(1. / x) : 0.;
}
double f_exp(double x, void *) { return std::exp(x); }                   // e - 1 on (0,1)
double f_lorentz(double x, void *p) {                                    // 4^-a / ((x - pi/4)^2 + 16^-a): QUADPACK family 2
  const double a = static_cast<P *>(p)->alpha;
  return std::pow(4., -a) / ((x - M_PI / 4.) * (x - M_PI / 4.) + std::pow(16., -a));
}
double f_cos(double x, void *p) {                                        // cos(2^a sin x) on (0, pi) = pi J0(2^a): family 3
  const double a = static_cast<P *>(p)->alpha;
  return std::cos(std::pow(2., a) * std::sin(x));
}
}  // namespace

extern "C" {
// which: 0 x^a ln(1/x) on (0,1); 1 exp on (0,1); 2 Lorentzian on (0,1); 3 cos(2^a sin x) on (0,pi)
int shim_qag(int which, double alpha, double epsabs, double epsrel, int key, double *result, double *abserr) {
  P par{alpha};
  gsl_function F;
  F.params = &par;
  double a = 0., b = 1.;
  switch (which) {
    case 0: F.function = f_xalog; break;
    case 1: F.function = f_exp; break;
    case 2: F.function = f_lorentz; break;
    default: F.function = f_cos; b = M_PI; break;
  }
  gsl_integration_workspace *w = gsl_integration_workspace_alloc(4096);
  const int rc = gsl_integration_qag(&F, a, b, epsabs, epsrel, 4096, key, w, result, abserr);
  gsl_integration_workspace_free(w);
  return rc;
}
// n raw outputs of gsl_rng_ran3 seeded with `seed`
void shim_ran3(unsigned long seed, int n, unsigned long *out, double *uniform_out) {
  gsl_rng *r = gsl_rng_alloc(gsl_rng_ran3);
  gsl_rng_set(r, seed);
  for (int i = 0; i < n; i++) out[i] = gsl_rng_get(r);
  gsl_rng_set(r, seed);
  for (int i = 0; i < n; i++) uniform_out[i] = gsl_rng_uniform(r);
  gsl_rng_free(r);
}
}
