// GSL shim, numerical part -- TEST INFRASTRUCTURE ONLY. See gsl/gslshim_all.h.
// QUADPACK QAG (Piessens, de Doncker-Kapenga, Ueberhuber, Kahaner 1983), Brent-Dekker root bracketing,
// LU with partial pivoting, Householder solve, Debye functions, linear interpolation.
#include <cstdio>
#include <cstring>
#include <vector>

#include "gsl/gslshim_all.h"
#include "gk_tables.inc"

extern "C" {

// ------------------------------------------------------------------ Gauss-Kronrod rule (QUADPACK qk)
static double rescale_error(double err, const double result_abs, const double result_asc) {
  err = fabs(err);
  if (result_asc != 0 && err != 0) {
    const double scale = pow((200 * err / result_asc), 1.5);
    err = (scale < 1) ? result_asc * scale : result_asc;
  }
  if (result_abs > GSL_DBL_MIN / (50 * GSL_DBL_EPSILON)) {
    const double min_err = 50 * GSL_DBL_EPSILON * result_abs;
    if (min_err > err) err = min_err;
  }
  return err;
}

static void qk(const int n, const double xgk[], const double wg[], const double wgk[], const gsl_function *f, double a,
               double b, double *result, double *abserr, double *resabs, double *resasc) {
  double fv1[32], fv2[32];
  const double center = 0.5 * (a + b);
  const double half_length = 0.5 * (b - a);
  const double abs_half_length = fabs(half_length);
  const double f_center = GSL_FN_EVAL(f, center);
  double result_gauss = 0;
  double result_kronrod = f_center * wgk[n - 1];
  double result_abs = fabs(result_kronrod);
  if (n % 2 == 0) result_gauss = f_center * wg[n / 2 - 1];
  for (int j = 0; j < (n - 1) / 2; j++) {
    const int jtw = j * 2 + 1;
    const double abscissa = half_length * xgk[jtw];
    const double fval1 = GSL_FN_EVAL(f, center - abscissa);
    const double fval2 = GSL_FN_EVAL(f, center + abscissa);
    const double fsum = fval1 + fval2;
    fv1[jtw] = fval1;
    fv2[jtw] = fval2;
    result_gauss += wg[j] * fsum;
    result_kronrod += wgk[jtw] * fsum;
    result_abs += wgk[jtw] * (fabs(fval1) + fabs(fval2));
  }
  for (int j = 0; j < n / 2; j++) {
    const int jtwm1 = j * 2;
    const double abscissa = half_length * xgk[jtwm1];
    const double fval1 = GSL_FN_EVAL(f, center - abscissa);
    const double fval2 = GSL_FN_EVAL(f, center + abscissa);
    fv1[jtwm1] = fval1;
    fv2[jtwm1] = fval2;
    result_kronrod += wgk[jtwm1] * (fval1 + fval2);
    result_abs += wgk[jtwm1] * (fabs(fval1) + fabs(fval2));
  }
  const double mean = result_kronrod * 0.5;
  double result_asc = wgk[n - 1] * fabs(f_center - mean);
  for (int j = 0; j < n - 1; j++) result_asc += wgk[j] * (fabs(fv1[j] - mean) + fabs(fv2[j] - mean));
  const double err = (result_kronrod - result_gauss) * half_length;
  result_kronrod *= half_length;
  result_abs *= abs_half_length;
  result_asc *= abs_half_length;
  *result = result_kronrod;
  *resabs = result_abs;
  *resasc = result_asc;
  *abserr = rescale_error(err, result_abs, result_asc);
}

static void qk_key(int key, const gsl_function *f, double a, double b, double *r, double *e, double *ra, double *rs) {
  switch (key) {
    case GSL_INTEG_GAUSS15: qk(8, xgk15, wg15, wgk15, f, a, b, r, e, ra, rs); break;
    case GSL_INTEG_GAUSS21: qk(11, xgk21, wg21, wgk21, f, a, b, r, e, ra, rs); break;
    case GSL_INTEG_GAUSS31: qk(16, xgk31, wg31, wgk31, f, a, b, r, e, ra, rs); break;
    case GSL_INTEG_GAUSS41: qk(21, xgk41, wg41, wgk41, f, a, b, r, e, ra, rs); break;
    case GSL_INTEG_GAUSS51: qk(26, xgk51, wg51, wgk51, f, a, b, r, e, ra, rs); break;
    default: qk(31, xgk61, wg61, wgk61, f, a, b, r, e, ra, rs); break;
  }
}

gsl_integration_workspace *gsl_integration_workspace_alloc(const size_t n) {
  gsl_integration_workspace *w = (gsl_integration_workspace *)calloc(1, sizeof(gsl_integration_workspace));
  w->limit = n;
  w->alist = (double *)malloc(n * sizeof(double));
  w->blist = (double *)malloc(n * sizeof(double));
  w->rlist = (double *)malloc(n * sizeof(double));
  w->elist = (double *)malloc(n * sizeof(double));
  w->order = (size_t *)malloc(n * sizeof(size_t));
  w->level = (size_t *)malloc(n * sizeof(size_t));
  return w;
}
void gsl_integration_workspace_free(gsl_integration_workspace *w) {
  if (!w) return;
  free(w->alist);
  free(w->blist);
  free(w->rlist);
  free(w->elist);
  free(w->order);
  free(w->level);
  free(w);
}

// Adaptive bisection of the interval with the largest error estimate (QUADPACK QAG). The "order" array of
// QUADPACK's qpsrt is replaced by a linear arg-max scan; the sequence of bisections is the same.
int gsl_integration_qag(const gsl_function *f, double a, double b, double epsabs, double epsrel, size_t limit, int key,
                        gsl_integration_workspace *w, double *result, double *abserr) {
  if (key < GSL_INTEG_GAUSS15) key = GSL_INTEG_GAUSS15;
  if (key > GSL_INTEG_GAUSS61) key = GSL_INTEG_GAUSS61;
  *result = 0;
  *abserr = 0;
  if (limit > w->limit) {
    gsl_error("iteration limit exceeds available workspace", __FILE__, __LINE__, GSL_EINVAL);
    return GSL_EINVAL;
  }
  if (epsabs <= 0 && (epsrel < 50 * GSL_DBL_EPSILON || epsrel < 0.5e-28)) {
    gsl_error("tolerance cannot be achieved with given epsabs and epsrel", __FILE__, __LINE__, 13);
    return 13;
  }
  double result0, abserr0, resabs0, resasc0;
  qk_key(key, f, a, b, &result0, &abserr0, &resabs0, &resasc0);
  w->size = 1;
  w->alist[0] = a;
  w->blist[0] = b;
  w->rlist[0] = result0;
  w->elist[0] = abserr0;
  double tolerance = GSL_MAX(epsabs, epsrel * fabs(result0));
  const double round_off = 50 * GSL_DBL_EPSILON * resabs0;
  if (abserr0 <= round_off && abserr0 > tolerance) {
    *result = result0;
    *abserr = abserr0;
    gsl_error("cannot reach tolerance because of roundoff error on first attempt", __FILE__, __LINE__, GSL_EROUND);
    return GSL_EROUND;
  } else if ((abserr0 <= tolerance && abserr0 != resasc0) || abserr0 == 0.0) {
    *result = result0;
    *abserr = abserr0;
    return GSL_SUCCESS;
  } else if (limit == 1) {
    *result = result0;
    *abserr = abserr0;
    gsl_error("a maximum of one iteration was insufficient", __FILE__, __LINE__, GSL_EMAXITER);
    return GSL_EMAXITER;
  }
  double area = result0, errsum = abserr0;
  size_t iteration = 1;
  int roundoff_type1 = 0, roundoff_type2 = 0, error_type = 0;
  do {
    size_t imax = 0;
    for (size_t i = 1; i < w->size; i++)
      if (w->elist[i] > w->elist[imax]) imax = i;
    const double a_i = w->alist[imax], b_i = w->blist[imax], r_i = w->rlist[imax], e_i = w->elist[imax];
    const double a1 = a_i, b1 = 0.5 * (a_i + b_i), a2 = b1, b2 = b_i;
    double area1, area2, error1, error2, resabs1, resabs2, resasc1, resasc2;
    qk_key(key, f, a1, b1, &area1, &error1, &resabs1, &resasc1);
    qk_key(key, f, a2, b2, &area2, &error2, &resabs2, &resasc2);
    const double area12 = area1 + area2, error12 = error1 + error2;
    errsum += (error12 - e_i);
    area += area12 - r_i;
    if (resasc1 != error1 && resasc2 != error2) {
      const double delta = r_i - area12;
      if (fabs(delta) <= 1.0e-5 * fabs(area12) && error12 >= 0.99 * e_i) roundoff_type1++;
      if (iteration >= 10 && error12 > e_i) roundoff_type2++;
    }
    tolerance = GSL_MAX(epsabs, epsrel * fabs(area));
    if (errsum > tolerance) {
      if (roundoff_type1 >= 6 || roundoff_type2 >= 20) error_type = 2;
      const double tmp = (1 + 100 * GSL_DBL_EPSILON) * (fabs(a2) + 1000 * GSL_DBL_MIN);
      if (fabs(a1) <= tmp && fabs(b2) <= tmp) error_type = 3;
    }
    // keep the sub-interval with the larger error in slot imax (as QUADPACK's update does)
    const size_t inew = w->size;
    if (error2 > error1) {
      w->alist[imax] = a2;
      w->blist[imax] = b2;
      w->rlist[imax] = area2;
      w->elist[imax] = error2;
      w->alist[inew] = a1;
      w->blist[inew] = b1;
      w->rlist[inew] = area1;
      w->elist[inew] = error1;
    } else {
      w->blist[imax] = b1;
      w->rlist[imax] = area1;
      w->elist[imax] = error1;
      w->alist[inew] = a2;
      w->blist[inew] = b2;
      w->rlist[inew] = area2;
      w->elist[inew] = error2;
    }
    w->size++;
    iteration++;
  } while (iteration < limit && !error_type && errsum > tolerance);
  double sum = 0;
  for (size_t i = 0; i < w->size; i++) sum += w->rlist[i];
  *result = sum;
  *abserr = errsum;
  if (errsum <= tolerance) return GSL_SUCCESS;
  if (error_type == 2) {
    gsl_error("roundoff error prevents tolerance from being achieved", __FILE__, __LINE__, GSL_EROUND);
    return GSL_EROUND;
  }
  if (error_type == 3) {
    gsl_error("bad integrand behavior found in the integration interval", __FILE__, __LINE__, GSL_ESING);
    return GSL_ESING;
  }
  if (iteration == limit) {
    gsl_error("maximum number of subdivisions reached", __FILE__, __LINE__, GSL_EMAXITER);
    return GSL_EMAXITER;
  }
  gsl_error("could not integrate function", __FILE__, __LINE__, GSL_EFAILED);
  return GSL_EFAILED;
}

// QAGS / QAGP / QNG are only reached from multi-bin radiation-field host code (radfield.cc:1836-1840); the shim
// maps them onto QAG-GK21 (no epsilon extrapolation), splitting at the user break points for QAGP.
int gsl_integration_qags(const gsl_function *f, double a, double b, double epsabs, double epsrel, size_t limit,
                         gsl_integration_workspace *w, double *result, double *abserr) {
  return gsl_integration_qag(f, a, b, epsabs, epsrel, limit, GSL_INTEG_GAUSS21, w, result, abserr);
}
int gsl_integration_qagp(const gsl_function *f, double *pts, size_t npts, double epsabs, double epsrel, size_t limit,
                         gsl_integration_workspace *w, double *result, double *abserr) {
  double sum = 0, err = 0;
  int status = GSL_SUCCESS;
  for (size_t i = 0; i + 1 < npts; i++) {
    double r = 0, e = 0;
    const int s = gsl_integration_qag(f, pts[i], pts[i + 1], epsabs / (npts - 1), epsrel, limit, GSL_INTEG_GAUSS21, w,
                                      &r, &e);
    if (s != GSL_SUCCESS) status = s;
    sum += r;
    err += e;
  }
  *result = sum;
  *abserr = err;
  return status;
}
int gsl_integration_qng(const gsl_function *f, double a, double b, double epsabs, double epsrel, double *result,
                        double *abserr, size_t *neval) {
  gsl_integration_workspace *w = gsl_integration_workspace_alloc(64);
  const int s = gsl_integration_qag(f, a, b, epsabs, epsrel, 64, GSL_INTEG_GAUSS21, w, result, abserr);
  if (neval) *neval = 21 * (2 * w->size - 1);
  gsl_integration_workspace_free(w);
  return s;
}

// ------------------------------------------------------------------ roots: Brent-Dekker
static const gsl_root_fsolver_type brent_type = {"brent"};
const gsl_root_fsolver_type *gsl_root_fsolver_brent = &brent_type;
static const gsl_root_fdfsolver_type newton_type = {"newton"};
const gsl_root_fdfsolver_type *gsl_root_fdfsolver_newton = &newton_type;
gsl_root_fdfsolver *gsl_root_fdfsolver_alloc(const gsl_root_fdfsolver_type *T) {
  gsl_root_fdfsolver *s = (gsl_root_fdfsolver *)calloc(1, sizeof(gsl_root_fdfsolver));
  s->type = T;
  return s;
}
gsl_root_fsolver *gsl_root_fsolver_alloc(const gsl_root_fsolver_type *T) {
  gsl_root_fsolver *s = (gsl_root_fsolver *)calloc(1, sizeof(gsl_root_fsolver));
  s->type = T;
  return s;
}
void gsl_root_fsolver_free(gsl_root_fsolver *s) { free(s); }
int gsl_root_fsolver_set(gsl_root_fsolver *s, gsl_function *f, double x_lower, double x_upper) {
  s->function = f;
  s->root = 0.5 * (x_lower + x_upper);
  s->x_lower = x_lower;
  s->x_upper = x_upper;
  const double f_lower = GSL_FN_EVAL(f, x_lower);
  const double f_upper = GSL_FN_EVAL(f, x_upper);
  s->a = x_lower;
  s->fa = f_lower;
  s->b = x_upper;
  s->fb = f_upper;
  s->c = x_upper;
  s->fc = f_upper;
  s->d = x_upper - x_lower;
  s->e = x_upper - x_lower;
  if ((f_lower < 0.0 && f_upper < 0.0) || (f_lower > 0.0 && f_upper > 0.0)) {
    gsl_error("endpoints do not straddle y=0", __FILE__, __LINE__, GSL_EINVAL);
    return GSL_EINVAL;
  }
  return GSL_SUCCESS;
}
int gsl_root_fsolver_iterate(gsl_root_fsolver *s) {
  double a = s->a, b = s->b, c = s->c, fa = s->fa, fb = s->fb, fc = s->fc, d = s->d, e = s->e;
  bool ac_equal = false;
  if ((fb < 0 && fc < 0) || (fb > 0 && fc > 0)) {
    ac_equal = true;
    c = a;
    fc = fa;
    d = b - a;
    e = b - a;
  }
  if (fabs(fc) < fabs(fb)) {
    ac_equal = true;
    a = b;
    b = c;
    c = a;
    fa = fb;
    fb = fc;
    fc = fa;
  }
  const double tol = 0.5 * GSL_DBL_EPSILON * fabs(b);
  const double m = 0.5 * (c - b);
  if (fb == 0) {
    s->root = b;
    s->x_lower = b;
    s->x_upper = b;
    return GSL_SUCCESS;
  }
  if (fabs(m) <= tol) {
    s->root = b;
    if (b < c) {
      s->x_lower = b;
      s->x_upper = c;
    } else {
      s->x_lower = c;
      s->x_upper = b;
    }
    return GSL_SUCCESS;
  }
  if (fabs(e) < tol || fabs(fa) <= fabs(fb)) {
    d = m;
    e = m;
  } else {
    double p, q, r;
    const double sv = fb / fa;
    if (ac_equal) {
      p = 2 * m * sv;
      q = 1 - sv;
    } else {
      q = fa / fc;
      r = fb / fc;
      p = sv * (2 * m * q * (q - r) - (b - a) * (r - 1));
      q = (q - 1) * (r - 1) * (sv - 1);
    }
    if (p > 0)
      q = -q;
    else
      p = -p;
    if (2 * p < GSL_MIN(3 * m * q - fabs(tol * q), fabs(e * q))) {
      e = d;
      d = p / q;
    } else {
      d = m;
      e = m;
    }
  }
  a = b;
  fa = fb;
  if (fabs(d) > tol)
    b += d;
  else
    b += (m > 0 ? +tol : -tol);
  fb = GSL_FN_EVAL(s->function, b);
  s->a = a;
  s->b = b;
  s->c = c;
  s->d = d;
  s->e = e;
  s->fa = fa;
  s->fb = fb;
  s->fc = fc;
  s->root = b;
  if ((fb < 0 && fc < 0) || (fb > 0 && fc > 0)) c = a;
  if (b < c) {
    s->x_lower = b;
    s->x_upper = c;
  } else {
    s->x_lower = c;
    s->x_upper = b;
  }
  return GSL_SUCCESS;
}
double gsl_root_fsolver_root(const gsl_root_fsolver *s) { return s->root; }
double gsl_root_fsolver_x_lower(const gsl_root_fsolver *s) { return s->x_lower; }
double gsl_root_fsolver_x_upper(const gsl_root_fsolver *s) { return s->x_upper; }
int gsl_root_test_interval(double x_lower, double x_upper, double epsabs, double epsrel) {
  const double abs_lower = fabs(x_lower), abs_upper = fabs(x_upper);
  double min_abs;
  if ((x_lower > 0.0 && x_upper > 0.0) || (x_lower < 0.0 && x_upper < 0.0))
    min_abs = GSL_MIN(abs_lower, abs_upper);
  else
    min_abs = 0;
  const double tolerance = epsabs + epsrel * min_abs;
  return (fabs(x_upper - x_lower) < tolerance) ? GSL_SUCCESS : GSL_CONTINUE;
}

// ------------------------------------------------------------------ permutation / linalg
gsl_permutation *gsl_permutation_alloc(const size_t n) {
  gsl_permutation *p = (gsl_permutation *)malloc(sizeof(gsl_permutation));
  p->size = n;
  p->data = (size_t *)malloc((n ? n : 1) * sizeof(size_t));
  return p;
}
gsl_permutation *gsl_permutation_calloc(const size_t n) {
  gsl_permutation *p = gsl_permutation_alloc(n);
  for (size_t i = 0; i < n; i++) p->data[i] = i;
  return p;
}
void gsl_permutation_free(gsl_permutation *p) {
  if (!p) return;
  free(p->data);
  free(p);
}
int gsl_linalg_LU_decomp(gsl_matrix *A, gsl_permutation *p, int *signum) {
  const size_t N = A->size1;
  *signum = 1;
  for (size_t i = 0; i < N; i++) p->data[i] = i;
  for (size_t j = 0; j + 1 < N; j++) {
    double max = fabs(gsl_matrix_get(A, j, j));
    size_t i_pivot = j;
    for (size_t i = j + 1; i < N; i++) {
      const double aij = fabs(gsl_matrix_get(A, i, j));
      if (aij > max) {
        max = aij;
        i_pivot = i;
      }
    }
    if (i_pivot != j) {
      for (size_t k = 0; k < A->size2; k++) {
        const double t = gsl_matrix_get(A, j, k);
        gsl_matrix_set(A, j, k, gsl_matrix_get(A, i_pivot, k));
        gsl_matrix_set(A, i_pivot, k, t);
      }
      const size_t t = p->data[j];
      p->data[j] = p->data[i_pivot];
      p->data[i_pivot] = t;
      *signum = -(*signum);
    }
    const double ajj = gsl_matrix_get(A, j, j);
    if (ajj != 0.0) {
      for (size_t i = j + 1; i < N; i++) {
        const double aij = gsl_matrix_get(A, i, j) / ajj;
        gsl_matrix_set(A, i, j, aij);
        double *rowi = A->data + i * A->tda;
        const double *rowj = A->data + j * A->tda;
        for (size_t k = j + 1; k < N; k++) rowi[k] -= aij * rowj[k];
      }
    }
  }
  return GSL_SUCCESS;
}
int gsl_linalg_LU_solve(const gsl_matrix *LU, const gsl_permutation *p, const gsl_vector *b, gsl_vector *x) {
  const size_t N = LU->size1;
  for (size_t i = 0; i < N; i++) {
    if (gsl_matrix_get(LU, i, i) == 0) {
      gsl_error("matrix is singular", __FILE__, __LINE__, GSL_EDOM);
      return GSL_EDOM;
    }
  }
  std::vector<double> y(N);
  for (size_t i = 0; i < N; i++) y[i] = gsl_vector_get(b, p->data[i]);
  for (size_t i = 0; i < N; i++) {
    double s = y[i];
    const double *row = LU->data + i * LU->tda;
    for (size_t k = 0; k < i; k++) s -= row[k] * y[k];
    y[i] = s;
  }
  for (size_t ii = N; ii-- > 0;) {
    double s = y[ii];
    const double *row = LU->data + ii * LU->tda;
    for (size_t k = ii + 1; k < N; k++) s -= row[k] * y[k];
    y[ii] = s / row[ii];
  }
  for (size_t i = 0; i < N; i++) gsl_vector_set(x, i, y[i]);
  return GSL_SUCCESS;
}
int gsl_linalg_LU_refine(const gsl_matrix *A, const gsl_matrix *LU, const gsl_permutation *p, const gsl_vector *b,
                         gsl_vector *x, gsl_vector *work) {
  gsl_vector_memcpy(work, b);
  gsl_blas_dgemv(CblasNoTrans, 1.0, A, x, -1.0, work);  // work = A x - b
  gsl_vector *corr = gsl_vector_alloc(x->size);
  const int status = gsl_linalg_LU_solve(LU, p, work, corr);
  for (size_t i = 0; i < x->size; i++) gsl_vector_set(x, i, gsl_vector_get(x, i) - gsl_vector_get(corr, i));
  gsl_vector_free(corr);
  return status;
}
int gsl_linalg_HH_solve(gsl_matrix *A, const gsl_vector *b, gsl_vector *x) {
  // the shim solves via LU (A is overwritten, as with the Householder routine)
  gsl_permutation *p = gsl_permutation_alloc(A->size1);
  int s;
  gsl_linalg_LU_decomp(A, p, &s);
  const int status = gsl_linalg_LU_solve(A, p, b, x);
  gsl_permutation_free(p);
  return status;
}
int gsl_linalg_balance_matrix(gsl_matrix *A, gsl_vector *D) {
  gsl_vector_set_all(D, 1.0);
  (void)A;
  return GSL_SUCCESS;
}

// ------------------------------------------------------------------ special functions
static double debye_n(const int n, const double x) {
  if (x <= 0) return 1.0;
  static const double B2k[] = {1. / 6,        -1. / 30,        1. / 42,      -1. / 30,   5. / 66,
                               -691. / 2730,  7. / 6,          -3617. / 510, 43867. / 798, -174611. / 330};
  if (x < 1.0) {
    double sum = 1.0 - n * x / (2.0 * (n + 1));
    double x2k = 1.0, fact = 1.0;
    for (int k = 1; k <= 10; k++) {
      x2k *= x * x;
      fact *= (2 * k - 1) * (2 * k);
      sum += n * B2k[k - 1] * x2k / ((2 * k + n) * fact);
    }
    return sum;
  }
  // D_n(x) = n/x^n [ n! zeta(n+1) - sum_k e^{-kx} sum_{j=0..n} n!/(n-j)! x^{n-j} / k^{j+1} ]
  const double zeta4 = 1.0823232337111381915, zeta5 = 1.0369277551433699263;
  double nfact = 1.0;
  for (int i = 2; i <= n; i++) nfact *= i;
  double total = nfact * (n == 3 ? zeta4 : zeta5);
  for (int k = 1; k < 200; k++) {
    const double ek = exp(-k * x);
    if (ek < 1e-20) break;
    double term = 0, coef = 1.0, xp = pow(x, n), kp = k;
    for (int j = 0; j <= n; j++) {
      term += coef * xp / kp;
      coef *= (n - j);
      xp /= x;
      kp *= k;
    }
    total -= ek * term;
  }
  return n * total / pow(x, n);
}
double gsl_sf_debye_3(const double x) { return debye_n(3, x); }
double gsl_sf_debye_4(const double x) { return debye_n(4, x); }
double gsl_sf_expint_E1(const double x) {
  if (x <= 0) return NAN;
  if (x < 1.0) {
    double sum = -0.57721566490153286061 - log(x), term = 1.0;
    for (int k = 1; k < 60; k++) {
      term *= -x / k;
      sum -= term / k;
    }
    return sum;
  }
  // continued fraction (modified Lentz)
  double b = x + 1.0, c = 1e300, d = 1.0 / b, h = d;
  for (int i = 1; i < 200; i++) {
    const double an = -i * (double)i;
    b += 2.0;
    d = 1.0 / (an * d + b);
    c = b + an / c;
    const double del = c * d;
    h *= del;
    if (fabs(del - 1.0) < 1e-16) break;
  }
  return h * exp(-x);
}

// ------------------------------------------------------------------ linear interpolation
static const gsl_interp_type linear_type = {"linear"};
const gsl_interp_type *gsl_interp_linear = &linear_type;
gsl_interp_accel *gsl_interp_accel_alloc(void) { return (gsl_interp_accel *)calloc(1, sizeof(gsl_interp_accel)); }
void gsl_interp_accel_free(gsl_interp_accel *a) { free(a); }
gsl_spline *gsl_spline_alloc(const gsl_interp_type *, size_t size) {
  gsl_spline *s = (gsl_spline *)malloc(sizeof(gsl_spline));
  s->size = size;
  s->x = (double *)malloc(size * sizeof(double));
  s->y = (double *)malloc(size * sizeof(double));
  return s;
}
int gsl_spline_init(gsl_spline *s, const double xa[], const double ya[], size_t size) {
  memcpy(s->x, xa, size * sizeof(double));
  memcpy(s->y, ya, size * sizeof(double));
  return GSL_SUCCESS;
}
double gsl_spline_eval(const gsl_spline *s, double x, gsl_interp_accel *) {
  if (x < s->x[0] || x > s->x[s->size - 1]) {
    gsl_error("interpolation error", __FILE__, __LINE__, GSL_EDOM);
    return NAN;
  }
  size_t lo = 0, hi = s->size - 1;
  while (hi - lo > 1) {
    const size_t mid = (lo + hi) / 2;
    if (s->x[mid] > x)
      hi = mid;
    else
      lo = mid;
  }
  const double dx = s->x[hi] - s->x[lo];
  return dx > 0 ? s->y[lo] + (x - s->x[lo]) / dx * (s->y[hi] - s->y[lo]) : s->y[lo];
}
void gsl_spline_free(gsl_spline *s) {
  if (!s) return;
  free(s->x);
  free(s->y);
  free(s);
}

}  // extern "C"
