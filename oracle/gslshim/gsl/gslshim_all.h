// GSL shim -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// GNU GSL is a dependency of the reference (Makefile:47-48 `pkg-config gsl`; CI installs Ubuntu 22.04's
// libgsl-dev 2.7.x, ci.yml:87,139) but is absent from this image and from /root/reference. This header
// declares the subset of the GSL API that the reference's translation units use, so that they can be compiled
// UNMODIFIED from where they lie (see oracle/Makefile). The implementations in gslshim.cc restate the published
// algorithms (Knuth's subtractive generator "ran3", QUADPACK QAG/QAGS, Brent-Dekker root bracketing, LU with
// partial pivoting, Householder solve, Debye-function series). Parity of these against real GSL is UNPINNED
// (no GSL here to compare with); parity of the hot path is defined against the reference built on this shim.
#ifndef GSLSHIM_ALL_H
#define GSLSHIM_ALL_H

#include <math.h>   // like the real gsl_math.h: in C++ this brings the float overloads of sqrt/exp/... into the global namespace
#include <stddef.h>
#include <stdlib.h>

#ifdef __cplusplus
extern "C" {
#endif

// ---- errno ----
enum {
  GSL_SUCCESS = 0,
  GSL_FAILURE = -1,
  GSL_CONTINUE = -2,
  GSL_EDOM = 1,
  GSL_ERANGE = 2,
  GSL_EINVAL = 4,
  GSL_EFAILED = 5,
  GSL_EMAXITER = 11,
  GSL_EROUND = 18,
  GSL_ESING = 21,
  GSL_EDIVERGE = 22
};
typedef void gsl_error_handler_t(const char *reason, const char *file, int line, int gsl_errno);
gsl_error_handler_t *gsl_set_error_handler(gsl_error_handler_t *new_handler);
gsl_error_handler_t *gsl_set_error_handler_off(void);
const char *gsl_strerror(const int gsl_errno);
void gsl_error(const char *reason, const char *file, int line, int gsl_errno);

// ---- math ----
struct gsl_function_struct {
  double (*function)(double x, void *params);
  void *params;
};
typedef struct gsl_function_struct gsl_function;
#define GSL_FN_EVAL(F, x) (*((F)->function))(x, (F)->params)
struct gsl_function_fdf_struct {
  double (*f)(double x, void *params);
  double (*df)(double x, void *params);
  void (*fdf)(double x, void *params, double *f, double *df);
  void *params;
};
typedef struct gsl_function_fdf_struct gsl_function_fdf;
#define GSL_DBL_EPSILON 2.2204460492503131e-16
#define GSL_DBL_MIN 2.2250738585072014e-308
#define GSL_DBL_MAX 1.7976931348623157e+308
#define GSL_MAX(a, b) ((a) > (b) ? (a) : (b))
#define GSL_MIN(a, b) ((a) < (b) ? (a) : (b))

// ---- rng ----
typedef struct {
  const char *name;
  unsigned long max, min;
  size_t size;
} gsl_rng_type;
typedef struct {
  const gsl_rng_type *type;
  unsigned int x, y;
  unsigned long buffer[56];
} gsl_rng;
extern const gsl_rng_type *gsl_rng_ran3;
gsl_rng *gsl_rng_alloc(const gsl_rng_type *T);
void gsl_rng_free(gsl_rng *r);
void gsl_rng_set(gsl_rng *r, unsigned long seed);
const char *gsl_rng_name(const gsl_rng *r);
unsigned long gsl_rng_get(gsl_rng *r);
double gsl_rng_uniform(gsl_rng *r);
double gsl_rng_uniform_pos(gsl_rng *r);

// ---- integration ----
enum { GSL_INTEG_GAUSS15 = 1, GSL_INTEG_GAUSS21 = 2, GSL_INTEG_GAUSS31 = 3, GSL_INTEG_GAUSS41 = 4, GSL_INTEG_GAUSS51 = 5, GSL_INTEG_GAUSS61 = 6 };
typedef struct {
  size_t limit, size, nrmax, i, maximum_level;
  double *alist, *blist, *rlist, *elist;
  size_t *order, *level;
} gsl_integration_workspace;
gsl_integration_workspace *gsl_integration_workspace_alloc(const size_t n);
void gsl_integration_workspace_free(gsl_integration_workspace *w);
int gsl_integration_qag(const gsl_function *f, double a, double b, double epsabs, double epsrel, size_t limit, int key,
                        gsl_integration_workspace *workspace, double *result, double *abserr);
int gsl_integration_qags(const gsl_function *f, double a, double b, double epsabs, double epsrel, size_t limit,
                         gsl_integration_workspace *workspace, double *result, double *abserr);
int gsl_integration_qagp(const gsl_function *f, double *pts, size_t npts, double epsabs, double epsrel, size_t limit,
                         gsl_integration_workspace *workspace, double *result, double *abserr);
int gsl_integration_qng(const gsl_function *f, double a, double b, double epsabs, double epsrel, double *result,
                        double *abserr, size_t *neval);

// ---- block / vector / matrix ----
typedef struct {
  size_t size;
  double *data;
} gsl_block;
typedef struct {
  size_t size, stride;
  double *data;
  gsl_block *block;
  int owner;
} gsl_vector;
typedef struct {
  gsl_vector vector;
} gsl_vector_view;
typedef const gsl_vector_view gsl_vector_const_view;
typedef struct {
  size_t size1, size2, tda;
  double *data;
  gsl_block *block;
  int owner;
} gsl_matrix;
typedef struct {
  gsl_matrix matrix;
} gsl_matrix_view;
typedef const gsl_matrix_view gsl_matrix_const_view;

gsl_vector *gsl_vector_alloc(const size_t n);
gsl_vector *gsl_vector_calloc(const size_t n);
void gsl_vector_free(gsl_vector *v);
static inline double gsl_vector_get(const gsl_vector *v, const size_t i) { return v->data[i * v->stride]; }
static inline void gsl_vector_set(gsl_vector *v, const size_t i, double x) { v->data[i * v->stride] = x; }
static inline double *gsl_vector_ptr(gsl_vector *v, const size_t i) { return v->data + i * v->stride; }
void gsl_vector_set_zero(gsl_vector *v);
void gsl_vector_set_all(gsl_vector *v, double x);
int gsl_vector_memcpy(gsl_vector *dest, const gsl_vector *src);
int gsl_vector_mul(gsl_vector *a, const gsl_vector *b);
int gsl_vector_add(gsl_vector *a, const gsl_vector *b);
int gsl_vector_scale(gsl_vector *a, const double x);
int gsl_vector_isnonneg(const gsl_vector *v);
gsl_vector_view gsl_vector_view_array(double *v, size_t n);
gsl_vector_view gsl_vector_subvector(gsl_vector *v, size_t i, size_t n);
gsl_vector_const_view gsl_vector_const_subvector(const gsl_vector *v, size_t i, size_t n);

gsl_matrix *gsl_matrix_alloc(const size_t n1, const size_t n2);
gsl_matrix *gsl_matrix_calloc(const size_t n1, const size_t n2);
void gsl_matrix_free(gsl_matrix *m);
static inline double gsl_matrix_get(const gsl_matrix *m, const size_t i, const size_t j) { return m->data[i * m->tda + j]; }
static inline void gsl_matrix_set(gsl_matrix *m, const size_t i, const size_t j, double x) { m->data[i * m->tda + j] = x; }
static inline double *gsl_matrix_ptr(gsl_matrix *m, const size_t i, const size_t j) { return m->data + (i * m->tda + j); }
void gsl_matrix_set_zero(gsl_matrix *m);
int gsl_matrix_memcpy(gsl_matrix *dest, const gsl_matrix *src);
int gsl_matrix_add(gsl_matrix *a, const gsl_matrix *b);
gsl_matrix_view gsl_matrix_view_array(double *base, size_t n1, size_t n2);
gsl_vector_view gsl_matrix_row(gsl_matrix *m, size_t i);
gsl_vector_view gsl_matrix_column(gsl_matrix *m, size_t j);
gsl_vector_const_view gsl_matrix_const_row(const gsl_matrix *m, size_t i);
gsl_vector_const_view gsl_matrix_const_column(const gsl_matrix *m, size_t j);

// ---- blas ----
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };
typedef enum CBLAS_TRANSPOSE CBLAS_TRANSPOSE_t;
void cblas_dcopy(const int N, const double *X, const int incX, double *Y, const int incY);
void cblas_daxpy(const int N, const double alpha, const double *X, const int incX, double *Y, const int incY);
double gsl_blas_dasum(const gsl_vector *X);
int gsl_blas_ddot(const gsl_vector *X, const gsl_vector *Y, double *result);
size_t gsl_blas_idamax(const gsl_vector *X);
int gsl_blas_dgemv(CBLAS_TRANSPOSE_t TransA, double alpha, const gsl_matrix *A, const gsl_vector *X, double beta,
                   gsl_vector *Y);
void gsl_blas_dscal(double alpha, gsl_vector *X);

// ---- permutation / linalg ----
typedef struct {
  size_t size;
  size_t *data;
} gsl_permutation;
gsl_permutation *gsl_permutation_alloc(const size_t n);
gsl_permutation *gsl_permutation_calloc(const size_t n);
void gsl_permutation_free(gsl_permutation *p);
int gsl_linalg_LU_decomp(gsl_matrix *A, gsl_permutation *p, int *signum);
int gsl_linalg_LU_solve(const gsl_matrix *LU, const gsl_permutation *p, const gsl_vector *b, gsl_vector *x);
int gsl_linalg_LU_refine(const gsl_matrix *A, const gsl_matrix *LU, const gsl_permutation *p, const gsl_vector *b,
                         gsl_vector *x, gsl_vector *work);
int gsl_linalg_HH_solve(gsl_matrix *A, const gsl_vector *b, gsl_vector *x);
int gsl_linalg_balance_matrix(gsl_matrix *A, gsl_vector *D);

// ---- roots ----
typedef struct {
  const char *name;
} gsl_root_fsolver_type;
typedef struct {
  const gsl_root_fsolver_type *type;
  gsl_function *function;
  double root, x_lower, x_upper;
  double a, b, c, d, e, fa, fb, fc;
} gsl_root_fsolver;
extern const gsl_root_fsolver_type *gsl_root_fsolver_brent;
gsl_root_fsolver *gsl_root_fsolver_alloc(const gsl_root_fsolver_type *T);
void gsl_root_fsolver_free(gsl_root_fsolver *s);
int gsl_root_fsolver_set(gsl_root_fsolver *s, gsl_function *f, double x_lower, double x_upper);
int gsl_root_fsolver_iterate(gsl_root_fsolver *s);
double gsl_root_fsolver_root(const gsl_root_fsolver *s);
double gsl_root_fsolver_x_lower(const gsl_root_fsolver *s);
double gsl_root_fsolver_x_upper(const gsl_root_fsolver *s);
int gsl_root_test_interval(double x_lower, double x_upper, double epsabs, double epsrel);
typedef struct {
  const char *name;
} gsl_root_fdfsolver_type;
typedef struct {
  const gsl_root_fdfsolver_type *type;
} gsl_root_fdfsolver;
extern const gsl_root_fdfsolver_type *gsl_root_fdfsolver_newton;
gsl_root_fdfsolver *gsl_root_fdfsolver_alloc(const gsl_root_fdfsolver_type *T);

// ---- special functions ----
double gsl_sf_debye_3(const double x);
double gsl_sf_debye_4(const double x);
double gsl_sf_expint_E1(const double x);

// ---- interpolation (linear only) ----
typedef struct {
  const char *name;
} gsl_interp_type;
extern const gsl_interp_type *gsl_interp_linear;
typedef struct {
  size_t cache;
} gsl_interp_accel;
typedef struct {
  size_t size;
  double *x, *y;
} gsl_spline;
gsl_interp_accel *gsl_interp_accel_alloc(void);
void gsl_interp_accel_free(gsl_interp_accel *a);
gsl_spline *gsl_spline_alloc(const gsl_interp_type *T, size_t size);
int gsl_spline_init(gsl_spline *spline, const double xa[], const double ya[], size_t size);
double gsl_spline_eval(const gsl_spline *spline, double x, gsl_interp_accel *a);
void gsl_spline_free(gsl_spline *spline);

#ifdef __cplusplus
}
#endif
#endif  // GSLSHIM_ALL_H
