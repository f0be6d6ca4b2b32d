// GSL shim implementation -- TEST INFRASTRUCTURE ONLY. See gsl/gslshim_all.h for scope and provenance.
// Restates published algorithms; GSL itself (2.7.x, unvendored dependency of the reference) is absent here,
// so parity against real GSL is UNPINNED.
#include "gsl/gslshim_all.h"

#include <cstdio>
#include <cstring>

extern "C" {

// ------------------------------------------------------------------ errno
static gsl_error_handler_t *g_handler = nullptr;
gsl_error_handler_t *gsl_set_error_handler(gsl_error_handler_t *h) {
  gsl_error_handler_t *old = g_handler;
  g_handler = h;
  return old;
}
static void handler_off(const char *, const char *, int, int) {}
gsl_error_handler_t *gsl_set_error_handler_off(void) { return gsl_set_error_handler(&handler_off); }
const char *gsl_strerror(const int e) {
  switch (e) {
    case GSL_SUCCESS: return "success";
    case GSL_FAILURE: return "failure";
    case GSL_CONTINUE: return "the iteration has not converged yet";
    case GSL_EMAXITER: return "exceeded max number of iterations";
    case GSL_EROUND: return "failed because of roundoff error";
    case GSL_ESING: return "apparent singularity detected";
    case GSL_EDIVERGE: return "integral or series is divergent";
    case GSL_EINVAL: return "invalid argument supplied by user";
    default: return "unknown error code";
  }
}
void gsl_error(const char *reason, const char *file, int line, int gsl_errno) {
  if (g_handler) {
    (*g_handler)(reason, file, line, gsl_errno);
    return;
  }
  fprintf(stderr, "gslshim: %s:%d: ERROR: %s (errno %d)\n", file, line, reason, gsl_errno);
  abort();
}
#define SHIM_ERROR(reason, code)                 \
  do {                                           \
    gsl_error(reason, __FILE__, __LINE__, code); \
    return code;                                 \
  } while (0)

// ------------------------------------------------------------------ rng: Knuth subtractive generator (ran3)
static const long M_BIG = 1000000000L;
static const long M_SEED = 161803398L;
static const gsl_rng_type ran3_type = {"ran3", (unsigned long)M_BIG, 0, sizeof(gsl_rng)};
const gsl_rng_type *gsl_rng_ran3 = &ran3_type;

void gsl_rng_set(gsl_rng *r, unsigned long s) {
  if (s == 0) s = 1;
  long j = (M_SEED - (long)s) % M_BIG;
  if (j < 0) j += M_BIG;
  r->buffer[0] = 0;
  r->buffer[55] = j;
  long k = 1;
  for (int i = 1; i < 55; i++) {
    const int n = (21 * i) % 55;
    r->buffer[n] = k;
    k = j - k;
    if (k < 0) k += M_BIG;
    j = (long)r->buffer[n];
  }
  for (int i1 = 0; i1 < 4; i1++) {
    for (int i = 1; i < 56; i++) {
      long t = (long)r->buffer[i] - (long)r->buffer[1 + (i + 30) % 55];
      if (t < 0) t += M_BIG;
      r->buffer[i] = t;
    }
  }
  r->x = 0;
  r->y = 31;
}
gsl_rng *gsl_rng_alloc(const gsl_rng_type *T) {
  gsl_rng *r = (gsl_rng *)calloc(1, sizeof(gsl_rng));
  r->type = T;
  gsl_rng_set(r, 0);
  return r;
}
void gsl_rng_free(gsl_rng *r) { free(r); }
const char *gsl_rng_name(const gsl_rng *r) { return r->type->name; }
unsigned long gsl_rng_get(gsl_rng *r) {
  r->x++;
  if (r->x == 56) r->x = 1;
  r->y++;
  if (r->y == 56) r->y = 1;
  long j = (long)r->buffer[r->x] - (long)r->buffer[r->y];
  if (j < 0) j += M_BIG;
  r->buffer[r->x] = j;
  return (unsigned long)j;
}
// Test hook (not part of GSL): the next gsl_rng_uniform call returns `u` instead of drawing. The reference's
// select_continuum_nu (ratecoeff.cc:643) draws its random number itself; with this its result becomes a deterministic
// function of (continuum, T_e, u) that the probe of oracle/refglue/hook.cc records.
static double g_forced_uniform = -1.;
extern "C" void gslshim_force_uniform(double u) { g_forced_uniform = u; }
double gsl_rng_uniform(gsl_rng *r) {
  if (g_forced_uniform >= 0.) {
    const double u = g_forced_uniform;
    g_forced_uniform = -1.;
    return u;
  }
  return gsl_rng_get(r) / (double)M_BIG;
}
double gsl_rng_uniform_pos(gsl_rng *r) {
  double x;
  do {
    x = gsl_rng_uniform(r);
  } while (x == 0);
  return x;
}

// ------------------------------------------------------------------ vector / matrix
static gsl_block *block_alloc(size_t n, bool zero) {
  gsl_block *b = (gsl_block *)malloc(sizeof(gsl_block));
  b->size = n;
  b->data = (double *)(zero ? calloc(n ? n : 1, sizeof(double)) : malloc((n ? n : 1) * sizeof(double)));
  return b;
}
gsl_vector *gsl_vector_alloc(const size_t n) {
  gsl_vector *v = (gsl_vector *)malloc(sizeof(gsl_vector));
  v->block = block_alloc(n, false);
  v->data = v->block->data;
  v->size = n;
  v->stride = 1;
  v->owner = 1;
  return v;
}
gsl_vector *gsl_vector_calloc(const size_t n) {
  gsl_vector *v = gsl_vector_alloc(n);
  memset(v->data, 0, n * sizeof(double));
  return v;
}
void gsl_vector_free(gsl_vector *v) {
  if (!v) return;
  if (v->owner) {
    free(v->block->data);
    free(v->block);
  }
  free(v);
}
void gsl_vector_set_zero(gsl_vector *v) {
  for (size_t i = 0; i < v->size; i++) v->data[i * v->stride] = 0.;
}
void gsl_vector_set_all(gsl_vector *v, double x) {
  for (size_t i = 0; i < v->size; i++) v->data[i * v->stride] = x;
}
int gsl_vector_memcpy(gsl_vector *d, const gsl_vector *s) {
  if (d->size != s->size) SHIM_ERROR("vector lengths are not equal", GSL_EINVAL);
  for (size_t i = 0; i < s->size; i++) d->data[i * d->stride] = s->data[i * s->stride];
  return GSL_SUCCESS;
}
int gsl_vector_mul(gsl_vector *a, const gsl_vector *b) {
  for (size_t i = 0; i < a->size; i++) a->data[i * a->stride] *= b->data[i * b->stride];
  return GSL_SUCCESS;
}
int gsl_vector_add(gsl_vector *a, const gsl_vector *b) {
  for (size_t i = 0; i < a->size; i++) a->data[i * a->stride] += b->data[i * b->stride];
  return GSL_SUCCESS;
}
int gsl_vector_scale(gsl_vector *a, const double x) {
  for (size_t i = 0; i < a->size; i++) a->data[i * a->stride] *= x;
  return GSL_SUCCESS;
}
int gsl_vector_isnonneg(const gsl_vector *v) {
  for (size_t i = 0; i < v->size; i++)
    if (v->data[i * v->stride] < 0.) return 0;
  return 1;
}
static gsl_vector_view make_vview(double *data, size_t n, size_t stride) {
  gsl_vector_view vv;
  vv.vector.data = data;
  vv.vector.size = n;
  vv.vector.stride = stride;
  vv.vector.block = nullptr;
  vv.vector.owner = 0;
  return vv;
}
gsl_vector_view gsl_vector_view_array(double *v, size_t n) { return make_vview(v, n, 1); }
gsl_vector_view gsl_vector_subvector(gsl_vector *v, size_t i, size_t n) {
  return make_vview(v->data + i * v->stride, n, v->stride);
}
gsl_vector_const_view gsl_vector_const_subvector(const gsl_vector *v, size_t i, size_t n) {
  return make_vview(v->data + i * v->stride, n, v->stride);
}
gsl_matrix *gsl_matrix_alloc(const size_t n1, const size_t n2) {
  gsl_matrix *m = (gsl_matrix *)malloc(sizeof(gsl_matrix));
  m->block = block_alloc(n1 * n2, false);
  m->data = m->block->data;
  m->size1 = n1;
  m->size2 = n2;
  m->tda = n2;
  m->owner = 1;
  return m;
}
gsl_matrix *gsl_matrix_calloc(const size_t n1, const size_t n2) {
  gsl_matrix *m = gsl_matrix_alloc(n1, n2);
  memset(m->data, 0, n1 * n2 * sizeof(double));
  return m;
}
void gsl_matrix_free(gsl_matrix *m) {
  if (!m) return;
  if (m->owner) {
    free(m->block->data);
    free(m->block);
  }
  free(m);
}
void gsl_matrix_set_zero(gsl_matrix *m) {
  for (size_t i = 0; i < m->size1; i++)
    for (size_t j = 0; j < m->size2; j++) m->data[i * m->tda + j] = 0.;
}
int gsl_matrix_memcpy(gsl_matrix *d, const gsl_matrix *s) {
  if (d->size1 != s->size1 || d->size2 != s->size2) SHIM_ERROR("matrix sizes are different", GSL_EINVAL);
  for (size_t i = 0; i < s->size1; i++)
    for (size_t j = 0; j < s->size2; j++) d->data[i * d->tda + j] = s->data[i * s->tda + j];
  return GSL_SUCCESS;
}
int gsl_matrix_add(gsl_matrix *a, const gsl_matrix *b) {
  for (size_t i = 0; i < a->size1; i++)
    for (size_t j = 0; j < a->size2; j++) a->data[i * a->tda + j] += b->data[i * b->tda + j];
  return GSL_SUCCESS;
}
gsl_matrix_view gsl_matrix_view_array(double *base, size_t n1, size_t n2) {
  gsl_matrix_view mv;
  mv.matrix.data = base;
  mv.matrix.size1 = n1;
  mv.matrix.size2 = n2;
  mv.matrix.tda = n2;
  mv.matrix.block = nullptr;
  mv.matrix.owner = 0;
  return mv;
}
gsl_vector_view gsl_matrix_row(gsl_matrix *m, size_t i) { return make_vview(m->data + i * m->tda, m->size2, 1); }
gsl_vector_view gsl_matrix_column(gsl_matrix *m, size_t j) { return make_vview(m->data + j, m->size1, m->tda); }
gsl_vector_const_view gsl_matrix_const_row(const gsl_matrix *m, size_t i) {
  return make_vview(m->data + i * m->tda, m->size2, 1);
}
gsl_vector_const_view gsl_matrix_const_column(const gsl_matrix *m, size_t j) {
  return make_vview(m->data + j, m->size1, m->tda);
}

// ------------------------------------------------------------------ blas
void cblas_dcopy(const int N, const double *X, const int incX, double *Y, const int incY) {
  for (int i = 0; i < N; i++) Y[i * incY] = X[i * incX];
}
void cblas_daxpy(const int N, const double alpha, const double *X, const int incX, double *Y, const int incY) {
  for (int i = 0; i < N; i++) Y[i * incY] += alpha * X[i * incX];
}
double gsl_blas_dasum(const gsl_vector *X) {
  double s = 0.;
  for (size_t i = 0; i < X->size; i++) s += fabs(X->data[i * X->stride]);
  return s;
}
int gsl_blas_ddot(const gsl_vector *X, const gsl_vector *Y, double *result) {
  double s = 0.;
  for (size_t i = 0; i < X->size; i++) s += X->data[i * X->stride] * Y->data[i * Y->stride];
  *result = s;
  return GSL_SUCCESS;
}
size_t gsl_blas_idamax(const gsl_vector *X) {
  size_t imax = 0;
  double vmax = -1.;
  for (size_t i = 0; i < X->size; i++) {
    const double v = fabs(X->data[i * X->stride]);
    if (v > vmax) {
      vmax = v;
      imax = i;
    }
  }
  return imax;
}
int gsl_blas_dgemv(CBLAS_TRANSPOSE_t TransA, double alpha, const gsl_matrix *A, const gsl_vector *X, double beta,
                   gsl_vector *Y) {
  const bool tr = (TransA != CblasNoTrans);
  const size_t ny = tr ? A->size2 : A->size1;
  const size_t nx = tr ? A->size1 : A->size2;
  for (size_t i = 0; i < ny; i++) {
    double s = 0.;
    for (size_t j = 0; j < nx; j++) {
      const double aij = tr ? A->data[j * A->tda + i] : A->data[i * A->tda + j];
      s += aij * X->data[j * X->stride];
    }
    Y->data[i * Y->stride] = alpha * s + beta * Y->data[i * Y->stride];
  }
  return GSL_SUCCESS;
}
void gsl_blas_dscal(double alpha, gsl_vector *X) { gsl_vector_scale(X, alpha); }

}  // extern "C"
