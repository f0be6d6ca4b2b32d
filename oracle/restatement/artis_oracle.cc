// artis_oracle.cc -- THE ORACLE'S CPU RESTATEMENT. TEST INFRASTRUCTURE: never shipped, never imported by artis_b200.
//
// Builds oracle/libartis_oracle.so (`make -C oracle restatement`): the restatement of the reference's update_packets
// path (artis_b200/csrc/*.cuh -- every function there cites the reference file:line it follows) compiled as plain
// host C++ and run one packet after another, behind the same C ABI as the product. Only tests/, bench.py's
// cpu_baseline leg and __graft_entry__.smoke() load it, and only as the checker.
//
// Pinning: compiled with the flags of the reference build (oracle/Makefile: -O3 -march=x86-64-v3) this restatement
// reproduces the outputs of the reference's own boundary_cross, closest_transition, calculate_kappa_rpkt_cont,
// get_event, get_levelpop, calculate_macroatom_transitionrates, calculate_kpkt_rates_ion, sig_comp,
// sig_photo_electric and sig_pair_prod BIT FOR BIT (<= 4e-16 relative for the macro-atom rates) on the probe sets
// stored in tests/golden/*.npz, which were produced by the unmodified reference sources built against the GSL shim
// (oracle/_ref, oracle/refglue/hook.cc). The full Monte Carlo step is pinned statistically against ensembles of
// reference runs (tests/statlib.py). See tests/test_oracle.py.
//
// The product library (artis_b200/csrc/transport.cu -> artis_b200/libartis_gpu.so) has no CPU path: without a CUDA
// device every entry point fails with ARTIS_ERR_NODEVICE.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../artis_b200/csrc/kernels.cuh"
#include "../../artis_b200/csrc/spectra.cuh"

namespace {
artis::Ctx g_ctx;
std::chrono::steady_clock::time_point g_t0;
bool be_available(std::string *) { return true; }
bool be_select_device(int, std::string *) { return true; }
void *be_alloc(size_t n) { return calloc(1, n ? n : 1); }
void be_free(void *p) { free(p); }
void be_h2d(void *d, const void *s, size_t n) { memcpy(d, s, n); }
void be_d2h(void *d, const void *s, size_t n) { memcpy(d, s, n); }
void be_zero(void *p, size_t n) { memset(p, 0, n); }
void be_set_ctx(const artis::Ctx &c) { g_ctx = c; }
int be_sync(std::string *) { return 0; }
void be_timer_begin() { g_t0 = std::chrono::steady_clock::now(); }
double be_timer_end_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - g_t0).count(); }
void be_run_levelpop(long long n) { for (long long i = 0; i < n; i++) artis::body_levelpop(g_ctx, i); }
void be_run_bf(long long n) { for (long long i = 0; i < n; i++) artis::body_bf(g_ctx, i); }
void be_run_ff_zsum(long long n, double *out) { for (long long i = 0; i < n; i++) artis::body_ff_zsum(g_ctx, i, out); }
void be_release() {}
size_t be_free_bytes() { return (size_t)1 << 46; }
long long be_run_bf_live_count(long long ncells, long long *off) {
  long long run = 0;
  for (long long tc = 0; tc < ncells; tc++) {
    off[tc] = run;
    for (int i = 0; i < g_ctx.st.nbfcontinua; i++) run += artis::body_bf_live_test(g_ctx, (int)tc, i) ? 1 : 0;
  }
  off[ncells] = run;
  return run;
}
void be_run_bf_live_fill(long long ncells, const long long *off, double *pair, artis::ContRec *rec, int *rank) {
  const int nbf = g_ctx.st.nbfcontinua;
  for (long long tc = 0; tc < ncells; tc++) {
    long long at = off[tc];
    for (int i = 0; i < nbf; i++) {
      rank[(size_t)tc * (nbf + 1) + i] = (int)(at - off[tc]);
      if (artis::body_bf_live_test(g_ctx, (int)tc, i)) artis::body_bf_live_store(g_ctx, (int)tc, i, at++, pair, rec);
    }
    rank[(size_t)tc * (nbf + 1) + nbf] = (int)(at - off[tc]);
  }
}
void be_run_corrphoto(long long n) { for (long long i = 0; i < n; i++) artis::body_corrphoto(g_ctx, i); }
void be_run_ma(long long n) { for (long long i = 0; i < n; i++) artis::body_ma(g_ctx, i); }
void be_run_cooling_rates(long long ncells, int nions, double *contrib, double *total) {
  for (long long i = 0; i < ncells * nions; i++) artis::body_cooling_rates_ion(g_ctx, i, contrib);
  for (long long i = 0; i < ncells; i++) artis::body_cooling_rates_total(g_ctx, i, contrib, total);
}
void be_run_cool(long long n) { for (long long i = 0; i < n; i++) artis::body_cool(g_ctx, i); }
using artis::PktCold;
using artis::PktHot;
void be_run_unpack(const ArtisPacket *aos, PktHot *hot, PktCold *cold, long long n) {
  for (long long i = 0; i < n; i++) artis::pkt_split(aos[i], hot[i], cold[i]);
}
void be_run_pack(const PktHot *hot, const PktCold *cold, ArtisPacket *aos, long long n) {
  for (long long i = 0; i < n; i++) artis::pkt_join(hot[i], cold[i], aos[i]);
}
void be_run_bin_escaped(const artis::BinSpec &b, const PktHot *hot, const PktCold *cold, long long n) {
  for (long long i = 0; i < n; i++) artis::body_bin_escaped(b, hot, cold, i, b.lc);
}
void be_run_bin_resolved(const artis::BinSpec &b, const artis::ResSpec &r, const PktHot *hot, const PktCold *cold, long long n, int *error_flag) {
  for (long long i = 0; i < n; i++) artis::body_bin_resolved(b, r, hot, cold, i, error_flag);
}
void be_run_normalise(const artis::NormSpec &n) {
  for (int i = 0; i < n.npts_model; i++) artis::body_normalise(n, i);
}
void be_run_escaped_count(const PktHot *hot, long long n, const unsigned char *reported, int only_new, int *counts) {
  const int nchunks = (int)((n + artis::ESC_CHUNK - 1) / artis::ESC_CHUNK);
  int run = 0;
  for (int c = 0; c < nchunks; c++) {
    counts[c] = run;
    for (long long i = (long long)c * artis::ESC_CHUNK; i < n && i < (long long)(c + 1) * artis::ESC_CHUNK; i++)
      run += artis::escaped_selected(hot, reported, only_new, i);
  }
  counts[nchunks] = run;
}
void be_run_escaped_scatter(const PktHot *hot, const PktCold *cold, long long n, unsigned char *reported, int only_new,
                            const int *offsets, ArtisPacket *out) {
  (void)offsets;
  int k = 0;
  for (long long i = 0; i < n; i++) {
    if (!artis::escaped_selected(hot, reported, only_new, i)) continue;
    artis::pkt_join(hot[i], cold[i], out[k++]);
    if (only_new) reported[i] = 1;
  }
}
void be_run_bucket(PktHot *hot, PktCold *cold, int n, int *buckets, int *kind_sizes, int *const *queues, bool fresh) {
  for (int i = 0; i < n; i++) artis::body_bucket_hist(g_ctx, hot, cold, i, buckets, fresh);
  artis::body_bucket_scan(g_ctx, buckets, kind_sizes);
  for (int i = 0; i < n; i++) artis::body_bucket_scatter(g_ctx, hot, i, buckets, queues);
}
void be_run_wavefront(PktHot *hot, PktCold *cold, const artis::QueueView &qv, int kind, long long maxev, int) {
  switch (kind) {  // one "thread" drains the whole queue
    case 0: artis::body_wavefront<0>(g_ctx, hot, cold, qv, maxev); break;
    case 1: artis::body_wavefront<1>(g_ctx, hot, cold, qv, maxev); break;
    case 2: artis::body_wavefront<2>(g_ctx, hot, cold, qv, maxev); break;
    case 4: artis::body_wavefront<4>(g_ctx, hot, cold, qv, maxev); break;
    case 5: artis::body_wavefront<5>(g_ctx, hot, cold, qv, maxev); break;
    default: artis::body_wavefront<3>(g_ctx, hot, cold, qv, maxev); break;
  }
}
// the barrier-free scheduler is a device-only construct; the host backend runs the same visits as kernel rounds
bool be_run_persistent(PktHot *, PktCold *, int *, int, unsigned long long *, unsigned long long *, int, long long) { return false; }
int be_persistent_kernels() { return 0; }
void be_run_vpkt_trace(const artis::VpktReq *, long long, unsigned long long *) {}  // the host build traces virtual packets where their events happen
bool be_vpkt_deferred() { return false; }
int be_scheduler_geometry(int *) { return 0; }
long long be_resident_lanes() { return 1; }
void be_fill_bytes(void *p, int byte, size_t n) { memset(p, byte, n); }
void be_set_estimator_block(double *, size_t) {}
void be_set_scheduler_kernels(int) {}
void be_set_idle_timeout_ms(long long) {}
double be_last_allreduce_ms() { return 0.; }
void be_run_probe_boundary(const ArtisPacket *p, int n, double *a, int *b, int *c) {
  for (int i = 0; i < n; i++) artis::body_probe_boundary(g_ctx, p, i, a, b, c);
}
void be_run_probe_kappa(const ArtisPacket *p, int n, double *k) { for (int i = 0; i < n; i++) artis::body_probe_kappa(g_ctx, p, i, k); }
void be_run_probe_closest(const double *nu, const int *nt, int n, int *o) {
  for (int i = 0; i < n; i++) artis::body_probe_closest(g_ctx, nu, nt, i, o);
}
void be_run_probe_get_event(const ArtisPacket *p, int n, const double *t, const double *ad, double *e, int *et, int *nt, int *ma) {
  for (int i = 0; i < n; i++) artis::body_probe_get_event(g_ctx, p, i, t, ad, e, et, nt, ma);
}
void be_run_probe_ma_rates(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  for (int i = 0; i < n; i++) artis::body_probe_ma_rates(g_ctx, a, b, c, d, i, o);
}
void be_run_probe_cooling(const int *a, const int *b, const int *c, int n, double *o) {
  for (int i = 0; i < n; i++) artis::body_probe_cooling(g_ctx, a, b, c, i, o);
}
void be_run_probe_gamma_kappa(const ArtisPacket *p, int n, double *o) {
  for (int i = 0; i < n; i++) artis::body_probe_gamma_kappa(g_ctx, p, i, o);
}
void be_run_probe_scnu(const int *a, const int *b, const int *c2, const int *d, const float *T, const double *z, int n, double *o) {
  for (int i = 0; i < n; i++) artis::body_probe_scnu(g_ctx, a, b, c2, d, T, z, i, o);
}
void be_run_probe_levelpop(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  for (int i = 0; i < n; i++) artis::body_probe_levelpop(g_ctx, a, b, c, d, i, o);
}
int be_comm_unique_id(char *, std::string *) { return ARTIS_ERR_NCCL; }
int be_comm_init(const char *, int, int, std::string *) { return ARTIS_ERR_NCCL; }
int be_comm_allreduce(double *, size_t, long long *, size_t, int *, size_t, std::string *) { return ARTIS_OK; }
}  // namespace

#include "../../artis_b200/csrc/engine.inc"
