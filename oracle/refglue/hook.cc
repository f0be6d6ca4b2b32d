// hook.cc -- stands at the reference's update_packets call site (sn3d.cc:574). TEST INFRASTRUCTURE.
//
// update_packets.cc is compiled with its entry point renamed to ref_update_packets (wrap_update_packets.cc), so the
// unmodified sn3d.cc calls THIS update_packets. Controlled by environment variables:
//   ARTIS_HOOK_SNAPSHOT_TS=a,b,c   dump <dir>/snap_ts<N>.bin (static tables once per file, cell state, packets
//                                  before and after the reference transport, the reference's estimators)
//   ARTIS_HOOK_DIR=<dir>           where to write (default ".")
//   ARTIS_HOOK_STOP_AFTER_TS=N     exit(0) after timestep N's transport (bounds the run)
//   ARTIS_HOOK_SKIP_POST=1         with snapshots: do not store the (large) post-transport packet array
// Always appends one line per timestep to <dir>/hook_timing.txt:
//   nts npkts seconds interactions cellcrossings escounter nesc threads
// This file never runs on the product path; the product is artis_b200/csrc behind include/artis_gpu.h.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <string>

#include "flatten.h"
#include "packet.h"
#include "update_packets.h"
#ifdef _OPENMP
#include <omp.h>
#endif

static_assert(sizeof(struct packet) == sizeof(ArtisPacket), "ArtisPacket must mirror struct packet (packet.h:28-73)");
static_assert(offsetof(struct packet, prop_time) == offsetof(ArtisPacket, prop_time), "layout");
static_assert(offsetof(struct packet, stokes) == offsetof(ArtisPacket, stokes), "layout");
static_assert(offsetof(struct packet, tdecay) == offsetof(ArtisPacket, tdecay), "layout");
static_assert(offsetof(struct packet, number) == offsetof(ArtisPacket, number), "layout");
static_assert(offsetof(struct packet, mastate) == offsetof(ArtisPacket, ma_element), "layout");
static_assert(offsetof(struct packet, trueemissionvelocity) == offsetof(ArtisPacket, trueemissionvelocity), "layout");

namespace {

template <typename T> const char *dtype_code();
template <> const char *dtype_code<int>() { return "i4"; }
template <> const char *dtype_code<float>() { return "f4"; }
template <> const char *dtype_code<double>() { return "f8"; }
template <> const char *dtype_code<long long>() { return "i8"; }

void write_record(FILE *f, const std::string &name, const char *dtype, const void *data, size_t count, size_t elsize) {
  char namebuf[64] = {0};
  char dt[8] = {0};
  strncpy(namebuf, name.c_str(), 63);
  strncpy(dt, dtype, 7);
  const unsigned long long n = count;
  fwrite(namebuf, 1, 64, f);
  fwrite(dt, 1, 8, f);
  fwrite(&n, sizeof(n), 1, f);
  if (count > 0) fwrite(data, elsize, count, f);
}

#define SNAP_S(type, name) write_record(f, prefix + #name, dtype_code<type>(), &s->name, 1, sizeof(type));
#define SNAP_A(type, name, count) \
  write_record(f, prefix + #name, dtype_code<type>(), s->name, (size_t)(count), sizeof(type));

void write_static(FILE *f, const ArtisStaticTables *s) {
  const std::string prefix = "static.";
  ARTIS_STATIC_FIELDS(SNAP_S, SNAP_A)
}
void write_cell(FILE *f, const ArtisCellState *s) {
  const std::string prefix = "cell.";
  ARTIS_CELL_FIELDS(SNAP_S, SNAP_A)
}
void write_estim(FILE *f, const ArtisEstimators *s) {
  const std::string prefix = "estim.";
  ARTIS_ESTIM_FIELDS(SNAP_S, SNAP_A)
}

std::set<int> parse_list(const char *env) {
  std::set<int> out;
  if (!env) return out;
  const char *p = env;
  while (*p) {
    char *end;
    const long v = strtol(p, &end, 10);
    if (end == p) break;
    out.insert((int)v);
    p = (*end == ',') ? end + 1 : end;
  }
  return out;
}

}  // namespace

void update_packets(const int my_rank, const int nts, struct packet *packets) {
  static const std::set<int> snap_ts = parse_list(getenv("ARTIS_HOOK_SNAPSHOT_TS"));
  static const std::string dir = getenv("ARTIS_HOOK_DIR") ? getenv("ARTIS_HOOK_DIR") : ".";
  static const int stop_after = getenv("ARTIS_HOOK_STOP_AFTER_TS") ? atoi(getenv("ARTIS_HOOK_STOP_AFTER_TS")) : -1;
  static const bool skip_post = getenv("ARTIS_HOOK_SKIP_POST") != nullptr;
  const bool snap = snap_ts.count(nts) > 0;
  const int npkts = globals::npkts;

  FILE *f = nullptr;
  if (snap) {
    char fname[512];
    snprintf(fname, sizeof(fname), "%s/snap_ts%d.bin", dir.c_str(), nts);
    f = fopen(fname, "wb");
    assert_always(f != nullptr);
    fwrite("ARTSNAP1", 1, 8, f);
    refglue::Storage st;
    ArtisStaticTables T;
    refglue::flatten_static(&T, &st);
    write_static(f, &T);
    ArtisCellState C;
    refglue::flatten_cellstate(&C, &st, nts);
    write_cell(f, &C);
    const int meta[4] = {my_rank, globals::nprocs, nts, npkts};
    write_record(f, "meta.rank_nprocs_nts_npkts", "i4", meta, 4, sizeof(int));
    write_record(f, "packets.pre", "pkt", packets, npkts, sizeof(struct packet));
  }

  const long long cross0 = stats::get_counter(stats::COUNTER_CELLCROSSINGS);
  const long long es0 = stats::get_counter(stats::COUNTER_ESCOUNTER);
  const auto t0 = std::chrono::steady_clock::now();
  ref_update_packets(my_rank, nts, packets);
  const auto t1 = std::chrono::steady_clock::now();
  const double seconds = std::chrono::duration<double>(t1 - t0).count();

  long long interactions = 0;
  for (int n = 0; n < npkts; n++) interactions += packets[n].interactions;
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  {
    const std::string tn = dir + "/hook_timing.txt";
    FILE *tf = fopen(tn.c_str(), "a");
    if (tf) {
      fprintf(tf, "%d %d %.6f %lld %lld %lld %d %d\n", nts, npkts, seconds, interactions,
              (long long)stats::get_counter(stats::COUNTER_CELLCROSSINGS) - cross0,
              (long long)stats::get_counter(stats::COUNTER_ESCOUNTER) - es0, globals::nesc, nthreads);
      fclose(tf);
    }
  }

  if (snap) {
    refglue::Storage st;
    ArtisEstimators E;
    refglue::flatten_estimators(&E, &st, nts);
    write_estim(f, &E);
    if (!skip_post) write_record(f, "packets.post", "pkt", packets, npkts, sizeof(struct packet));
    const double secs[1] = {seconds};
    write_record(f, "meta.ref_seconds", "f8", secs, 1, sizeof(double));
    fclose(f);
  }
  if (stop_after >= 0 && nts >= stop_after) {
    printout("hook: stopping after timestep %d as requested\n", nts);
    exit(0);
  }
}
