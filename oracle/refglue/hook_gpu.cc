// hook_gpu.cc -- the drop-in, compiled: the reference driver with its update_packets call site (sn3d.cc:574) served by
// libartis_gpu.so through the C ABI of include/artis_gpu.h. TEST INFRASTRUCTURE and integration example (INTEGRATION.md
// 1.2 quotes this file): linked with the UNMODIFIED reference objects into oracle/_ref/<mode>/sn3d_gpu by oracle/Makefile.
//
// Per timestep, at the place of update_packets(my_rank, nts, packets):
//   first call   flatten the reference's globals into ArtisStaticTables (flatten.h) -> artis_gpu_init
//   every call   flatten the cell state update_grid has just produced -> artis_gpu_upload_cellstate;
//                artis_gpu_zero_estimators; artis_gpu_update_packets on the driver's own packet array;
//                artis_gpu_download_estimators -> ADDED to the reference's estimator globals (which the driver zeroed at
//                sn3d.cc:568), so that everything downstream -- normalisation, update_grid of the next timestep, the
//                output writers -- runs on them exactly as after the host transport.
// Option sets served here: those whose estimators are all process globals or reachable through refglue.h accessors
// (classic, kilonova LTE); the multi-bin radiation field and detailed bf-rate estimators of the nebular mode would need
// two more write accessors in wrap_radfield.cc.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "flatten.h"
#include "packet.h"
#include "stats.h"
#include "update_packets.h"

static_assert(sizeof(struct packet) == sizeof(ArtisPacket), "ArtisPacket must mirror struct packet (packet.h:28-73)");
static_assert(!MULTIBIN_RADFIELD_MODEL_ON && !DETAILED_BF_ESTIMATORS_ON, "see the header comment: write accessors missing");

namespace {
void check(int rc, const char *what) {
  if (rc == ARTIS_OK) return;
  char buf[1024];
  artis_gpu_last_error(buf, sizeof(buf));
  printout("[fatal] %s failed with code %d: %s\n", what, rc, buf);  // the reference's own failure convention (sn3d.h:50-62)
  fprintf(stderr, "[fatal] %s failed with code %d: %s\n", what, rc, buf);
  abort();
}
unsigned long long seed_from_input() {  // pre_zseed, the first entry of input.txt (input.cc:1877-1893)
  if (const char *e = getenv("ARTIS_GPU_SEED")) return strtoull(e, nullptr, 10);
  FILE *f = fopen("input.txt", "r");
  long long v = 1281360349;
  if (f) {
    char line[512];
    while (fgets(line, sizeof(line), f)) {
      if (line[0] == '#' || line[0] == '\n') continue;
      v = atoll(line);
      break;
    }
    fclose(f);
  }
  return v > 0 ? (unsigned long long)v : 1281360349ull;
}
}  // namespace

void update_packets(const int my_rank, const int nts, struct packet *packets) {
  static bool inited = false;
  static refglue::Storage st_static;
  static const int stop_after = getenv("ARTIS_HOOK_STOP_AFTER_TS") ? atoi(getenv("ARTIS_HOOK_STOP_AFTER_TS")) : -1;
  const int npkts = globals::npkts;
  const auto t0 = std::chrono::steady_clock::now();
  if (!inited) {
    ArtisStaticTables T;
    refglue::flatten_static(&T, &st_static);
    const int device = getenv("ARTIS_GPU_DEVICE") ? atoi(getenv("ARTIS_GPU_DEVICE")) : 0;
    check(artis_gpu_init(&T, device, my_rank, globals::nprocs, seed_from_input() + 13ull * (unsigned long long)my_rank), "artis_gpu_init");
    inited = true;
  }
  refglue::Storage st;
  ArtisCellState C;
  refglue::flatten_cellstate(&C, &st, nts);
  check(artis_gpu_upload_cellstate(&C), "artis_gpu_upload_cellstate");
  check(artis_gpu_zero_estimators(), "artis_gpu_zero_estimators");
  check(artis_gpu_update_packets(my_rank, nts, reinterpret_cast<ArtisPacket *>(packets), npkts), "artis_gpu_update_packets");
  check(artis_gpu_reduce_estimators(nts), "artis_gpu_reduce_estimators");  // no-op for one shard

  // device estimators -> the reference's globals
  ArtisEstimators E;
  memset(&E, 0, sizeof(E));
  const int NM = grid::get_npts_model() + 1, nel = get_nelements(), mi = get_max_nions();
  const size_t NG = (size_t)NM * nel * mi;
  std::vector<double> J(NM), nuJ(NM), ff(NM), col(NM), re(NM), ga(NG), bf(NG), ts(16);
  std::vector<long long> ctr(ARTIS_CTR_COUNT);
  std::vector<int> ec(globals::nlines > 0 ? globals::nlines : 1), ac(ec.size());
  E.npts_model = NM - 1;
  E.nelements = nel;
  E.max_nions = mi;
  E.nlines = globals::nlines;
  E.J = J.data();
  E.nuJ = nuJ.data();
  E.ffheatingestimator = ff.data();
  E.colheatingestimator = col.data();
  E.rpkt_emiss = re.data();
  E.gammaestimator = ga.data();
  E.bfheatingestimator = bf.data();
  E.ts_scalars = ts.data();
  E.counters = ctr.data();
  E.ecounter = ec.data();
  E.acounter = ac.data();
  check(artis_gpu_download_estimators(&E), "artis_gpu_download_estimators");
  for (int m = 0; m < NM - 1; m++) {
    refglue::radfield_J()[m] += J[m];
    refglue::radfield_nuJ()[m] += nuJ[m];
    globals::ffheatingestimator[m] += ff[m];
    globals::colheatingestimator[m] += col[m];
    globals::rpkt_emiss[m] += re[m];
  }
#if (!NO_LUT_PHOTOION)
  for (size_t i = 0; i < (size_t)(NM - 1) * nel * mi; i++) globals::gammaestimator[i] += ga[i];
#endif
#if (!NO_LUT_BFHEATING)
  for (size_t i = 0; i < (size_t)(NM - 1) * nel * mi; i++) globals::bfheatingestimator[i] += bf[i];
#endif
  struct time &T = globals::time_step[nts];
  T.gamma_dep += ts[ARTIS_TS_GAMMA_DEP];
  T.positron_dep += ts[ARTIS_TS_POSITRON_DEP];
  T.electron_dep += ts[ARTIS_TS_ELECTRON_DEP];
  T.electron_emission += ts[ARTIS_TS_ELECTRON_EMISSION];
  T.alpha_dep += ts[ARTIS_TS_ALPHA_DEP];
  T.alpha_emission += ts[ARTIS_TS_ALPHA_EMISSION];
  T.gamma_emission += ts[ARTIS_TS_GAMMA_EMISSION];
  T.cmf_lum += ts[ARTIS_TS_CMF_LUM];
  T.pellet_decays += (int)ctr[ARTIS_CTR_PELLET_DECAYS];
  nonthermal::nt_energy_deposited += ts[ARTIS_TS_NT_ENERGY_DEPOSITED];
  globals::nesc += (int)ctr[ARTIS_CTR_NESC];
  for (int i = 0; i < stats::COUNTER_COUNT; i++)  // stats has no add: the log lines only need the totals to be right
    for (long long k = 0; k < ctr[i] && i != stats::COUNTER_NESC; k++) stats::increment((enum stats::eventcounters)i);
#ifdef RECORD_LINESTAT
  for (int i = 0; i < globals::nlines; i++) {
    globals::ecounter[i] += ec[i];
    globals::acounter[i] += ac[i];
  }
#endif
  const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  long long interactions = 0;
  for (int n = 0; n < npkts; n++) interactions += packets[n].interactions;
  printout("artis_gpu: timestep %d, %d packets, %lld interactions, %.3f s in the C-ABI calls\n", nts, npkts, interactions, seconds);
  if (FILE *tf = fopen("hook_timing.txt", "a")) {
    fprintf(tf, "%d %d %.6f %lld gpu\n", nts, npkts, seconds, interactions);
    fclose(tf);
  }
  if (stop_after >= 0 && nts >= stop_after) {
    printout("hook: stopping after timestep %d as requested\n", nts);
    artis_gpu_finalize();
    exit(0);
  }
}
