// hook.cc -- stands at the reference's update_packets call site (sn3d.cc:574). TEST INFRASTRUCTURE.
//
// update_packets.cc is compiled with its entry point renamed to ref_update_packets (wrap_update_packets.cc), so the
// unmodified sn3d.cc calls THIS update_packets. Controlled by environment variables:
//   ARTIS_HOOK_SNAPSHOT_TS=a,b,c   dump <dir>/snap_ts<N>.bin (static tables once per file, cell state, packets
//                                  before and after the reference transport, the reference's estimators)
//   ARTIS_HOOK_DIR=<dir>           where to write (default ".")
//   ARTIS_HOOK_STOP_AFTER_TS=N     exit(0) after timestep N's transport (bounds the run)
//   ARTIS_HOOK_PROBES=N            with snapshots: also store N deterministic probes (inputs + reference outputs)
//   ARTIS_HOOK_SPECTRA=1           with snapshots: also store the light curves and the spectrum that the reference's
//                                  add_to_lc_res / add_to_spec_res make of the post-transport packets (spectra.*)
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
#include "boundary.h"
#include "ltepop.h"
#include "photo_electric.h"
#include "grey_emissivities.h"
#include "radfield.h"
#include "ratecoeff.h"
#include "rpkt.h"
#include "update_grid.h"
#include "vectors.h"
#include "exspec.h"
#include "light_curve.h"
#include "spectrum.h"
#include <vector>
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


// ---- deterministic probes: the reference's own functions called on packet states built
// from packets.pre; inputs and outputs go into the snapshot so that the CUDA probes (artis_gpu_probe_*) can be
// compared entry by entry. Runs before the reference transport; touches only the per-thread caches.
extern "C" void gslshim_force_uniform(double u);  // oracle/gslshim/gslshim.cc (test hook of the GSL stand-in)
double hash01(unsigned long long x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

void write_probes(FILE *f, const struct packet *packets, const int npkts, const int nts, const int nprobe_max) {
  std::vector<struct packet> q;
  std::vector<struct packet> gq;
  const double t2 = globals::time_step[nts].start + globals::time_step[nts].width;
  const int stride = (nprobe_max > 0 && npkts > nprobe_max) ? (npkts + nprobe_max - 1) / nprobe_max : 1;  // sample every cell
  for (int n = 0; n < npkts && (int)q.size() < nprobe_max; n += stride) {
    struct packet p = packets[n];
    if (p.type == TYPE_ESCAPE || !(p.prop_time < t2)) continue;
    const unsigned long long h = 1000003ULL * (unsigned long long)p.number + 17ULL * nts;
    if (p.type != TYPE_RPKT) {
      // pellet / gamma / k-packet position re-used for a synthetic r-packet
      const double mu = -1 + 2 * hash01(h + 1), phi = 2 * 3.141592653589793 * hash01(h + 2);
      const double st = sqrt(1 - mu * mu);
      p.dir[0] = st * cos(phi);
      p.dir[1] = st * sin(phi);
      p.dir[2] = mu;
      p.nu_rf = globals::nu_min_r * pow(globals::nu_max_r / globals::nu_min_r, hash01(h + 3));
      p.e_rf = p.e_cmf;
      p.type = TYPE_RPKT;
      p.last_cross = NONE;
      p.next_trans = 0;
      const double dopplerfactor = doppler_packet_nucmf_on_nurf(&p);
      p.nu_cmf = p.nu_rf * dopplerfactor;
      p.e_cmf = p.e_rf * dopplerfactor;
      // the same position as a gamma packet of 0.05 - 3 MeV
      struct packet g = p;
      g.type = TYPE_GAMMA;
      g.nu_rf = (0.05 * pow(60., hash01(h + 4))) * MEV / H;
      g.nu_cmf = g.nu_rf * dopplerfactor;
      gq.push_back(g);
    }
    q.push_back(p);
  }
  const int N = q.size();
  std::vector<double> sdist(N), kappa(5 * (size_t)N, 0.), tau(N), abortd(N), edist(N, 0.);
  std::vector<int> snext(N), lcross(N), closest(N), closest0(N), valid(N, 0), etype(N, -1), ntout(N, 0), ma(4 * (size_t)N, 0);
  for (int i = 0; i < N; i++) {
    struct packet p = q[i];
    int sn = 0;
    sdist[i] = boundary_cross(&p, &sn);
    snext[i] = sn;
    lcross[i] = p.last_cross;
    closest[i] = closest_transition(q[i].nu_cmf, q[i].next_trans);
    closest0[i] = closest_transition(q[i].nu_cmf, 0);
    const int mgi = grid::get_cell_modelgridindex(q[i].where);
    tau[i] = -log(1. - hash01(7919ULL * i + 3));
    const double tdist = (t2 - q[i].prop_time) * CLIGHT_PROP;
    abortd[i] = (sdist[i] > 0 && sdist[i] < tdist) ? sdist[i] : tdist;
    if (mgi != grid::get_npts_model() && grid::modelgrid[mgi].thick != 1) {
      valid[i] = 1;
      cellhistory_reset(mgi, false);
      globals::kappa_rpkt_cont[tid].recalculate_required = true;
      struct packet p2 = q[i];
      calculate_kappa_rpkt_cont(&p2, &globals::kappa_rpkt_cont[tid]);
      kappa[5 * i + 0] = globals::kappa_rpkt_cont[tid].total;
      kappa[5 * i + 1] = globals::kappa_rpkt_cont[tid].es;
      kappa[5 * i + 2] = globals::kappa_rpkt_cont[tid].ff;
      kappa[5 * i + 3] = globals::kappa_rpkt_cont[tid].bf;
      kappa[5 * i + 4] = globals::kappa_rpkt_cont[tid].ffheating;
      struct packet p3 = q[i];
      int et = -1;
      edist[i] = refglue::probe_get_event(mgi, &p3, &et, tau[i], abortd[i]);
      etype[i] = (et == RPKT_EVENTTYPE_BB) ? 1 : (et == RPKT_EVENTTYPE_CONT ? 2 : -1);
      ntout[i] = p3.next_trans;
      if (etype[i] == 1) {
        ma[4 * i + 0] = p3.mastate.element;
        ma[4 * i + 1] = p3.mastate.ion;
        ma[4 * i + 2] = p3.mastate.level;
        ma[4 * i + 3] = p3.mastate.activatingline;
      }
    }
  }
  write_record(f, "probe.packets", "pkt", q.data(), N, sizeof(struct packet));
  write_record(f, "probe.sdist", "f8", sdist.data(), N, 8);
  write_record(f, "probe.snext", "i4", snext.data(), N, 4);
  write_record(f, "probe.last_cross", "i4", lcross.data(), N, 4);
  write_record(f, "probe.closest", "i4", closest.data(), N, 4);
  write_record(f, "probe.closest0", "i4", closest0.data(), N, 4);
  write_record(f, "probe.valid", "i4", valid.data(), N, 4);
  write_record(f, "probe.kappa", "f8", kappa.data(), 5 * (size_t)N, 8);
  write_record(f, "probe.tau_rnd", "f8", tau.data(), N, 8);
  write_record(f, "probe.abort_dist", "f8", abortd.data(), N, 8);
  write_record(f, "probe.edist", "f8", edist.data(), N, 8);
  write_record(f, "probe.eventtype", "i4", etype.data(), N, 4);
  write_record(f, "probe.next_trans_out", "i4", ntout.data(), N, 4);
  write_record(f, "probe.mastate", "i4", ma.data(), 4 * (size_t)N, 4);
  // gamma opacities
  const int G = gq.size();
  std::vector<double> gk(3 * (size_t)G);
  for (int i = 0; i < G; i++) {
    gk[3 * i + 0] = refglue::probe_sig_comp(&gq[i]);
    gk[3 * i + 1] = sig_photo_electric(&gq[i]);
    gk[3 * i + 2] = sig_pair_prod(&gq[i]);
  }
  write_record(f, "probe.gpackets", "pkt", gq.data(), G, sizeof(struct packet));
  write_record(f, "probe.gkappa", "f8", gk.data(), 3 * (size_t)G, 8);
  // level populations, macro-atom process rates and per-ion cooling sums for every level / ion of a few cells
  std::vector<int> cells;
  for (int m = 0; m < grid::get_npts_model() && cells.size() < 4; m++)
    if (grid::get_numassociatedcells(m) > 0 && grid::modelgrid[m].thick != 1) cells.push_back(m);
  {
    int last = -1;
    for (int m = grid::get_npts_model() - 1; m >= 0; m--)
      if (grid::get_numassociatedcells(m) > 0 && grid::modelgrid[m].thick != 1) {
        last = m;
        break;
      }
    if (last >= 0 && (cells.empty() || cells.back() != last)) cells.push_back(last);
  }
  std::vector<int> lm, le, li, ll, cm, ce, ci;
  std::vector<double> pop, rates, cool;
  const double t_mid = globals::time_step[nts].mid;
  for (int m : cells) {
    cellhistory_reset(m, true);
    for (int el = 0; el < get_nelements(); el++)
      for (int ion = 0; ion < get_nions(el); ion++) {
        cm.push_back(m);
        ce.push_back(el);
        ci.push_back(ion);
        cool.push_back(refglue::probe_cooling_ion(m, el, ion));
        for (int lev = 0; lev < get_nlevels(el, ion); lev++) {
          lm.push_back(m);
          le.push_back(el);
          li.push_back(ion);
          ll.push_back(lev);
          pop.push_back(get_levelpop(m, el, ion, lev));
          double r9[9];
          refglue::probe_macroatom_rates(m, el, ion, lev, t_mid, r9);
          rates.insert(rates.end(), r9, r9 + 9);
        }
      }
  }
  write_record(f, "probe.lev_mgi", "i4", lm.data(), lm.size(), 4);
  write_record(f, "probe.lev_element", "i4", le.data(), le.size(), 4);
  write_record(f, "probe.lev_ion", "i4", li.data(), li.size(), 4);
  write_record(f, "probe.lev_level", "i4", ll.data(), ll.size(), 4);
  write_record(f, "probe.levelpop", "f8", pop.data(), pop.size(), 8);
  write_record(f, "probe.ma_rates", "f8", rates.data(), rates.size(), 8);
  write_record(f, "probe.cool_mgi", "i4", cm.data(), cm.size(), 4);
  write_record(f, "probe.cool_element", "i4", ce.data(), ce.size(), 4);
  write_record(f, "probe.cool_ion", "i4", ci.data(), ci.size(), 4);
  write_record(f, "probe.cool_total", "f8", cool.data(), cool.size(), 8);
  // select_continuum_nu (ratecoeff.cc:628-684) with its random number injected through the GSL stand-in: a sample of
  // bound-free continua x the T_e of the probed cells x a spread of draws (the function returns the frequency at which
  // the tail integral of the emission profile equals zrand = 1 - u of the total)
  {
    std::vector<int> se, si, sl, su;
    std::vector<float> sT;
    std::vector<double> sz, snu;
    const double draws[] = {0.0, 0.004, 0.05, 0.3, 0.5, 0.75, 0.93, 0.995};
    const int nbf = globals::nbfcontinua;
    const int stride_c = nbf > 160 ? nbf / 160 : 1;
    for (int i = 0; i < nbf; i += stride_c) {
      const int el = globals::allcont[i].element, ion = globals::allcont[i].ion, lev = globals::allcont[i].level;
      const int upper = get_phixsupperlevel(el, ion, lev, globals::allcont[i].phixstargetindex);
      for (size_t ci2 = 0; ci2 < cells.size(); ci2 += (cells.size() > 2 ? cells.size() - 1 : 1)) {
        const float T_e = grid::get_Te(cells[ci2]);
        for (double u : draws) {
          gslshim_force_uniform(u);
          const double nu = select_continuum_nu(el, ion, lev, upper, T_e);
          se.push_back(el);
          si.push_back(ion);
          sl.push_back(lev);
          su.push_back(upper);
          sT.push_back(T_e);
          sz.push_back(1. - u);
          snu.push_back(nu);
        }
      }
    }
    write_record(f, "probe.scnu_element", "i4", se.data(), se.size(), 4);
    write_record(f, "probe.scnu_ion", "i4", si.data(), si.size(), 4);
    write_record(f, "probe.scnu_level", "i4", sl.data(), sl.size(), 4);
    write_record(f, "probe.scnu_upper", "i4", su.data(), su.size(), 4);
    write_record(f, "probe.scnu_Te", "f4", sT.data(), sT.size(), 4);
    write_record(f, "probe.scnu_zrand", "f8", sz.data(), sz.size(), 8);
    write_record(f, "probe.scnu_nu", "f8", snu.data(), snu.size(), 8);
  }
  // leave the per-thread caches as update_packets expects to find them
  globals::cellhistory[tid].cellnumber = -99;
  globals::kappa_rpkt_cont[tid].recalculate_required = true;
}

// the packet loop of write_partial_lightcurve_spectra (spectrum.cc:672-682) with the reference's own callees, plus the
// Stokes spectra that exspec passes to the same function
void write_spectra(FILE *f, const struct packet *pkts, const int npkts) {
  globals::nnubins = MNUBINS;  // spectrum.cc:649
  const int ntstep = globals::ntstep, nnubins = globals::nnubins;
  struct spec *sp[4];
  for (auto &x : sp) {
    x = alloc_spectra(false);
    init_spectra(x, globals::nu_min_r, globals::nu_max_r, false);
  }
  std::vector<double> lc(4 * (size_t)ntstep, 0.);
  for (int ii = 0; ii < npkts; ii++) {
    if (pkts[ii].type != TYPE_ESCAPE) continue;
    if (pkts[ii].escape_type == TYPE_RPKT) {
      add_to_lc_res(&pkts[ii], -1, &lc[0], &lc[ntstep]);
      add_to_spec_res(&pkts[ii], -1, sp[0], sp[1], sp[2], sp[3]);
    } else if (pkts[ii].escape_type == TYPE_GAMMA) {
      add_to_lc_res(&pkts[ii], -1, &lc[2 * (size_t)ntstep], &lc[3 * (size_t)ntstep]);
    }
  }
  std::vector<double> flux(4 * (size_t)ntstep * nnubins);
  for (int k = 0; k < 4; k++)
    for (int nt = 0; nt < ntstep; nt++)
      for (int nnu = 0; nnu < nnubins; nnu++) flux[((size_t)k * ntstep + nt) * nnubins + nnu] = sp[k]->timesteps[nt].flux[nnu];
  const double params[3] = {globals::tmax, (double)globals::nprocs, (double)nnubins};
  write_record(f, "spectra.params", "f8", params, 3, 8);
  write_record(f, "spectra.lc", "f8", lc.data(), lc.size(), 8);
  write_record(f, "spectra.flux", "f8", flux.data(), (size_t)ntstep * nnubins, 8);
  write_record(f, "spectra.stokes", "f8", flux.data() + (size_t)ntstep * nnubins, 3 * (size_t)ntstep * nnubins, 8);
  write_record(f, "spectra.delta_freq", "f4", sp[0]->delta_freq, nnubins, 4);
  for (auto &x : sp) free_spectra(x);
}

// emission / absorption-resolved angle-averaged spectrum (add_to_spec with do_emission_res, spectrum.cc:339-452) and the
// angle-resolved spectra and light curves of all MABINS escape-direction bins (add_to_spec_res / add_to_lc_res with
// current_abin >= 0, spectrum.cc:602-620, light_curve.cc:55-62; vectors.h:158-192), by the reference's own functions
void write_spectra_resolved(FILE *f, const struct packet *pkts, const int npkts) {
  globals::nnubins = MNUBINS;
  const int ntstep = globals::ntstep, nnubins = globals::nnubins;
  const int proccount = 2 * get_nelements() * get_max_nions() + 1, ioncount = get_nelements() * get_max_nions();
  struct spec *sp = alloc_spectra(true);
  init_spectra(sp, globals::nu_min_r, globals::nu_max_r, true);
  std::vector<struct spec *> spa(MABINS);
  for (auto &x : spa) {
    x = alloc_spectra(false);
    init_spectra(x, globals::nu_min_r, globals::nu_max_r, false);
  }
  std::vector<double> lcres((size_t)MABINS * ntstep, 0.), lcdummy((size_t)ntstep, 0.);
  for (int ii = 0; ii < npkts; ii++) {
    if (pkts[ii].type != TYPE_ESCAPE || pkts[ii].escape_type != TYPE_RPKT) continue;
    add_to_spec_res(&pkts[ii], -1, sp, nullptr, nullptr, nullptr);
    for (int a = 0; a < MABINS; a++) {
      add_to_lc_res(&pkts[ii], a, &lcres[(size_t)a * ntstep], lcdummy.data());
      add_to_spec_res(&pkts[ii], a, spa[a], nullptr, nullptr, nullptr);
    }
  }
  std::vector<double> em((size_t)ntstep * nnubins * proccount), tem(em.size()), ab((size_t)ntstep * nnubins * ioncount);
  for (int nt = 0; nt < ntstep; nt++) {
    memcpy(&em[(size_t)nt * nnubins * proccount], sp->timesteps[nt].emission, sizeof(double) * nnubins * proccount);
    memcpy(&tem[(size_t)nt * nnubins * proccount], sp->timesteps[nt].trueemission, sizeof(double) * nnubins * proccount);
    memcpy(&ab[(size_t)nt * nnubins * ioncount], sp->timesteps[nt].absorption, sizeof(double) * nnubins * ioncount);
  }
  std::vector<double> fluxres((size_t)MABINS * ntstep * nnubins);
  for (int a = 0; a < MABINS; a++)
    for (int nt = 0; nt < ntstep; nt++)
      memcpy(&fluxres[((size_t)a * ntstep + nt) * nnubins], spa[a]->timesteps[nt].flux, sizeof(double) * nnubins);
  const double params[6] = {globals::tmax, (double)globals::nprocs, (double)nnubins, globals::syn_dir[0], globals::syn_dir[1],
                            globals::syn_dir[2]};
  write_record(f, "spectra.res_params", "f8", params, 6, 8);
  write_record(f, "spectra.res_emission", "f8", em.data(), em.size(), 8);
  write_record(f, "spectra.res_trueemission", "f8", tem.data(), tem.size(), 8);
  write_record(f, "spectra.res_absorption", "f8", ab.data(), ab.size(), 8);
  write_record(f, "spectra.res_lc", "f8", lcres.data(), lcres.size(), 8);
  write_record(f, "spectra.res_flux", "f8", fluxres.data(), fluxres.size(), 8);
  free_spectra(sp);
  for (auto &x : spa) free_spectra(x);
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
    if (getenv("ARTIS_HOOK_PROBES")) write_probes(f, packets, npkts, nts, atoi(getenv("ARTIS_HOOK_PROBES")));
  }

#ifdef VPKT_ON
  if (snap) refglue::vpkt_zero_accumulators();  // the run-long virtual-packet sums: dump this step's contribution only
#endif
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
    if (getenv("ARTIS_HOOK_SPECTRA")) write_spectra(f, packets, npkts);
    if (getenv("ARTIS_HOOK_SPECTRA_RES")) write_spectra_resolved(f, packets, npkts);
    if (getenv("ARTIS_HOOK_NORMALISE")) {
      // what update_grid makes of the raw estimators at the start of the next timestep (update_grid.cc:1041-1134) and
      // normalise_grey (grey_emissivities.cc:124-139), by the reference's own functions. Destructive: combine with
      // ARTIS_HOOK_STOP_AFTER_TS.
      const int nm = grid::get_npts_model();
      std::vector<double> vol(nm, 0.), nJ(nm, 0.), nnuJ(nm, 0.), nff(nm, 0.), ncol(nm, 0.), nre(nm, 0.), nTJ(nm, 0.);
      const double deltat = globals::time_step[nts].width;
      for (int mgi = 0; mgi < nm; mgi++) {
        vol[mgi] = grid::vol_init_modelcell(mgi);
        if (grid::get_numassociatedcells(mgi) <= 0) continue;
        const double deltaV = grid::vol_init_modelcell(mgi) * pow(globals::time_step[nts].mid / globals::tmin, 3);
        const double estimator_normfactor = 1 / deltaV / deltat / globals::nprocs;
        const double estimator_normfactor_over4pi = ONEOVER4PI * estimator_normfactor;
        radfield::normalise_J(mgi, estimator_normfactor_over4pi);
        nTJ[mgi] = radfield::get_T_J_from_J(mgi);
        radfield::normalise_nuJ(mgi, estimator_normfactor_over4pi);
        globals::ffheatingestimator[mgi] *= estimator_normfactor;
        globals::colheatingestimator[mgi] *= estimator_normfactor;
        nJ[mgi] = refglue::radfield_J()[mgi];
        nnuJ[mgi] = refglue::radfield_nuJ()[mgi];
        nff[mgi] = globals::ffheatingestimator[mgi];
        ncol[mgi] = globals::colheatingestimator[mgi];
      }
      normalise_grey(nts);
      for (int mgi = 0; mgi < nm; mgi++) nre[mgi] = globals::rpkt_emiss[mgi];
      const double scal[4] = {globals::time_step[nts].gamma_dep_pathint, (double)globals::nprocs, MINTEMP, MAXTEMP};
      write_record(f, "norm.vol_init", "f8", vol.data(), nm, 8);
      write_record(f, "norm.J", "f8", nJ.data(), nm, 8);
      write_record(f, "norm.nuJ", "f8", nnuJ.data(), nm, 8);
      write_record(f, "norm.ff", "f8", nff.data(), nm, 8);
      write_record(f, "norm.col", "f8", ncol.data(), nm, 8);
      write_record(f, "norm.rpkt_emiss", "f8", nre.data(), nm, 8);
      write_record(f, "norm.TJ", "f8", nTJ.data(), nm, 8);
      write_record(f, "norm.scalars", "f8", scal, 4, 8);
    }
    const double secs[1] = {seconds};
    write_record(f, "meta.ref_seconds", "f8", secs, 1, sizeof(double));
    fclose(f);
  }
  if (stop_after >= 0 && nts >= stop_after) {
    printout("hook: stopping after timestep %d as requested\n", nts);
    exit(0);
  }
}
