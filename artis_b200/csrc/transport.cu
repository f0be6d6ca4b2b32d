// transport.cu -- CUDA backend of the ARTIS packet-transport C ABI (include/artis_gpu.h), sm_100a.
//
// Builds libartis_gpu.so (see __graft_entry__.build()). Kernel bodies live in kernels.cuh; this file holds the
// __global__ wrappers, the launch geometry and the CUDA implementations of the backend primitives that engine.inc
// (the host side of the ABI) is written against. There is no CPU path: every entry point fails with
// ARTIS_ERR_NODEVICE when no CUDA device is present.
//
// Launch geometry: B200 has 148 SMs. Transport threads carry a whole packet (304 B) plus the continuum-opacity
// scratch in registers/local memory, so blocks are small (128 threads) and the grid is a grid-stride loop sized to a
// multiple of 148 * resident blocks per SM; the table kernels are plain one-thread-per-entry grids.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <string>

#include "kernels.cuh"
#include "spectra.cuh"

namespace {

__constant__ artis::Ctx d_ctx;

constexpr int SM_COUNT_B200 = 148;


__global__ void k_levelpop(long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_levelpop(d_ctx, i);
}
__global__ void k_bf(long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_bf(d_ctx, i);
}
__global__ void k_ff_zsum(long long n, double *out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_ff_zsum(d_ctx, i, out);
}
// live continua of every cell (rpkt.cuh bf_is_live): count per cell, offsets, ordered compaction. One block per table cell.
__global__ void __launch_bounds__(256) k_bf_live_count(long long *counts) {
  __shared__ int warp_n[8];
  const int tc = blockIdx.x;
  const int nbf = d_ctx.st.nbfcontinua;
  int n = 0;
  for (int i = threadIdx.x; i < nbf; i += 256) n += artis::body_bf_live_test(d_ctx, tc, i) ? 1 : 0;
  for (int off = 16; off >= 1; off >>= 1) n += __shfl_down_sync(0xffffffffu, n, off);
  if ((threadIdx.x & 31) == 0) warp_n[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; w++) t += warp_n[w];
    counts[tc] = t;
  }
}
__global__ void __launch_bounds__(1024) k_scan_exclusive_ll(long long *a, const int n) {  // one block; a[n] receives the total
  __shared__ long long warp_tot[32];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    const long long v = i < n ? a[i] : 0;
    long long x = v;
    for (int d = 1; d < 32; d <<= 1) {
      const long long y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    if (w == 0) {
      long long t = warp_tot[lane];
      for (int d = 1; d < 32; d <<= 1) {
        const long long y = __shfl_up_sync(0xffffffffu, t, d);
        if (lane >= d) t += y;
      }
      warp_tot[lane] = t;
    }
    __syncthreads();
    const long long before = carry + (w > 0 ? warp_tot[w - 1] : 0) + x - v;
    if (i < n) a[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) a[n] = carry;
}
__global__ void __launch_bounds__(256) k_bf_live_fill(const long long *off, double *pair, artis::ContRec *rec, int *rank) {
  __shared__ int warp_n[8];
  __shared__ long long base;
  const int tc = blockIdx.x;
  const int nbf = d_ctx.st.nbfcontinua;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) base = off[tc];
  __syncthreads();
  for (int start = 0; start < nbf; start += 256) {
    const int i = start + threadIdx.x;
    const bool live = i < nbf && artis::body_bf_live_test(d_ctx, tc, i);
    const unsigned m = __ballot_sync(0xffffffffu, live);
    if (lane == 0) warp_n[w] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
    for (int k = 0; k < 8; k++) {
      if (k < w) before += warp_n[k];
      total += warp_n[k];
    }
    const long long at = base + before + __popc(m & ((1u << lane) - 1u));
    if (live) artis::body_bf_live_store(d_ctx, tc, i, at, pair, rec);
    if (i < nbf) rank[(size_t)tc * (nbf + 1) + i] = (int)(at - off[tc]);
    __syncthreads();
    if (threadIdx.x == 0) base += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) rank[(size_t)tc * (nbf + 1) + nbf] = (int)(base - off[tc]);
}
// one warp per (cell, continuum): the lanes split the knot intervals of the photoionisation integral
__global__ void k_corrphoto(long long n) {
  using namespace artis;
  const Ctx &c = d_ctx;
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long idx = warp; idx < n; idx += nwarps) {
    const int tc = (int)(idx / c.st.nbfcontinua);
    const int i = (int)(idx % c.st.nbfcontinua);
    const int mgi = c.mgi_of_tabcell[tc];
    if (!cell_uses_tables(c, mgi)) continue;
    bool needs_sum;
    double v = build_corrphoto_part(c, mgi, i, lane, 32, &needs_sum);  // warp-uniform branch inside
    if (needs_sum) {
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
      v *= corrphoto_scale(c, i);
    }
    if (lane == 0) c.tb.corrphoto[(size_t)tc * c.st.nbfcontinua + i] = v;
  }
}
// dense macro-atom tables: one warp per (cell, level), a lane per transition (celltables.cuh coop_build_ma_into; bit for bit
// the rows of the one-thread builder body_ma, which the host build and the rows-on-demand path's single lanes use). A level
// of a 1e6-line list has thousands of transitions: one thread per level left most of the device waiting for the longest rows.
__global__ void __launch_bounds__(128) k_ma(long long n) {
  using namespace artis;
  const Ctx &c = d_ctx;
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long idx = warp; idx < n; idx += nwarps) {
    const size_t tc = (size_t)(idx / c.st.nlevels_total);
    const int gl = (int)(idx % c.st.nlevels_total);
    const int mgi = c.mgi_of_tabcell[tc];
    if (!cell_uses_tables(c, mgi)) continue;
    const int doff = c.st.lev_downoffset[gl], uoff = c.st.lev_upoffset[gl];
    coop_build_ma_into(c, mgi, (int)tc, gl, lane, c.tb.ma_rates + (tc * c.st.nlevels_total + gl) * MA_ACTION_COUNT,
                       c.tb.ma_cum_raddeexc + tc * c.st.ndowntrans_total + doff,
                       c.tb.ma_cum_intdown + tc * c.st.ndowntrans_total + doff, c.tb.ma_cum_intup + tc * c.st.nuptrans_total + uoff);
  }
}
__global__ void k_cooling_rates_ion(long long n, double *contrib) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_cooling_rates_ion(d_ctx, i, contrib);
}
__global__ void k_cooling_rates_total(long long n, const double *contrib, double *total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_cooling_rates_total(d_ctx, i, contrib, total);
}
// dense cooling tables: one warp per (cell, ion) (celltables.cuh coop_cooling_ion_into; the rows of body_cool bit for bit).
// One thread per (cell, ion) was 2000 threads for the classic model, each on a serial loop over ~2000 terms: 5.2 of the 5.5 ms
// of the per-timestep table build.
__global__ void __launch_bounds__(128) k_cool(long long n) {
  using namespace artis;
  const Ctx &c = d_ctx;
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long idx = warp; idx < n; idx += nwarps) {
    const size_t tc = (size_t)(idx / c.st.includedions);
    const int mgi = c.mgi_of_tabcell[tc];
    if (!cell_uses_tables(c, mgi)) continue;
    coop_cooling_ion_into(c, mgi, (int)(idx % c.st.includedions), lane, c.tb.cool_cum + tc * c.st.ncoolingterms,
                          c.tb.collexc_cum + tc * c.st.nuptrans_total);
  }
}
// light curves + angle-averaged spectrum of the resident packets (spectra.cuh). The 4 x ntstep light-curve sums are a few
// hundred addresses that every escaped packet of the ensemble adds to: they are accumulated per block in shared memory
// and flushed once; the ntstep x nnubins spectrum bins are spread thinly and take global FP64 reductions directly.
__global__ void __launch_bounds__(256) k_bin_escaped(const artis::BinSpec b, const artis::PktHot *hot, const artis::PktCold *cold,
                                                     long long n, int lc_in_smem) {
  extern __shared__ double lc_smem[];
  const int nlc = 4 * b.ntstep;
  double *lc_local = b.lc;
  if (lc_in_smem) {
    for (int i = threadIdx.x; i < nlc; i += blockDim.x) lc_smem[i] = 0.;
    __syncthreads();
    lc_local = lc_smem;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_bin_escaped(b, hot, cold, i, lc_local);
  if (lc_in_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < nlc; i += blockDim.x)
      if (lc_smem[i] != 0.) atomicAdd(&b.lc[i], lc_smem[i]);
  }
}
__global__ void __launch_bounds__(256) k_bin_resolved(const artis::BinSpec b, const artis::ResSpec r, const artis::PktHot *hot,
                                                      const artis::PktCold *cold, long long n, int *error_flag) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_bin_resolved(b, r, hot, cold, i, error_flag);
}
__global__ void k_normalise(const artis::NormSpec n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n.npts_model; i += gridDim.x * blockDim.x) artis::body_normalise(n, i);
}
// ABI record <-> device records (common.cuh pkt_split / pkt_join), one thread per packet. The 304-byte host records of a
// warp are consecutive, so its 16-byte loads sweep 9.5 KB of contiguous memory; the hot lines are written whole.
__global__ void __launch_bounds__(256) k_unpack(const ArtisPacket *aos, artis::PktHot *hot, artis::PktCold *cold, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    ArtisPacket a;
    const uint4 *src = reinterpret_cast<const uint4 *>(aos + i);
    uint4 *dst = reinterpret_cast<uint4 *>(&a);
#pragma unroll
    for (int w = 0; w < (int)(sizeof(ArtisPacket) / 16); w++) dst[w] = src[w];
    artis::PktHot h;
    artis::PktCold c;
    artis::pkt_split(a, h, c);
    hot[i] = h;
    cold[i] = c;
  }
}
__global__ void __launch_bounds__(256) k_pack(const artis::PktHot *hot, const artis::PktCold *cold, ArtisPacket *aos, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const artis::PktHot h = hot[i];
    const artis::PktCold c = cold[i];
    ArtisPacket a;
    artis::pkt_join(h, c, a);
    const uint4 *src = reinterpret_cast<const uint4 *>(&a);
    uint4 *dst = reinterpret_cast<uint4 *>(aos + i);
#pragma unroll
    for (int w = 0; w < (int)(sizeof(ArtisPacket) / 16); w++) dst[w] = src[w];
  }
}
// escaped-packet compaction in array order: per-chunk counts, one scan, each selected packet joined into its ABI record
__global__ void __launch_bounds__(artis::ESC_CHUNK) k_escaped_count(const artis::PktHot *hot, long long n, const unsigned char *reported,
                                                                   int only_new, int *counts) {
  const long long i = blockIdx.x * (long long)artis::ESC_CHUNK + threadIdx.x;
  const int sel = i < n && artis::escaped_selected(hot, reported, only_new, i);
  const int c = __syncthreads_count(sel);
  if (threadIdx.x == 0) counts[blockIdx.x] = c;
}
// exclusive scan of nchunks counts in place; counts[nchunks] = total. One block: the array is n/256 ints.
__global__ void __launch_bounds__(1024) k_escaped_scan(int *counts, int nchunks) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int base = 0; base < nchunks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < nchunks ? counts[i] : 0;
    int x = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    if (w == 0) {
      int t = warp_tot[lane];
      for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, t, d);
        if (lane >= d) t += y;
      }
      warp_tot[lane] = t;
    }
    __syncthreads();
    const int before = carry + (w > 0 ? warp_tot[w - 1] : 0) + x - v;
    if (i < nchunks) counts[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[nchunks] = carry;
}
__global__ void __launch_bounds__(artis::ESC_CHUNK) k_escaped_scatter(const artis::PktHot *hot, const artis::PktCold *cold, long long n,
                                                                     unsigned char *reported, int only_new, const int *offsets,
                                                                     ArtisPacket *out) {
  __shared__ int warp_base[artis::ESC_CHUNK / 32];
  const long long i = blockIdx.x * (long long)artis::ESC_CHUNK + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const bool sel = i < n && artis::escaped_selected(hot, reported, only_new, i);
  const unsigned m = __ballot_sync(0xffffffffu, sel);
  if (lane == 0) warp_base[w] = __popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = offsets[blockIdx.x];
    for (int k = 0; k < artis::ESC_CHUNK / 32; k++) {
      const int c = warp_base[k];
      warp_base[k] = run;
      run += c;
    }
  }
  __syncthreads();
  if (sel) {
    ArtisPacket a;
    artis::pkt_join(hot[i], cold[i], a);
    const uint4 *src = reinterpret_cast<const uint4 *>(&a);
    uint4 *dst = reinterpret_cast<uint4 *>(out + warp_base[w] + __popc(m & ((1u << lane) - 1u)));
#pragma unroll
    for (int q = 0; q < (int)(sizeof(ArtisPacket) / 16); q++) dst[q] = src[q];
    if (only_new) reported[i] = 1;
  }
}

// ---- barrier-free scheduler ------------------------------------------------------------------------------------------
// Persistent kernels, one ring of packet indices per event kind, warps specialised by kind. A warp that has no packet left
// picks the kind with the longest backlog (among the kinds its kernel serves) and keeps that role until it runs dry; its
// lanes claim ring entries (warp-aggregated), advance their packet by events of that kind only and, when the packet changes
// kind, write it back and push its index on the ring of the new kind. Nothing waits for the slowest packet of a round: the
// step ends when the longest single packet history ends, and a warp never executes two kinds' code in one iteration.
//
// Two kernels run side by side and exchange packets through the rings:
//   k_transport_light       the two roles that are chains of table reads on almost no packet state: macro-atom jumps (32
//                           bytes of the packet, one rate row and one prefix-sum search per jump) and the continuum-opacity
//                           sums of thin-cell r-packet steps (cell and frequency; ~900 bound-free terms per sum, staged by
//                           TMA) -- a small register budget and little code, hence many resident warps to hide the L2
//                           round trips and the instruction fetches
//   k_transport_persistent  every other role (r-packet steps around their opacity sum, gamma packets, pellets, macro-atom
//                           events, k-packets) on the whole hot line; big register budget, few warps
// (k_transport_persistent<ROLES_ALL> also serves the jump role itself: the single-kernel variant, kept for comparisons.)
//
// Ring protocol (capacity >= number of packets, so at most `cap` entries are ever outstanding): a slot holds -1 when
// empty. Producers take a ticket from tail (atomicAdd) and CAS their index into slot ticket % cap; consumers CAS head
// forward only over tickets below the tail they observed and then exchange the slot back to -1, spinning for the few
// cycles a producer may still need to fill it.
// Memory ordering without L1 invalidations: the producer's CAS on the slot is an atom.release.gpu (MEMBAR.ALL.GPU: every
// store of the packet's records has reached L2 before the index becomes visible); the consumer's exchange is relaxed and
// everything it then reads of the packet comes through L2 (ld.global.cg; common.cuh ldcold) at an address that depends on
// the index it took. __threadfence() on either side would add a CCTL.IVALL -- an invalidation of the SM's whole L1, where
// the read-only atomic data and per-cell table rows of all resident warps live -- to every hand-over (measured in round 1
// without knowing it: 81 % L1 hit rate with 4e7 hand-overs per timestep).
using artis::NRING;
using artis::PktCold;
using artis::PktHot;
constexpr unsigned ROLES_HEAVY = (1u << artis::KIND_MAK) | (1u << artis::KIND_RPKT_THIN) | (1u << artis::KIND_GAMMA) |
                                 (1u << artis::KIND_OTHER) | (1u << artis::KIND_RPKT_GREY);
constexpr unsigned ROLES_ALL = ROLES_HEAVY | (1u << artis::KIND_MAJUMP);
// register budgets of the two persistent kernels, as resident 128-thread blocks per SM that each must allow on its own
// (__launch_bounds__: 65536 / (128 * blocks) registers per thread): 3 -> 168 registers, 2 -> 255, 8 -> 64, 6 -> 80
#ifndef ARTIS_HEAVY_MINBLOCKS
#define ARTIS_HEAVY_MINBLOCKS 3
#endif
// ARTIS_HEAVY_MAXNREG: an explicit register cap for the big kernel instead of one derived from blocks per SM (160: two of its
// blocks leave room for three blocks of the jump kernel; measurement builds)
#ifdef ARTIS_HEAVY_MAXNREG
#define ARTIS_HEAVY_BOUNDS __maxnreg__(ARTIS_HEAVY_MAXNREG)
#else
#define ARTIS_HEAVY_BOUNDS __launch_bounds__(128, ARTIS_HEAVY_MINBLOCKS)
#endif
#ifndef ARTIS_JUMP_MINBLOCKS
#define ARTIS_JUMP_MINBLOCKS 8
#endif

struct PersistView {
  int *ring[NRING];
  unsigned long long *head;  // [NRING]
  unsigned long long *tail;  // [NRING]
  int *n_active;             // packets that have not finished the timestep
  int n_initial;             // its value at launch
  unsigned long long *t_start;  // globaltimer at launch
  int *partner_started;         // two-kernel scheduler: set by the jump kernel when it starts (nullptr: one kernel)
  int cap;
  unsigned long long *pend;  // [cap] macro-atom action a jump has drawn and left for the event role: (action + 1) << 32 | selidx
  double *kcache;            // [cap][kstride] continuum opacity delivered for a packet: KappaCache scalars + gamma_contr
  double *stepstate;         // [cap][3] RpktStepState of a step that waits for its opacity
  int kstride;
};

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p) {
  return *reinterpret_cast<const volatile unsigned long long *>(p);
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// hot record <-> working set, through L2
__device__ __forceinline__ void pkt_load_cg(artis::Pkt &p, artis::Rng &rng, const artis::Ctx &c, const PktHot *src, PktCold *cold) {
  PktHot h;
  const uint4 *s = reinterpret_cast<const uint4 *>(src);
  uint4 *d = reinterpret_cast<uint4 *>(&h);
#pragma unroll
  for (int i = 0; i < (int)(sizeof(PktHot) / 16); i++) d[i] = __ldcg(s + i);
  artis::pkt_from_hot(p, h, cold);
  rng = artis::rng_init(c, h.number, (unsigned long long)h.rngctr);
}
__device__ __forceinline__ void pkt_store_cg(PktHot *dst, const artis::Pkt &p, const artis::Rng &rng) {
  PktHot h;
  artis::pkt_to_hot(h, p, (unsigned)rng.ctr);
  uint4 *d = reinterpret_cast<uint4 *>(dst);
  const uint4 *s = reinterpret_cast<const uint4 *>(&h);
#pragma unroll
  for (int i = 0; i < (int)(sizeof(PktHot) / 16); i++) __stcg(d + i, s[i]);
}
// opacity hand-over record: [0] nu, [1] total, [2] es, [3] ff, [4] bf, [5] ffheating, [6] mgi, [8 + g] gamma_contr[g]
__device__ __forceinline__ void kcache_load(artis::KappaCache &k, const double *kc, const int ng) {
  k.nu = __ldcg(kc + 0);
  k.total = __ldcg(kc + 1);
  k.es = __ldcg(kc + 2);
  k.ff = __ldcg(kc + 3);
  k.bf = __ldcg(kc + 4);
  k.ffheating = __ldcg(kc + 5);
  k.mgi = (int)__ldcg(kc + 6);
  k.valid = true;
  for (int g = 0; g < ng; g++) k.gamma_contr[g] = __ldcg(kc + 8 + g);
}
// ring primitives (protocol in the comment above)
constexpr int RMS = artis::RING_META_STRIDE;
// 1: the two-kernel scheduler hands the opacity sums of the thin-cell steps to the light kernel (KIND_KAPPA role); 0: the
// thin role's own warp evaluates them between the two halves of the step (measured: profiles/r2_notes.md)
#ifndef ARTIS_KAPPA_OFFLOAD
#define ARTIS_KAPPA_OFFLOAD 0
#endif
// Virtual packets in the barrier-free scheduler. ARTIS_VPKT_DEFER 1 (default): events record their requests and k_vpkt_trace
// traces the list when the scheduler has finished; 0: the rays are traced inside the transport kernel, right after the event --
// ARTIS_VPKT_COOP 1: by the whole warp when few lanes have a request (vpkt.cuh coop_vpkt_flush), 0: by the lane alone.
// ARTIS_VPKT_TRACE_MODE: how k_vpkt_trace works through the list -- 2 (default): a lane per request, the warp in lock step through
// the phases of a cell (vpkt.cuh vpkt_trace_lockstep); 0: a thread per request at its own pace; 1: a warp per request.
#ifndef ARTIS_VPKT_DEFER
#define ARTIS_VPKT_DEFER 1
#endif
#ifndef ARTIS_VPKT_COOP
#define ARTIS_VPKT_COOP 1
#endif
#ifndef ARTIS_VPKT_TRACE_MODE
#define ARTIS_VPKT_TRACE_MODE 2
#endif
#ifndef ARTIS_VPKT_TRACE_MINBLOCKS
#define ARTIS_VPKT_TRACE_MINBLOCKS 4
#endif
// rows on demand: 1 = the event role's warps build the cooling rows their k-packets are about to search with all 32 lanes
// before the event; 0 = the lane that misses a row builds it alone (A/B builds)
#ifndef ARTIS_COOP_COOL_ROWS
#define ARTIS_COOP_COOL_ROWS 1
#endif
#ifndef ARTIS_REFILL_BACKOFF
#define ARTIS_REFILL_BACKOFF 10
#endif
constexpr int REFILL_BACKOFF = ARTIS_REFILL_BACKOFF;
__device__ __forceinline__ long long ring_backlog(const PersistView &pv, const int r) {
  const unsigned long long h = ld_volatile_u64(pv.head + r * RMS), t = ld_volatile_u64(pv.tail + r * RMS);
  return t > h ? (long long)(t - h) : 0;
}
// lane 0: claim up to `want` entries of ring r; returns the count and the first ticket
__device__ __forceinline__ int ring_claim(const PersistView &pv, const int r, const int want, unsigned long long *base) {
  for (int tries = 0; tries < 4; tries++) {
    const unsigned long long h = ld_volatile_u64(pv.head + r * RMS), t = ld_volatile_u64(pv.tail + r * RMS);
    const long long avail = t > h ? (long long)(t - h) : 0;
    const int take = avail < want ? (int)avail : want;
    if (take == 0) return 0;
    if (atomicCAS(pv.head + r * RMS, h, h + (unsigned long long)take) == h) {
      *base = h;
      return take;
    }
  }
  return 0;
}
__device__ __forceinline__ int ring_take(const PersistView &pv, const int r, const unsigned long long ticket) {
  int *slot = pv.ring[r] + (size_t)(ticket % (unsigned long long)pv.cap);
  int v;
  while ((v = atomicExch(slot, -1)) < 0) {
  }
  return v;
}
__device__ __forceinline__ void ring_put(const PersistView &pv, const int r, const unsigned long long ticket, const int idx) {
  int *slot = pv.ring[r] + (size_t)(ticket % (unsigned long long)pv.cap);
  int old;
  do {
    asm volatile("atom.release.gpu.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(slot), "r"(-1), "r"(idx) : "memory");
  } while (old != -1);
}
// The warp's finished lanes (done) push their packets on the rings of their next kinds (k2 >= 0) or retire them (k2 < 0).
__device__ __forceinline__ void push_done(const artis::Ctx &c, const PersistView &pv, const bool done, const int k2, const int idx) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int kk = 0; kk < NRING; kk++) {
    const unsigned m = __ballot_sync(full, done && k2 == kk);
    if (m != 0u) {
      const int leader = __ffs(m) - 1;
      unsigned long long b = 0ull;
      if (lane == leader) b = atomicAdd(pv.tail + kk * RMS, (unsigned long long)__popc(m));
      b = __shfl_sync(full, b, leader);
      if (done && k2 == kk) ring_put(pv, kk, b + (unsigned long long)__popc(m & lt_mask), idx);
    }
  }
  const unsigned fin = __ballot_sync(full, done && k2 < 0);
  if (fin != 0u && lane == __ffs(fin) - 1) {
    const int nfin = __popc(fin);
    const int before = atomicSub(pv.n_active, nfin), after = before - nfin;
    // drain profile: when the active count falls through 10%, 1% and 0 of its initial value (ns since launch)
    const int n0 = pv.n_initial;
    const long long now = (long long)(global_ns() - *reinterpret_cast<volatile unsigned long long *>(pv.t_start));
    if (before > n0 / 10 && after <= n0 / 10) c.es.counters[ARTIS_CTR_DRAIN_NS_10PCT] = now;
    if (before > n0 / 100 && after <= n0 / 100) c.es.counters[ARTIS_CTR_DRAIN_NS_1PCT] = now;
    if (after <= 0) c.es.counters[ARTIS_CTR_DRAIN_NS_DONE] = now;
  }
}
// An idle warp's poll: returns true when the timestep is over (no active packet left, or the hang guard fired)
__device__ __forceinline__ bool idle_poll(const artis::Ctx &c, const PersistView &pv, unsigned long long &idle_since, int &last_active,
                                          const unsigned long long idle_timeout_ns) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  int left = 0;
  if (lane == 0) left = *reinterpret_cast<volatile int *>(pv.n_active);
  left = __shfl_sync(full, left, 0);
  if (left <= 0) return true;  // every packet has reached the end of the timestep
  if (lane == 0 && c.error_flag != nullptr) left = (*reinterpret_cast<volatile int *>(c.error_flag) != 0) ? -1 : left;
  left = __shfl_sync(full, left, 0);
  if (left < 0) return true;   // somebody has given the timestep up
  unsigned long long now = 0ull;
  if (lane == 0) now = global_ns();
  now = __shfl_sync(full, now, 0);
  // Hang guard: the timestep is given up (code 900) only when the number of active packets has not changed for the whole
  // timeout -- a warp that merely finds its rings empty while a few long histories finish elsewhere keeps waiting.
  if (idle_since == 0ull || left != last_active) {
    idle_since = now;
    last_active = left;
  }
  // The two kernels of the pair need each other. Under a profiler, a debugger or CUDA_LAUNCH_BLOCKING the runtime runs
  // kernels one after the other: this kernel then waits for a partner that cannot start before it ends. One second without
  // progress and without the partner's start flag: give up with code 901; the host switches to the one-kernel scheduler and
  // launches the unfinished packets again (engine.inc).
  if (pv.partner_started != nullptr) {
    int give_up = 0;
    if (lane == 0 && now - idle_since > 1000000000ull && *reinterpret_cast<volatile int *>(pv.partner_started) == 0) {
      if (c.error_flag) atomicCAS(c.error_flag, 0, 901);
      give_up = 1;
    }
    if (__shfl_sync(full, give_up, 0) != 0) return true;
  }
  if (now - idle_since > idle_timeout_ns) {  // a lost packet would otherwise hang the device
    if (lane == 0 && c.error_flag) atomicCAS(c.error_flag, 0, 900);
    return true;
  }
  __nanosleep(1000);
  return false;
}

// ---- warp-cooperative continuum opacity with staged tiles (K2) ---------------------------------------------------------
// Every lane of a thin-cell r-packet warp that is about to need kappa_cont for its own packet and misses its cache
// ("owner") gets its bound-free sum evaluated by all 32 lanes, a tile of 32 consecutive continua at a time. The result is
// left in the owner's cache exactly as calculate_kappa_rpkt_cont would leave it, so the unchanged get_event that follows
// finds a hit. Summation order = rpkt.cuh BF_TILE trees (tile = 32 consecutive terms reduced as a binary tree, tile sums
// added in order), the same as the per-lane loop and the host build: identical bits.
//
// What a term reads: the cell's (level population, departure ratio) pair -- 16 bytes of the row bf_cell[cell][.] -- its
// static 32-byte ContRec, and two neighbouring floats of its cross-section table. The first two are CONTIGUOUS over the
// terms of one sum (rows indexed by continuum), ~900 terms = 43 KB per evaluation: they are brought into shared memory by
// the TMA engine (cp.async.bulk, one 1-D bulk copy per array and chunk, completion on an mbarrier), two chunks ahead of the
// arithmetic and across the boundary between one owner's sum and the next, so that the lanes never wait for them and
// issue no load instructions for them; the dependent table gather is the only global load left in the loop. In round 1
// the same sum issued 3-4 dependent L2 loads per term from two warps per scheduler and was 61 % of the timestep.
static_assert(artis::BF_TILE == 32, "one continuum per lane");
#ifndef ARTIS_KAPPA_CHUNK_TILES
#define ARTIS_KAPPA_CHUNK_TILES 2
#endif
constexpr int KCH_TILES = ARTIS_KAPPA_CHUNK_TILES;          // tiles per staged chunk
constexpr int KCH = KCH_TILES * artis::BF_TILE;            // terms per chunk
constexpr int KSTAGES = 2;
struct alignas(16) KappaStage {
  double2 cell[KSTAGES][KCH];          // bf_cell pairs
  artis::ContRec rec[KSTAGES][KCH];    // static records
  unsigned long long bar[KSTAGES];     // one mbarrier per stage (this warp's)
};
constexpr size_t KAPPA_SMEM_PER_BLOCK = 4 * sizeof(KappaStage);  // 128-thread blocks
constexpr size_t LIGHT_SMEM = ARTIS_KAPPA_OFFLOAD != 0 ? KAPPA_SMEM_PER_BLOCK : 0;  // dynamic shared memory of the pair's kernels
constexpr size_t HEAVY_SMEM = ARTIS_KAPPA_OFFLOAD != 0 ? 0 : KAPPA_SMEM_PER_BLOCK;
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, const unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, const unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, const unsigned bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, const unsigned parity) {
  unsigned ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (ok == 0u);
}
// called once per warp before the first evaluation
__device__ __forceinline__ void kappa_stage_init(KappaStage *ks) {
  if ((threadIdx.x & 31) == 0) {
    for (int s2 = 0; s2 < KSTAGES; s2++) mbar_init(&ks->bar[s2], 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();
}
// `nchunk`: chunks this warp has staged since the kernel began (stage = nchunk % KSTAGES, mbarrier parity from it).
// need / nu_cmf / mgi: this lane's request; returns its bound-free sum; gc(slot, v) is called on the owner lane for every
// ground-continuum contribution, in index order like the serial loop (rpkt.cc:1158-1166).
// -DARTIS_KAPPA_PROFILE: where the cycles of coop_kappa_sum go (measurement builds): counters[54..56], [59], [60] and [55]...
#ifdef ARTIS_KAPPA_PROFILE
#define KP_BEGIN() long long kp_t = clock64(), kp_acc[6] = {0, 0, 0, 0, 0, 0}
#define KP_ADD(i)                 \
  do {                            \
    const long long kp_n = clock64(); \
    kp_acc[i] += kp_n - kp_t;     \
    kp_t = kp_n;                  \
  } while (0)
#define KP_END()                                                                                           \
  do {                                                                                                     \
    if ((threadIdx.x & 31) == 0) {                                                                         \
      const int kp_slot[6] = {54, 55, 56, 58, 59, 60};                                                     \
      for (int kp_i = 0; kp_i < 6; kp_i++)                                                                 \
        atomicAdd((unsigned long long *)&c.es.counters[kp_slot[kp_i]], (unsigned long long)kp_acc[kp_i]); \
    }                                                                                                      \
  } while (0)
#else
#define KP_BEGIN()
#define KP_ADD(i)
#define KP_END()
#endif
template <class GC>
__device__ __forceinline__ double coop_kappa_sum(const artis::Ctx &c, const bool need, const int mgi, const double nu_cmf,
                                                 artis::Tally &tl, KappaStage *ks, unsigned &nchunk, GC gc) {
  using namespace artis;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned owners = __ballot_sync(full, need);
  if (owners == 0u) return 0.;
  KP_BEGIN();
  // every owner prepares its own sum in parallel: the range of continua, exp(-h nu / k T_e)
  double kappa_bf = 0., my_expfac = 0.;
  int my_ilo = 0, my_i0 = 0;
  const double *my_pair = nullptr;  // the list the sum runs over (rpkt.cuh bf_list): the cell's live continua
  const ContRec *my_rec = nullptr;
  if (need) {
    const int tc = tabcell(c, mgi);
    const BfList l = bf_list(c, tc);
    my_pair = l.pair;
    my_rec = l.rec;
    my_expfac = exp(-HOVERKB * nu_cmf / (double)c.cs.Te[mgi]);
    int listed;
    bf_list_range(c, tc, nu_cmf, &my_ilo, &my_i0, &listed);
    tl.bf_evaluated += my_i0 - my_ilo;
    tl.bf_listed += listed;  // what the reference's loop would visit: the roofline accounting reports both
  }
  // the issue cursor runs KSTAGES chunks ahead of the arithmetic, across owners
  unsigned iss_rest = owners;
  int iss_j = 0, iss_nch = 0, iss_ilo = 0, iss_i0 = 0;
  const double *iss_pair = nullptr;
  const ContRec *iss_rec = nullptr;
  unsigned iss_count = nchunk;
  auto iss_next_owner = [&]() {  // warp-uniform
    iss_nch = 0;
    while (iss_rest != 0u && iss_nch == 0) {
      const int o = __ffs(iss_rest) - 1;
      iss_rest &= iss_rest - 1u;
      iss_ilo = __shfl_sync(full, my_ilo, o);
      iss_i0 = __shfl_sync(full, my_i0, o);
      iss_pair = reinterpret_cast<const double *>(__shfl_sync(full, reinterpret_cast<unsigned long long>(my_pair), o));
      iss_rec = reinterpret_cast<const ContRec *>(__shfl_sync(full, reinterpret_cast<unsigned long long>(my_rec), o));
      iss_j = 0;
      iss_nch = (iss_i0 - iss_ilo + KCH - 1) / KCH;
    }
  };
  auto issue_one = [&]() {  // warp-uniform; lane 0 talks to the TMA engine
    if (iss_nch == 0) return;
    if (lane == 0) {
      const int st = (int)(iss_count % KSTAGES);
      const int base = iss_ilo + iss_j * KCH;
      const int n = (iss_i0 - base) < KCH ? (iss_i0 - base) : KCH;
      // the lanes have finished reading this stage (__syncwarp before the call): order their reads before the engine's writes
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_expect_tx(&ks->bar[st], (unsigned)n * 48u);
      bulk_g2s(&ks->cell[st][0], iss_pair + 2 * (size_t)base, (unsigned)n * 16u, &ks->bar[st]);
      bulk_g2s(&ks->rec[st][0], iss_rec + base, (unsigned)n * 32u, &ks->bar[st]);
    }
    iss_count++;
    iss_j++;
    if (iss_j == iss_nch) iss_next_owner();
  };
  KP_ADD(0);  // prologue: every owner's range, exp, list pointers
  iss_next_owner();
#pragma unroll
  for (int s2 = 0; s2 < KSTAGES; s2++) issue_one();
  KP_ADD(3);
  unsigned rest = owners;
  while (rest != 0u) {
    const int owner = __ffs(rest) - 1;
    rest &= rest - 1u;
    const double nu = __shfl_sync(full, nu_cmf, owner);
    const double expfac = __shfl_sync(full, my_expfac, owner);
    const int ilo = __shfl_sync(full, my_ilo, owner);
    const int i0 = __shfl_sync(full, my_i0, owner);
    double sum = 0.;
    for (int base = ilo; base < i0; base += KCH) {
      const int st = (int)(nchunk % KSTAGES);
      KP_ADD(4);  // owner set-up (shuffles) and loop overhead
      mbar_wait(&ks->bar[st], (nchunk / KSTAGES) & 1u);
      KP_ADD(1);
      nchunk++;
      double t[KCH_TILES], gcontr[KCH_TILES];
      int slot[KCH_TILES];
#pragma unroll
      for (int u = 0; u < KCH_TILES; u++) {
        const int q = u * BF_TILE + lane;
        t[u] = 0.;
        gcontr[u] = 0.;
        slot[u] = -1;
        if (base + q < i0) {
          const double2 cellv = ks->cell[st][q];
          const ContRec r = ks->rec[st][q];
          t[u] = kappa_bf_term_fast(c, cellv.x, cellv.y, r, nu, expfac, &gcontr[u]);
          if (gcontr[u] != 0.) slot[u] = r.ground_slot;
        }
      }
      __syncwarp(full);  // every lane has its operands out of the stage: it can be refilled
      KP_ADD(2);  // term arithmetic
      issue_one();
      KP_ADD(3);  // copy issue
#pragma unroll
      for (int off = 1; off < BF_TILE; off <<= 1) {  // adjacent pairs first: the tree of rpkt.cuh PairTree
#pragma unroll
        for (int u = 0; u < KCH_TILES; u++) t[u] += __shfl_down_sync(full, t[u], off);
      }
#pragma unroll
      for (int u = 0; u < KCH_TILES; u++) {
        sum += t[u];  // lane 0 holds each tile sum in tree order; only lane 0's running sum is used
        unsigned g = __ballot_sync(full, slot[u] >= 0);
        while (g != 0u) {  // ground-continuum contributions, in index order like the serial loop
          const int src = __ffs(g) - 1;
          g &= g - 1u;
          const double v = __shfl_sync(full, gcontr[u], src);
          const int sl = __shfl_sync(full, slot[u], src);
          if (lane == owner) gc(sl, v);
        }
      }
      KP_ADD(5);
    }
    KP_ADD(5);  // reductions and ground-continuum contributions (of the owner's last chunk; the others fall under 4)
    sum = __shfl_sync(full, sum, 0);
    if (lane == owner) kappa_bf = sum;
  }
  KP_END();
  return kappa_bf;
}
// the single-kernel variant: the thin role's own warp evaluates the sums of its lanes and leaves them in their caches
__device__ __forceinline__ void coop_kappa_warm(const artis::Ctx &c, const bool wants, const int mgi, const double nu_cmf,
                                                artis::KappaCache &k, artis::Tally &tl, KappaStage *ks, unsigned &nchunk) {
  using namespace artis;
  const bool need = wants && c.st.do_r_lc && c.st.opacity_case == 4 && !kappa_cache_hit(k, mgi, nu_cmf);
  double sigma = 0., kappa_ff = 0.;
  if (need) {
    tl.kappa_evals++;
    sigma = SIGMA_T * c.cs.nne[mgi];
    kappa_ff = calculate_kappa_ff(c, mgi, nu_cmf);
    const int ng = c.st.nbfcontinua_ground;
    for (int g = 0; g < ng; g++) k.gamma_contr[g] = 0.;
  }
  const double kappa_bf = coop_kappa_sum(c, need, mgi, nu_cmf, tl, ks, nchunk, [&](const int sl, const double v) { k.gamma_contr[sl] += v; });
  if (need) kappa_store(c, k, mgi, nu_cmf, sigma, kappa_ff, kappa_bf, kappa_ff, tl);
}

// Warp-cooperative bf-rate estimator sweep (DETAILED_BF_ESTIMATORS_ON, radfield.cc:764-829): an r-packet step in a
// non-grey cell adds gamma_contr[i] * distance_e_cmf / nu to the estimator of EVERY continuum whose table covers nu --
// about a thousand dependent table reads and FP64 atomics on one lane, in a role whose warps are sparsely filled in the
// optically thin nebular models (2-6 busy lanes). update_estimators leaves the request in the lane's KappaCache and the
// whole warp serves the requests one after another, lane l taking continua l, l + 32, ...: the same terms and products,
// so the estimator sums differ from the serial sweep by the (already unordered) order of the atomics only.
__device__ __forceinline__ void coop_bfest_flush(const artis::Ctx &c, artis::KappaCache &k, artis::Tally &tl) {
  using namespace artis;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  unsigned owners = __ballot_sync(full, k.bfest_factor != 0.);
  while (owners != 0u) {
    const int owner = __ffs(owners) - 1;
    owners &= owners - 1u;
    const double knu = __shfl_sync(full, k.nu, owner);
    const double nu = __shfl_sync(full, k.bfest_nu, owner);
    const double factor = __shfl_sync(full, k.bfest_factor, owner);
    const int tc = __shfl_sync(full, k.bfest_tc, owner);
    const int mgi = __shfl_sync(full, k.mgi, owner);
    bfest_sweep(c, tc, mgi, knu, nu, factor, lane, 32, tl);
  }
  k.bfest_factor = 0.;
}

// rows on demand (celltables.cuh row_get): true once the row arena has passed its soft limit. The warps then put the
// packets they hold back into the hot array and leave; the host forgets the rows and launches the scheduler again on the
// packets that have not finished (engine.inc artis_gpu_transport_resident). Warp-uniform.
// The same exit serves the virtual-packet request list when it has passed its soft limit (vpkt.cuh vpkt_enqueue).
__device__ __forceinline__ bool row_arena_stop(const artis::Ctx &c, const int lane) {
#if !defined(ARTIS_NO_VPKT) && ARTIS_VPKT_DEFER
  if (ARTIS_OPT_VPKT(c)) {
    int vstop = 0;
    if (lane == 0) vstop = *reinterpret_cast<volatile int *>(c.vreq_stop);
    if (__shfl_sync(0xffffffffu, vstop, 0) != 0) return true;
  }
#endif
  if (!artis::rows_on_demand(c)) return false;
  int stop = 0;
  if (lane == 0) stop = *reinterpret_cast<volatile int *>(c.tb.row_stop);
  return __shfl_sync(0xffffffffu, stop, 0) != 0;
}

// ROLES: bit mask of the kinds the warps of this kernel may take up
template <unsigned ROLES>
__global__ void ARTIS_HEAVY_BOUNDS k_transport_persistent(PktHot *hot, PktCold *cold, PersistView pv, long long max_events,
                                                                       unsigned long long idle_timeout_ns) {
  using namespace artis;
  const Ctx &c = d_ctx;
  const double t2 = c.t2;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  constexpr bool HAS_THIN = (ROLES & (1u << KIND_RPKT_THIN)) != 0;
  constexpr bool HAS_JUMP = (ROLES & (1u << KIND_MAJUMP)) != 0;
  constexpr bool KAPPA_HERE = HAS_JUMP || ARTIS_KAPPA_OFFLOAD == 0;  // this kernel's thin role evaluates its own opacity sums
  Tally tl = {};
  KappaCache k;
  k.valid = false;
  k.defer_bfest = HAS_THIN && ARTIS_OPT_DETAILED_BF(c);
  bool have = false, resumed = false;
  int idx = 0, role = 0;
  long long events = 0;
  Pkt p;
  Rng rng;
  RpktStepState st;
  st.mgi = 0;
  unsigned long long idle_since = 0ull;
  int last_active = -1;  // active-packet count at the last idle poll: the hang guard looks for progress, not for work
  int refill_skip = 0;   // iterations left during which a partly busy warp does not look at its (just found empty) ring
  stats_block_begin();
  extern __shared__ __align__(16) unsigned char kappa_smem[];
  KappaStage *const ks = reinterpret_cast<KappaStage *>(kappa_smem) + (threadIdx.x >> 5);
  unsigned nchunk = 0u;
  if (KAPPA_HERE) kappa_stage_init(ks);
  if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long *>(pv.t_start) = global_ns();
  long long fill_iters = 0, fill_lanes = 0;  // iterations run in kind `lane`, and the busy lanes summed over them
  // where the warp's cycles go (lane k keeps kind k's event cycles; lane 0 the scheduler's own)
  long long cyc_kind = 0, cyc_refill = 0, cyc_push = 0, cyc_idle = 0;
  long long cyc_thin_begin = 0, cyc_thin_kappa = 0;  // lane 0: the thin role's cycles before and inside the opacity evaluation
  while (true) {
    __syncwarp(full);
    if (row_arena_stop(c, lane)) {
      if (have) pkt_store_cg(hot + idx, p, rng);
      break;
    }
    const long long tick0 = clock64();
    const unsigned need = __ballot_sync(full, !have);
    if (need != full && refill_skip > 0) {
      refill_skip--;
    } else if (need != 0u) {
      if (need == full) {  // the whole warp is out of work: choose the kind with the longest backlog
        long long backlog = 0;
        if (lane < NKIND && ((ROLES >> lane) & 1u) != 0) backlog = ring_backlog(pv, lane);
        if (HAS_THIN && !KAPPA_HERE) {  // the thin role drains two rings
          long long extra = 0;
          if (lane == KIND_RPKT_THIN) extra = ring_backlog(pv, KIND_THIN_RESUME);
          backlog += extra;
        }
        long long best = backlog;
        int best_kind = lane;
#pragma unroll
        for (int off = 4; off >= 1; off >>= 1) {
          const long long ob = __shfl_down_sync(full, best, off);
          const int ok = __shfl_down_sync(full, best_kind, off);
          if (ob > best) {
            best = ob;
            best_kind = ok;
          }
        }
        best = __shfl_sync(full, best, 0);
        best_kind = __shfl_sync(full, best_kind, 0);
        if (best <= 0) {
          if (idle_poll(c, pv, idle_since, last_active, idle_timeout_ns)) break;
          cyc_idle += clock64() - tick0;
          continue;
        }
        role = best_kind;
        idle_since = 0ull;
      }
      // refill the idle lanes from the role's ring(s); for the thin role the resumed steps come first
      const int want = __popc(need);
      int take_a = 0, take_b = 0;
      unsigned long long base_a = 0ull, base_b = 0ull;
      if (lane == 0) {
        if (HAS_THIN && !KAPPA_HERE && role == KIND_RPKT_THIN) take_a = ring_claim(pv, KIND_THIN_RESUME, want, &base_a);
        if (take_a < want) take_b = ring_claim(pv, role, want - take_a, &base_b);
      }
      take_a = __shfl_sync(full, take_a, 0);
      take_b = __shfl_sync(full, take_b, 0);
      base_a = __shfl_sync(full, base_a, 0);
      base_b = __shfl_sync(full, base_b, 0);
      const int take = take_a + take_b;
      const int rank = __popc(need & lt_mask);
      if (!have && rank < take) {
        const bool from_resume = rank < take_a;
        idx = from_resume ? ring_take(pv, KIND_THIN_RESUME, base_a + (unsigned long long)rank)
                          : ring_take(pv, role, base_b + (unsigned long long)(rank - take_a));
        pkt_load_cg(p, rng, c, hot + idx, cold + idx);
        k.valid = false;
        resumed = false;
        if (HAS_THIN && !KAPPA_HERE && from_resume) {
          kcache_load(k, pv.kcache + (size_t)idx * pv.kstride, c.st.nbfcontinua_ground);
          const double *ss = pv.stepstate + (size_t)idx * 3;
          st.tau_next = __ldcg(ss + 0);
          st.sdist = __ldcg(ss + 1);
          const long long packed = __double_as_longlong(__ldcg(ss + 2));
          st.snext = (int)(packed >> 32);
          st.mgi = (int)(packed & 0xffffffffll);
          st.mgi_next = rpkt_step_mgi_next(c, p, st);
          resumed = true;
        }
        events = 0;
        have = true;
      }
      if (need == full && take == 0) continue;  // lost the race for the last entries
      // a claim costs the warp two L2 round trips on the busiest words of the device: after an empty answer the busy
      // lanes run a few iterations before the warp asks again (a warp without any work asks at once)
      refill_skip = (take < want) ? REFILL_BACKOFF : 0;
    }
    __syncwarp(full);
    const long long tick1 = clock64();
    cyc_refill += tick1 - tick0;
    {
      const unsigned busy = __ballot_sync(full, have);
      if (lane == role && lane < NKIND) {  // lane k keeps the occupancy statistics of kind k
        fill_iters++;
        fill_lanes += __popc(busy);
      }
    }
    bool parked = false;  // thin lane that has handed its step to the opacity role
    if (HAS_THIN && role == KIND_RPKT_THIN) {  // warp-uniform: the step in two halves around the opacity evaluation
      bool proceed = false;
      if (have) {
        if (resumed) {
          proceed = true;
        } else {
          tl.packet_updates++;
          proceed = rpkt_step_begin<1>(c, p, rng, tl, st);
        }
      }
      __syncwarp(full);
      const long long tick_a = clock64();
      const bool wants = proceed && !resumed && c.st.do_r_lc && c.st.opacity_case == 4 && !kappa_cache_hit(k, st.mgi, p.nu_cmf);
      if constexpr (KAPPA_HERE) {
        // the sum is evaluated here, by this warp
        coop_kappa_warm(c, wants, st.mgi, p.nu_cmf, k, tl, ks, nchunk);
      } else if (wants) {
        // two-kernel scheduler: the step waits on the KIND_KAPPA ring while a warp of the light kernel evaluates its sum
        double *ss = pv.stepstate + (size_t)idx * 3;
        __stcg(ss + 0, st.tau_next);
        __stcg(ss + 1, st.sdist);
        __stcg(ss + 2, __longlong_as_double(((long long)st.snext << 32) | (long long)(unsigned)st.mgi));
        parked = true;
      }
      __syncwarp(full);
      if (lane == 0) {
        const long long tick_b = clock64();
        cyc_thin_begin += tick_a - tick1;
        cyc_thin_kappa += tick_b - tick_a;
      }
      if (proceed && !parked) rpkt_step_finish<1>(c, p, t2, k, rng, tl, st);
      if (k.defer_bfest) coop_bfest_flush(c, k, tl);  // warp-uniform: the flag is an option of the build mode
      resumed = false;
      if (have && !parked && p.type == ARTIS_TYPE_ESCAPE) est_add(&c.es.ts_scalars[ARTIS_TS_CMF_LUM], p.e_cmf);
    } else {
#if ARTIS_ROWS_ON_DEMAND && ARTIS_COOP_COOL_ROWS
      if (role == KIND_MAK && rows_on_demand(c)) {  // warp-uniform
        // cooling rows the lanes' k-packets are about to search and nobody has built yet: the warp builds them one after the
        // other with all 32 lanes (a lane that built one alone walked ~2000 terms while its warp waited)
        int b_mgi = 0, b_tc = -1, b_ui = -1;
        bool miss = false;
        if (have && p.type == ARTIS_TYPE_KPKT) {
          b_mgi = cell_mgi(c, p.where);
          b_tc = tabcell(c, b_mgi);
          if (b_tc >= 0 && c.cs.thick[b_mgi] != 1) b_ui = kpkt_peek_cooling_ion(c, p, t2, rng);  // (a grey cell: do_kpkt_bb)
          if (b_ui >= 0)
            miss = __ldcg(c.tb.row_ptr + (size_t)b_tc * (c.st.nlevels_total + c.st.includedions) + c.st.nlevels_total + b_ui) == 0u;
        }
        unsigned todo = __ballot_sync(full, miss);
        while (todo != 0u) {
          const int src = __ffs(todo) - 1;
          todo &= todo - 1u;
          coop_build_cool_row(c, __shfl_sync(full, b_mgi, src), __shfl_sync(full, b_tc, src), __shfl_sync(full, b_ui, src), lane);
        }
      }
#endif
    }
    if (!(HAS_THIN && role == KIND_RPKT_THIN) && have) {
      switch (role) {  // warp-uniform
        case KIND_MAK:
          if constexpr ((ROLES & (1u << KIND_MAK)) != 0) {
            if (p.type == ARTIS_TYPE_MA) {  // the rest of an iteration whose action the jump role has drawn
              const unsigned long long pd = __ldcg(pv.pend + idx);
              ma_event(c, p, (int)(pd >> 32) - 1, (int)(unsigned)(pd & 0xffffffffull), rng, tl);
            } else {
              do_packet_kind<KIND_MAK>(c, p, t2, k, rng, tl);
            }
          }
          break;
        case KIND_MAJUMP:
          if constexpr (HAS_JUMP) do_packet_kind<KIND_MAJUMP>(c, p, t2, k, rng, tl);
          break;
        case KIND_GAMMA:
          if constexpr ((ROLES & (1u << KIND_GAMMA)) != 0) do_packet_kind<KIND_GAMMA>(c, p, t2, k, rng, tl);
          break;
        case KIND_RPKT_GREY:
          if constexpr ((ROLES & (1u << KIND_RPKT_GREY)) != 0) do_packet_kind<KIND_RPKT_GREY>(c, p, t2, k, rng, tl);
          break;
        default:
          if constexpr ((ROLES & (1u << KIND_OTHER)) != 0) do_packet_kind<KIND_OTHER>(c, p, t2, k, rng, tl);
          break;
      }
    }
#if !defined(ARTIS_NO_VPKT) && ARTIS_VPKT_DEFER
    // virtual packets: the event has left a request; it is recorded for k_vpkt_trace and the packet goes on (vpkt.cuh)
    if (ARTIS_OPT_VPKT(c) && vpkt_enqueue(c, have, p, tl)) k.valid = false;
#elif !defined(ARTIS_NO_VPKT) && ARTIS_VPKT_COOP
    if (ARTIS_OPT_VPKT(c)) coop_vpkt_flush(c, have, p, k, tl);  // A/B builds: the rays traced here, by the whole warp / side by side
#else
    if (ARTIS_OPT_VPKT(c) && have) vpkt_flush_request(c, p, k, tl);  // A/B builds: one lane traces its rays alone
#endif
    bool done = false;
    int k2 = -1;
    if (have && parked) {
      pkt_store_cg(hot + idx, p, rng);
      done = true;
      k2 = KIND_KAPPA;
      have = false;
      tl.handoffs++;
    } else if (have) {
      events++;
      const bool active = (p.type != ARTIS_TYPE_ESCAPE) && (p.prop_time < t2);
      const int knext = active ? kind_of(c, p) : -1;
      // a macro-atom that stays activated after an event of the MAK role goes back to the jump role
      if (!(active && knext == role && events < max_events)) {
        pkt_store_cg(hot + idx, p, rng);
        done = true;
        k2 = knext;
        have = false;
        if (active) tl.handoffs++;
      }
    }
    __syncwarp(full);
    const long long tick2 = clock64();
    if (lane == role) cyc_kind += tick2 - tick1;
    if (__ballot_sync(full, done) != 0u) {
      push_done(c, pv, done, k2, idx);
      cyc_push += clock64() - tick2;
    }
  }
  tally_flush(c, tl);
  if (lane < NKIND && fill_iters > 0) {
    ctr_add(&c.es.counters[ARTIS_CTR_WARP_ITERS0 + lane], fill_iters);
    ctr_add(&c.es.counters[ARTIS_CTR_BUSY_LANES0 + lane], fill_lanes);
    ctr_add(&c.es.counters[ARTIS_CTR_WARP_CYCLES0 + lane], cyc_kind);
  }
  if (lane == 0) {
    ctr_add(&c.es.counters[ARTIS_CTR_REFILL_CYCLES], cyc_refill);
    ctr_add(&c.es.counters[ARTIS_CTR_PUSH_CYCLES], cyc_push);
    ctr_add(&c.es.counters[ARTIS_CTR_IDLE_CYCLES], cyc_idle);
    if (cyc_thin_begin) ctr_add(&c.es.counters[ARTIS_CTR_THIN_BEGIN_CYCLES], cyc_thin_begin);
    if (cyc_thin_kappa) ctr_add(&c.es.counters[ARTIS_CTR_THIN_KAPPA_CYCLES], cyc_thin_kappa);
  }
  stats_block_flush(c);
}

// The light kernel: the two roles that need almost no packet state and are chains of table reads -- a small register budget,
// hence many resident warps to hide the L2 round trips and the instruction fetches.
//   KIND_MAJUMP  macro-atom jumps within an ion. A lane holds 32 bytes of its packet (cell, macro-atom state, interaction
//                count) and the packet's Philox stream; an iteration is ma_jump: one row of process rates, one or two draws,
//                one search over a prefix-sum row. Jumps that stay inside the ion (98 % of them) only change the level; any
//                other action is left in pend[] for the event role of k_transport_persistent and the packet moves to that ring.
//   KIND_KAPPA   continuum-opacity sums for thin-cell r-packet steps that k_transport_persistent has begun: a lane reads the
//                packet's cell and frequency, the warp evaluates the requests of its lanes one after the other with all 32
//                lanes (coop_kappa_sum, tiles staged by TMA), the result goes to the packet's kcache row and the packet to
//                the KIND_THIN_RESUME ring, where a thin-role warp finishes the step.
struct MaView {
  int where, ma_element, ma_ion, ma_level, interactions;
};
__global__ void __launch_bounds__(128, ARTIS_JUMP_MINBLOCKS) k_transport_light(PktHot *hot, PersistView pv, long long max_events,
                                                                              unsigned long long idle_timeout_ns) {
  using namespace artis;
  const Ctx &c = d_ctx;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  Tally tl = {};
  bool have = false;
  int idx = 0, role = KIND_MAJUMP;
  int events = 0;
  const int event_cap = max_events > 0x7fffffffLL ? 0x7fffffff : (int)max_events;
  MaView v = {0, 0, 0, 0, 0};
  Rng rng = {};
  if (blockIdx.x == 0 && threadIdx.x == 0 && pv.partner_started != nullptr) *reinterpret_cast<volatile int *>(pv.partner_started) = 1;
  unsigned long long idle_since = 0ull;
  int last_active = -1;
  int refill_skip = 0;
  // statistics in 32-bit words (this kernel lives on its small register budget): iterations, busy lanes, and cycles in units
  // of 64 (lanes KIND_MAJUMP and KIND_KAPPA: the role's own cycles; lane 0: refill, push, idle)
  unsigned fill_iters = 0, fill_lanes = 0, cyc_a = 0, cyc_b = 0, cyc_c = 0;
  stats_block_begin();
  extern __shared__ __align__(16) unsigned char kappa_smem[];
  KappaStage *const ks = reinterpret_cast<KappaStage *>(kappa_smem) + (threadIdx.x >> 5);
  unsigned nchunk = 0u;
  if (ARTIS_KAPPA_OFFLOAD != 0) kappa_stage_init(ks);
  while (true) {
    __syncwarp(full);
    if (row_arena_stop(c, lane)) {
      if (have && role == KIND_MAJUMP) {
        __stcg(&hot[idx].ma_el_ion, pkt_pack2(v.ma_element, v.ma_ion));
        __stcg(&hot[idx].ma_level, v.ma_level);
        __stcg(&hot[idx].interactions, v.interactions);
        __stcg(&hot[idx].rngctr, (unsigned)rng.ctr);
      }
      break;
    }
    const unsigned tick0 = (unsigned)clock();
    const unsigned need = __ballot_sync(full, !have);
    if (need != full && refill_skip > 0) {
      refill_skip--;
    } else if (need != 0u) {
      if (need == full) {  // out of work: the role with the longer backlog
        long long bj = 0, bk = 0;
        if (lane == 0) {
          bj = ring_backlog(pv, KIND_MAJUMP);
          bk = ARTIS_KAPPA_OFFLOAD != 0 ? ring_backlog(pv, KIND_KAPPA) : 0;
        }
        bj = __shfl_sync(full, bj, 0);
        bk = __shfl_sync(full, bk, 0);
        if (bj <= 0 && bk <= 0) {
          if (idle_poll(c, pv, idle_since, last_active, idle_timeout_ns)) break;
          if (lane == 0) cyc_c += ((unsigned)clock() - tick0) >> 6;
          continue;
        }
        role = (bk > 0 && bk >= bj) ? KIND_KAPPA : KIND_MAJUMP;
        idle_since = 0ull;
      }
      const int want = __popc(need);
      int take = 0;
      unsigned long long base = 0ull;
      if (lane == 0) take = ring_claim(pv, role, want, &base);
      take = __shfl_sync(full, take, 0);
      base = __shfl_sync(full, base, 0);
      if (need == full && take == 0) continue;  // lost the race for the last entries
      const int rank = __popc(need & ((1u << lane) - 1u));
      if (!have && rank < take) {
        idx = ring_take(pv, role, base + (unsigned long long)rank);
        if (role == KIND_MAJUMP) {
          // sector 0 of the hot record and the (draw counter, number) pair of sector 1
          const uint4 a = __ldcg(reinterpret_cast<const uint4 *>(hot + idx));
          const uint4 b = __ldcg(reinterpret_cast<const uint4 *>(hot + idx) + 1);
          const uint2 r = __ldcg(reinterpret_cast<const uint2 *>(&hot[idx].rngctr));
          v.where = (int)a.x;
          v.ma_element = pkt_lo16(a.z);
          v.ma_ion = pkt_hi16(a.z);
          v.ma_level = (int)a.w;
          v.interactions = (int)b.y;
          rng = rng_init(c, (int)r.y, (unsigned long long)r.x);
        }
        events = 0;
        have = true;
      }
      refill_skip = (take < want) ? REFILL_BACKOFF : 0;
    }
    __syncwarp(full);
    const unsigned tick1 = (unsigned)clock();
    if (lane == 0) cyc_a += (tick1 - tick0) >> 6;
    {
      const unsigned busy = __ballot_sync(full, have);
      if (lane == role) {
        fill_iters++;
        fill_lanes += __popc(busy);
      }
    }
#if ARTIS_ROWS_ON_DEMAND
    if (role == KIND_MAJUMP && rows_on_demand(c)) {  // warp-uniform
      // rows the lanes' jumps need and nobody has built yet: the warp builds them one after the other with all 32 lanes
      int b_mgi = 0, b_tc = -1, b_gl = 0;
      bool miss = false;
      if (have) {
        b_mgi = cell_mgi(c, v.where);
        b_tc = tabcell(c, b_mgi);
        if (b_tc >= 0 && c.cs.thick[b_mgi] != 1) {
          b_gl = c.st.ion_leveloffset[uion_of(c, v.ma_element, v.ma_ion)] + v.ma_level;
          miss = __ldcg(c.tb.row_ptr + (size_t)b_tc * (c.st.nlevels_total + c.st.includedions) + b_gl) == 0u;
        }
      }
      unsigned todo = __ballot_sync(full, miss);
      while (todo != 0u) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1u;
        coop_build_ma_row(c, __shfl_sync(full, b_mgi, src), __shfl_sync(full, b_tc, src), __shfl_sync(full, b_gl, src), lane);
      }
    }
#endif
    bool done = false;
    int k2 = -1;
    if (ARTIS_KAPPA_OFFLOAD != 0 && role == KIND_KAPPA) {  // warp-uniform
      // rpkt.cc:1209-1309 calculate_kappa_rpkt_cont for the packets of the warp's lanes, written to their kcache rows
      int mgi = 0;
      double nu_cmf = 0., sigma = 0., kappa_ff = 0.;
      double *kc = nullptr;
      if (have) {
        mgi = cell_mgi(c, __ldcg(&hot[idx].where));
        nu_cmf = __ldcg(&hot[idx].nu_cmf);
        kc = pv.kcache + (size_t)idx * pv.kstride;
        tl.kappa_evals++;
        sigma = SIGMA_T * c.cs.nne[mgi];
        kappa_ff = calculate_kappa_ff(c, mgi, nu_cmf);
        const int ng = c.st.nbfcontinua_ground;
        for (int g = 0; g < ng; g++) __stcg(kc + 8 + g, 0.);
      }
      const double kappa_bf = coop_kappa_sum(c, have, mgi, nu_cmf, tl, ks, nchunk, [&](const int sl, const double val) {
        __stcg(kc + 8 + sl, __ldcg(kc + 8 + sl) + val);
      });
      if (have) {
        double total = sigma + kappa_bf + kappa_ff, es = sigma, ff = kappa_ff, bf = kappa_bf;
        if (!isfinite(total)) {  // rpkt.cc:1280-1294 soft fallback, as rpkt.cuh kappa_store
          if (isfinite(es)) {
            ff = 0.;
            bf = 0.;
            total = es;
            tl.soft_repairs++;
          } else {
            fatal(c, 101);
          }
        }
        __stcg(kc + 0, nu_cmf);
        __stcg(kc + 1, total);
        __stcg(kc + 2, es);
        __stcg(kc + 3, ff);
        __stcg(kc + 4, bf);
        __stcg(kc + 5, kappa_ff);
        __stcg(kc + 6, (double)mgi);
        tl.handoffs++;
        done = true;
        k2 = KIND_THIN_RESUME;
        have = false;
      }
    } else if (have) {
      tl.packet_updates++;
      int selidx;
      const int action = ma_jump(c, v, rng, tl, &selidx);
      events++;
      if (action >= 0 || events >= event_cap) {
        // the words of sector 0 a jump changes, and the draw counter
        __stcg(&hot[idx].ma_el_ion, pkt_pack2(v.ma_element, v.ma_ion));
        __stcg(&hot[idx].ma_level, v.ma_level);
        __stcg(&hot[idx].interactions, v.interactions);
        __stcg(&hot[idx].rngctr, (unsigned)rng.ctr);
        if (action >= 0) {
          __stcg(pv.pend + idx, ((unsigned long long)(unsigned)(action + 1) << 32) | (unsigned long long)(unsigned)selidx);
          k2 = KIND_MAK;
        } else {
          k2 = KIND_MAJUMP;
        }
        tl.handoffs++;
        done = true;
        have = false;
      }
    }
    __syncwarp(full);
    const unsigned tick2 = (unsigned)clock();
    if (lane == role) cyc_b += (tick2 - tick1) >> 6;
    if (__ballot_sync(full, done) != 0u) {
      push_done(c, pv, done, k2, idx);
      if (lane == 0) cyc_b += ((unsigned)clock() - tick2) >> 6;
    }
  }
  tally_flush(c, tl);
  // a warp may have served both roles: its role cycles are filed under the role it ends in (statistics only)
  if ((lane == KIND_MAJUMP || lane == KIND_KAPPA) && fill_iters > 0) {
    ctr_add(&c.es.counters[ARTIS_CTR_WARP_ITERS0 + lane], fill_iters);
    ctr_add(&c.es.counters[ARTIS_CTR_BUSY_LANES0 + lane], fill_lanes);
    ctr_add(&c.es.counters[ARTIS_CTR_WARP_CYCLES0 + lane], (long long)cyc_b << 6);
  }
  if (lane == 0) {
    ctr_add(&c.es.counters[ARTIS_CTR_REFILL_CYCLES], (long long)cyc_a << 6);
    ctr_add(&c.es.counters[ARTIS_CTR_PUSH_CYCLES], (long long)cyc_b << 6);
    ctr_add(&c.es.counters[ARTIS_CTR_IDLE_CYCLES], (long long)cyc_c << 6);
  }
  stats_block_flush(c);
}

// VPKT_ON: the virtual packets of the requests the scheduler's events have recorded (vpkt.cuh). One request per thread (its
// rays towards the observers one after the other, the start cell's opacity shared between them), or per warp.
#ifndef ARTIS_NO_VPKT
__global__ void __launch_bounds__(128, ARTIS_VPKT_TRACE_MINBLOCKS) k_vpkt_trace(const artis::VpktReq *req, const long long n,
                                                                               unsigned long long *cursor) {
  using namespace artis;
  const Ctx &c = d_ctx;
  Tally tl = {};
  stats_block_begin();
#if ARTIS_VPKT_TRACE_MODE == 2
  vpkt_trace_lockstep(c, req, n, cursor, tl);  // a lane per request, the warp through the phases of a cell together
#else
  constexpr bool COOP = ARTIS_VPKT_TRACE_MODE == 1;  // 1: a warp per request; 0: a thread per request, each at its own pace
  const long long unit = COOP ? 32 : 1;
  const long long stride = (long long)gridDim.x * blockDim.x / unit;
  for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / unit; i < n; i += stride) {
    VpktReq r;
    const uint4 *src = reinterpret_cast<const uint4 *>(req + i);
    uint4 *dst = reinterpret_cast<uint4 *>(&r);
#pragma unroll
    for (int q = 0; q < 8; q++) dst[q] = __ldcg(src + q);
    vpkt_trace_request<COOP>(c, r, tl);
  }
#endif
  tally_flush(c, tl);
  stats_block_flush(c);
}
#endif
__global__ void k_bucket_hist(PktHot *hot, PktCold *cold, int n, int *counts, bool fresh) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    artis::body_bucket_hist(d_ctx, hot, cold, i, counts, fresh);
}
// exclusive prefix sum over the (kind, cell) counts, one 1024-thread block per kind (a 3D model has a model cell per
// propagation cell: 1e6 counts per kind); also the size of each kind's queue
__global__ void __launch_bounds__(1024) k_bucket_scan(int *counts, int *kind_sizes) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  const int per = d_ctx.st.npts_model + 1;
  int *row = counts + (size_t)blockIdx.x * per;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int base = 0; base < per; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < per ? row[i] : 0;
    int x = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    if (w == 0) {
      int t = warp_tot[lane];
      for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, t, d);
        if (lane >= d) t += y;
      }
      warp_tot[lane] = t;
    }
    __syncthreads();
    const int before = carry + (w > 0 ? warp_tot[w - 1] : 0) + x - v;
    if (i < per) row[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) kind_sizes[blockIdx.x] = carry;
}
__global__ void k_bucket_scatter(const PktHot *hot, int n, int *offsets, int *const *queues) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    artis::body_bucket_scatter(d_ctx, hot, i, offsets, queues);
}
__global__ void k_probe_boundary(const ArtisPacket *p, int n, double *a, int *b, int *c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_boundary(d_ctx, p, i, a, b, c);
}
__global__ void k_probe_kappa(const ArtisPacket *p, int n, double *k) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_kappa(d_ctx, p, i, k);
}
__global__ void k_probe_closest(const double *nu, const int *nt, int n, int *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_closest(d_ctx, nu, nt, i, o);
}
__global__ void k_probe_get_event(const ArtisPacket *p, int n, const double *t, const double *ad, double *e, int *et, int *nt,
                                  int *ma) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_get_event(d_ctx, p, i, t, ad, e, et, nt, ma);
}
__global__ void k_probe_ma_rates(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_ma_rates(d_ctx, a, b, c, d, i, o);
}
__global__ void k_probe_cooling(const int *a, const int *b, const int *c, int n, double *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_cooling(d_ctx, a, b, c, i, o);
}
__global__ void k_probe_gamma_kappa(const ArtisPacket *p, int n, double *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_gamma_kappa(d_ctx, p, i, o);
}
__global__ void k_probe_scnu(const int *a, const int *b, const int *c2, const int *d, const float *T, const double *z, int n, double *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_scnu(d_ctx, a, b, c2, d, T, z, i, o);
}
__global__ void k_probe_levelpop(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_levelpop(d_ctx, a, b, c, d, i, o);
}

cudaStream_t g_stream = nullptr;
cudaEvent_t g_ev0 = nullptr, g_ev1 = nullptr;
int g_sms = SM_COUNT_B200;
// resident blocks per SM of the persistent kernels: the single kernel with every role on its own, and the pair
int g_blocks_all = 1, g_blocks_heavy = 1, g_blocks_jump = 0;
int g_scheduler_kernels = 2;  // 2: jump kernel + kernel with every other role side by side; 1: one kernel with every role
cudaStream_t g_stream2 = nullptr;
cudaEvent_t g_ev_fork = nullptr, g_ev_join = nullptr;
bool g_persistent_warmed = false;
unsigned long long g_idle_timeout_ns = 20ull * 1000000000ull;  // hang guard of the persistent kernels (artis_gpu_set_tuning)
double *g_est_block = nullptr;
size_t g_est_ndoubles = 0;
int g_nbfcontinua_ground = 0;
double *g_kcache = nullptr, *g_stepstate = nullptr;
size_t g_kcache_doubles = 0;
std::string g_async_err;

void note(cudaError_t e, const char *what) {
  if (e != cudaSuccess && g_async_err.empty()) g_async_err = std::string(what) + ": " + cudaGetErrorString(e);
}

bool be_available(std::string *why) {
  int n = 0;
  const cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    *why = (e != cudaSuccess) ? cudaGetErrorString(e) : "cudaGetDeviceCount returned 0 devices";
    return false;
  }
  return true;
}
#ifndef ARTIS_HEAVY_BLOCKS
#define ARTIS_HEAVY_BLOCKS 0  // resident blocks per SM of the every-other-role kernel next to the jump kernel; 0: as many as fit with >= 2 jump blocks
#endif
bool be_select_device(int device, std::string *why) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    *why = std::string("cudaSetDevice: ") + cudaGetErrorString(e);
    return false;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) g_sms = prop.multiProcessorCount;
  // Every block of the persistent grids must be resident (idle warps wait for the others' packets): the grids are sized
  // from the registers each kernel was compiled to. Registers are allocated per warp in units of 256 (8 per thread).
  // The two kernels of the pair must be able to share an SM: the same L1 / shared-memory split for both (an SM cannot
  // hold blocks of kernels with different carve-outs -- the second kernel would wait for the first to drain, which never
  // happens here because each waits for the other's packets).
  constexpr int CARVEOUT_PERCENT = 50;  // of the 228 KB: room for the opacity stages of four light blocks, the rest stays L1
  note(cudaFuncSetAttribute(k_transport_persistent<ROLES_ALL>, cudaFuncAttributePreferredSharedMemoryCarveout, CARVEOUT_PERCENT), "carve-out");
  note(cudaFuncSetAttribute(k_transport_persistent<ROLES_HEAVY>, cudaFuncAttributePreferredSharedMemoryCarveout, CARVEOUT_PERCENT), "carve-out");
  note(cudaFuncSetAttribute(k_transport_light, cudaFuncAttributePreferredSharedMemoryCarveout, CARVEOUT_PERCENT), "carve-out");
  note(cudaFuncSetAttribute(k_transport_persistent<ROLES_ALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KAPPA_SMEM_PER_BLOCK), "dynamic smem");
  note(cudaFuncSetAttribute(k_transport_light, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KAPPA_SMEM_PER_BLOCK), "dynamic smem");
  note(cudaFuncSetAttribute(k_transport_persistent<ROLES_HEAVY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KAPPA_SMEM_PER_BLOCK), "dynamic smem");
  {
    cudaFuncAttributes aa{}, ah{}, aj{};
    int regs_per_sm = 65536;
    cudaDeviceGetAttribute(&regs_per_sm, cudaDevAttrMaxRegistersPerMultiprocessor, device);
    auto block_regs = [](const cudaFuncAttributes &a) { return ((a.numRegs + 7) / 8) * 8 * 128; };
    // blocks of the big kernel that the shared-memory carve-out holds (opacity stages + 1 KB the system keeps per block)
    const int smem_blocks = (int)((size_t)CARVEOUT_PERCENT * 228 * 1024 / 100 / (KAPPA_SMEM_PER_BLOCK + 1024));
    g_scheduler_kernels = 1;
    if (cudaFuncGetAttributes(&aa, k_transport_persistent<ROLES_ALL>) == cudaSuccess && aa.numRegs > 0) {
      g_blocks_all = regs_per_sm / block_regs(aa);
      if (g_blocks_all > smem_blocks) g_blocks_all = smem_blocks;
      if (g_blocks_all < 1) g_blocks_all = 1;
    }
    if (cudaFuncGetAttributes(&ah, k_transport_persistent<ROLES_HEAVY>) == cudaSuccess &&
        cudaFuncGetAttributes(&aj, k_transport_light) == cudaSuccess && ah.numRegs > 0 && aj.numRegs > 0) {
      const int heavy = block_regs(ah), jump = block_regs(aj);
      int hb = ARTIS_HEAVY_BLOCKS;
      if (hb <= 0) {  // as many blocks of the big kernel as leave room for at least two blocks (8 warps) of jumps
        hb = (regs_per_sm - 2 * jump) / heavy;
      }
      if (hb < 1) hb = 1;
      int jb = (regs_per_sm - hb * heavy) / jump;
      if (LIGHT_SMEM != 0 && jb > smem_blocks) jb = smem_blocks;  // whichever kernel holds the opacity stages
      if (HEAVY_SMEM != 0 && hb > smem_blocks) hb = smem_blocks;
      if ((hb + jb) * 128 > prop.maxThreadsPerMultiProcessor) jb = prop.maxThreadsPerMultiProcessor / 128 - hb;
      // measurement knob: ARTIS_GEOMETRY="<blocks of the big kernel>,<blocks of the jump kernel>" per SM (must fit the SM)
      if (const char *geo = getenv("ARTIS_GEOMETRY")) {
        int a = 0, b2 = 0;
        if (sscanf(geo, "%d,%d", &a, &b2) == 2 && a >= 1 && b2 >= 1 && a * heavy + b2 * jump <= regs_per_sm &&
            (a + b2) * 128 <= prop.maxThreadsPerMultiProcessor && (LIGHT_SMEM != 0 ? b2 : a) <= smem_blocks) {
          hb = a;
          jb = b2;
        }
      }
      if (jb >= 1) {
        g_blocks_heavy = hb;
        g_blocks_jump = jb;
        g_scheduler_kernels = 2;
      }
    }
  }
  if (g_stream == nullptr) {
    note(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking), "cudaStreamCreate");
    note(cudaStreamCreateWithFlags(&g_stream2, cudaStreamNonBlocking), "cudaStreamCreate");
    note(cudaEventCreate(&g_ev0), "cudaEventCreate");
    note(cudaEventCreate(&g_ev1), "cudaEventCreate");
    note(cudaEventCreateWithFlags(&g_ev_fork, cudaEventDisableTiming), "cudaEventCreate");
    note(cudaEventCreateWithFlags(&g_ev_join, cudaEventDisableTiming), "cudaEventCreate");
  }
  g_async_err.clear();
  return true;
}
void *be_alloc(size_t n) {
  void *p = nullptr;
  if (cudaMalloc(&p, n ? n : 1) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void be_free(void *p) { cudaFree(p); }
// artis_gpu_finalize: the buffers this backend grows on demand (the streams and events stay: a later init reuses them)
void be_release() {
  if (g_kcache) cudaFree(g_kcache);
  if (g_stepstate) cudaFree(g_stepstate);
  g_kcache = nullptr;
  g_stepstate = nullptr;
  g_kcache_doubles = 0;
}
size_t be_free_bytes() {
  size_t fr = 0, tot = 0;
  if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return fr;
}
void be_h2d(void *d, const void *s, size_t n) {
  note(cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, g_stream), "H2D copy");
  note(cudaStreamSynchronize(g_stream), "H2D sync");  // the caller's buffer may be pageable / reused right away
}
void be_d2h(void *d, const void *s, size_t n) {
  note(cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, g_stream), "D2H copy");
  note(cudaStreamSynchronize(g_stream), "D2H sync");
}
void be_zero(void *p, size_t n) { note(cudaMemsetAsync(p, 0, n, g_stream), "memset"); }
void be_set_ctx(const artis::Ctx &c) {
  g_nbfcontinua_ground = c.st.nbfcontinua_ground;
  note(cudaMemcpyToSymbolAsync(d_ctx, &c, sizeof(c), 0, cudaMemcpyHostToDevice, g_stream), "ctx upload");
  note(cudaStreamSynchronize(g_stream), "ctx sync");
}
int be_sync(std::string *why) {
  note(cudaStreamSynchronize(g_stream), "stream sync");
  note(cudaGetLastError(), "kernel launch");
  if (!g_async_err.empty()) {
    *why = g_async_err;
    g_async_err.clear();
    return 1;
  }
  return 0;
}
void be_timer_begin() { note(cudaEventRecord(g_ev0, g_stream), "event record"); }
double be_timer_end_ms() {
  float ms = 0.f;
  note(cudaEventRecord(g_ev1, g_stream), "event record");
  note(cudaEventSynchronize(g_ev1), "event sync");
  note(cudaEventElapsedTime(&ms, g_ev0, g_ev1), "event elapsed");
  return ms;
}
int grid_for(long long n, int block, int blocks_per_sm) {
  long long g = (n + block - 1) / block;
  const long long cap = (long long)g_sms * blocks_per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}
void be_run_levelpop(long long n) { if (n > 0) k_levelpop<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(n); }
void be_run_bf(long long n) { if (n > 0) k_bf<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(n); }
void be_run_ff_zsum(long long n, double *out) { if (n > 0) k_ff_zsum<<<grid_for(n, 128, 8), 128, 0, g_stream>>>(n, out); }
// live continua: per-cell counts -> offsets in off[0 .. ncells] (device); returns the total (synchronises)
long long be_run_bf_live_count(long long ncells, long long *off) {
  if (ncells <= 0) return 0;
  k_bf_live_count<<<(unsigned)ncells, 256, 0, g_stream>>>(off);
  k_scan_exclusive_ll<<<1, 1024, 0, g_stream>>>(off, (int)ncells);
  long long total = 0;
  note(cudaMemcpyAsync(&total, off + ncells, sizeof(total), cudaMemcpyDeviceToHost, g_stream), "D2H");
  note(cudaStreamSynchronize(g_stream), "D2H sync");
  return total;
}
void be_run_bf_live_fill(long long ncells, const long long *off, double *pair, artis::ContRec *rec, int *rank) {
  if (ncells > 0) k_bf_live_fill<<<(unsigned)ncells, 256, 0, g_stream>>>(off, pair, rec, rank);
}
void be_run_corrphoto(long long n) { if (n > 0) k_corrphoto<<<grid_for(n * 32, 128, 8), 128, 0, g_stream>>>(n); }
void be_run_ma(long long n) { if (n > 0) k_ma<<<grid_for(n * 32, 128, 8), 128, 0, g_stream>>>(n); }
void be_run_cooling_rates(long long ncells, int nions, double *contrib, double *total) {
  if (ncells <= 0) return;
  k_cooling_rates_ion<<<grid_for(ncells * nions, 64, 8), 64, 0, g_stream>>>(ncells * nions, contrib);
  k_cooling_rates_total<<<grid_for(ncells, 128, 8), 128, 0, g_stream>>>(ncells, contrib, total);
}
void be_run_cool(long long n) { if (n > 0) k_cool<<<grid_for(n * 32, 128, 8), 128, 0, g_stream>>>(n); }
void be_run_unpack(const ArtisPacket *aos, PktHot *hot, PktCold *cold, long long n) {
  if (n > 0) k_unpack<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(aos, hot, cold, n);
}
void be_run_pack(const PktHot *hot, const PktCold *cold, ArtisPacket *aos, long long n) {
  if (n > 0) k_pack<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(hot, cold, aos, n);
}
void be_run_bin_escaped(const artis::BinSpec &b, const PktHot *hot, const PktCold *cold, long long n) {
  if (n <= 0) return;
  const size_t smem = (size_t)4 * b.ntstep * sizeof(double);
  const int in_smem = smem <= 32768;
  k_bin_escaped<<<grid_for(n, 256, 8), 256, in_smem ? smem : 0, g_stream>>>(b, hot, cold, n, in_smem);
}
void be_run_bin_resolved(const artis::BinSpec &b, const artis::ResSpec &r, const PktHot *hot, const PktCold *cold, long long n, int *error_flag) {
  if (n > 0) k_bin_resolved<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(b, r, hot, cold, n, error_flag);
}
void be_run_normalise(const artis::NormSpec &n) {
  if (n.npts_model > 0) k_normalise<<<grid_for(n.npts_model, 256, 8), 256, 0, g_stream>>>(n);
}
// counts must hold n / ESC_CHUNK + 2 ints; afterwards counts[0..nchunks) are the chunks' output offsets, counts[nchunks] the total
void be_run_escaped_count(const PktHot *hot, long long n, const unsigned char *reported, int only_new, int *counts) {
  const int nchunks = (int)((n + artis::ESC_CHUNK - 1) / artis::ESC_CHUNK);
  if (nchunks > 0) k_escaped_count<<<nchunks, artis::ESC_CHUNK, 0, g_stream>>>(hot, n, reported, only_new, counts);
  k_escaped_scan<<<1, 1024, 0, g_stream>>>(counts, nchunks);
}
void be_run_escaped_scatter(const PktHot *hot, const PktCold *cold, long long n, unsigned char *reported, int only_new,
                            const int *offsets, ArtisPacket *out) {
  const int nchunks = (int)((n + artis::ESC_CHUNK - 1) / artis::ESC_CHUNK);
  if (nchunks > 0) k_escaped_scatter<<<nchunks, artis::ESC_CHUNK, 0, g_stream>>>(hot, cold, n, reported, only_new, offsets, out);
}
void be_run_bucket(PktHot *hot, PktCold *cold, int n, int *buckets, int *kind_sizes, int *const *queues, bool fresh) {
  k_bucket_hist<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(hot, cold, n, buckets, fresh);
  k_bucket_scan<<<artis::NKIND, 1024, 0, g_stream>>>(buckets, kind_sizes);
  k_bucket_scatter<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(hot, n, buckets, queues);
}
// the kernel-rounds scheduler is the host build's (oracle) way of running the same visits; the device has none
void be_run_wavefront(PktHot *, PktCold *, const artis::QueueView &, int, long long, int) {}
// barrier-free scheduler: rings = NRING rings of `cap` ints (initial entries written by be_run_bucket, -1 elsewhere),
// pmeta = head[NRING], tail[NRING], n_active (int), t_start, each on its own RING_META_STRIDE words; pend = [cap] words.
bool be_run_persistent(PktHot *hot, PktCold *cold, int *rings, int cap, unsigned long long *pmeta, unsigned long long *pend,
                       int n_initial, long long maxev) {
  PersistView pv;
  for (int k = 0; k < NRING; k++) pv.ring[k] = rings + (size_t)k * cap;
  pv.head = pmeta;
  pv.tail = pmeta + NRING * RMS;
  pv.n_active = reinterpret_cast<int *>(pmeta + 2 * NRING * RMS);
  pv.n_initial = n_initial;
  pv.t_start = pmeta + (2 * NRING + 1) * RMS;
  pv.partner_started = nullptr;  // set below for the pair
  pv.cap = cap;
  pv.pend = pend;
  // opacity hand-over rows of the two-kernel scheduler (8 scalars + the ground-continuum contributions per packet)
  const int kstride = (8 + g_nbfcontinua_ground + 1) & ~1;
  if ((size_t)cap * kstride > g_kcache_doubles) {
    if (g_kcache) cudaFree(g_kcache);
    if (g_stepstate) cudaFree(g_stepstate);
    g_kcache = nullptr;
    g_stepstate = nullptr;
    g_kcache_doubles = 0;
    if (cudaMalloc(&g_kcache, (size_t)cap * kstride * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&g_stepstate, (size_t)cap * 3 * sizeof(double)) != cudaSuccess) {
      note(cudaGetLastError(), "cudaMalloc (opacity hand-over buffers)");
      return true;  // be_sync reports the error
    }
    g_kcache_doubles = (size_t)cap * kstride;
  }
  pv.kcache = g_kcache;
  pv.stepstate = g_stepstate;
  pv.kstride = kstride;
  const unsigned long long timeout = g_idle_timeout_ns;
  if (!g_persistent_warmed) {
    // First launch of each persistent kernel with nothing to do (n_active == 0: every warp leaves at once). A first
    // launch can make the driver grow the local-memory pool, which waits for running kernels -- with the other
    // persistent kernel already spinning that would never end.
    unsigned long long *scratch = nullptr;
    if (cudaMalloc(&scratch, artis::RING_META_WORDS * sizeof(unsigned long long)) == cudaSuccess) {
      note(cudaMemsetAsync(scratch, 0, artis::RING_META_WORDS * sizeof(unsigned long long), g_stream), "memset");
      PersistView w = pv;
      w.head = scratch;
      w.tail = scratch + NRING * RMS;
      w.n_active = reinterpret_cast<int *>(scratch + 2 * NRING * RMS);
      w.t_start = scratch + (2 * NRING + 1) * RMS;
      w.partner_started = nullptr;
      w.n_initial = 0;
      k_transport_persistent<ROLES_ALL><<<g_sms * g_blocks_all, 128, KAPPA_SMEM_PER_BLOCK, g_stream>>>(hot, cold, w, maxev, timeout);
      k_transport_persistent<ROLES_HEAVY><<<g_sms * g_blocks_heavy, 128, HEAVY_SMEM, g_stream>>>(hot, cold, w, maxev, timeout);
      if (g_blocks_jump > 0) k_transport_light<<<g_sms * g_blocks_jump, 128, LIGHT_SMEM, g_stream>>>(hot, w, maxev, timeout);
      note(cudaStreamSynchronize(g_stream), "warm-up sync");
      cudaFree(scratch);
    }
    g_persistent_warmed = true;
  }
  if (g_scheduler_kernels != 2) {
    k_transport_persistent<ROLES_ALL><<<g_sms * g_blocks_all, 128, KAPPA_SMEM_PER_BLOCK, g_stream>>>(hot, cold, pv, maxev, timeout);
    return true;
  }
  // Both kernels must be resident on every SM at once. The kernel that needs shared memory (the opacity stages) is
  // launched first, so the L1 / shared-memory split it brings to an SM also fits the blocks of the other kernel, which
  // need next to none -- launched the other way round, the split chosen for the first kernel leaves no room for the
  // second one's blocks and they wait for the first kernel to drain, which it never does (each waits for the other's
  // packets). The side stream waits for what is queued on the main stream (ring setup), the main stream for the side kernel.
  pv.partner_started = reinterpret_cast<int *>(pmeta + (2 * NRING + 2) * RMS);
  note(cudaEventRecord(g_ev_fork, g_stream), "event record");
  note(cudaStreamWaitEvent(g_stream2, g_ev_fork, 0), "stream wait");
  if (LIGHT_SMEM != 0) k_transport_light<<<g_sms * g_blocks_jump, 128, LIGHT_SMEM, g_stream2>>>(hot, pv, maxev, timeout);
  k_transport_persistent<ROLES_HEAVY><<<g_sms * g_blocks_heavy, 128, HEAVY_SMEM, g_stream>>>(hot, cold, pv, maxev, timeout);
  if (LIGHT_SMEM == 0) k_transport_light<<<g_sms * g_blocks_jump, 128, LIGHT_SMEM, g_stream2>>>(hot, pv, maxev, timeout);
  note(cudaEventRecord(g_ev_join, g_stream2), "event record");
  note(cudaStreamWaitEvent(g_stream, g_ev_join, 0), "stream wait");
  return true;
}
void be_fill_bytes(void *p, int byte, size_t n) { note(cudaMemsetAsync(p, byte, n, g_stream), "memset"); }
// VPKT_ON: trace the n recorded requests (on the transport stream, inside the timed region of the timestep)
void be_run_vpkt_trace(const artis::VpktReq *req, long long n, unsigned long long *cursor) {
#ifndef ARTIS_NO_VPKT
  if (n <= 0) return;
  const long long unit = ARTIS_VPKT_TRACE_MODE == 1 ? 32 : 1;
  long long blocks = (n * unit + 127) / 128;
  const long long most = (long long)g_sms * (ARTIS_VPKT_TRACE_MODE == 2 ? ARTIS_VPKT_TRACE_MINBLOCKS : 16);  // grid-stride beyond that
  if (blocks > most) blocks = most;
  k_vpkt_trace<<<(unsigned)blocks, 128, 0, g_stream>>>(req, n, cursor);
#endif
}
bool be_vpkt_deferred() {
#if !defined(ARTIS_NO_VPKT) && ARTIS_VPKT_DEFER
  return true;
#else
  return false;
#endif
}
void be_set_scheduler_kernels(int n) {
  if (n == 1 || (n == 2 && g_blocks_jump >= 1)) g_scheduler_kernels = n;
}
int be_persistent_kernels() { return g_scheduler_kernels == 2 ? 2 : 1; }
// threads the scheduler keeps resident (bound of the rows that can be under construction at once)
long long be_resident_lanes() {
  const int per_sm = g_blocks_all > g_blocks_heavy + g_blocks_jump ? g_blocks_all : g_blocks_heavy + g_blocks_jump;
  return (long long)g_sms * per_sm * 128;
}
int be_scheduler_geometry(int *out) {  // kernels, blocks per SM of: single kernel, big kernel, jump kernel
  out[0] = g_scheduler_kernels;
  out[1] = g_blocks_all;
  out[2] = g_blocks_heavy;
  out[3] = g_blocks_jump;
  return 4;
}
void be_set_idle_timeout_ms(long long ms) { g_idle_timeout_ns = (unsigned long long)ms * 1000000ull; }
void be_set_estimator_block(double *block, size_t ndoubles) {
  g_est_block = block;
  g_est_ndoubles = ndoubles;
}
#define PROBE_GRID(n) ((n) + 127) / 128, 128, 0, g_stream
void be_run_probe_boundary(const ArtisPacket *p, int n, double *a, int *b, int *c) { k_probe_boundary<<<PROBE_GRID(n)>>>(p, n, a, b, c); }
void be_run_probe_kappa(const ArtisPacket *p, int n, double *k) { k_probe_kappa<<<PROBE_GRID(n)>>>(p, n, k); }
void be_run_probe_closest(const double *nu, const int *nt, int n, int *o) { k_probe_closest<<<PROBE_GRID(n)>>>(nu, nt, n, o); }
void be_run_probe_get_event(const ArtisPacket *p, int n, const double *t, const double *ad, double *e, int *et, int *nt, int *ma) {
  k_probe_get_event<<<PROBE_GRID(n)>>>(p, n, t, ad, e, et, nt, ma);
}
void be_run_probe_ma_rates(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  k_probe_ma_rates<<<PROBE_GRID(n)>>>(a, b, c, d, n, o);
}
void be_run_probe_cooling(const int *a, const int *b, const int *c, int n, double *o) { k_probe_cooling<<<PROBE_GRID(n)>>>(a, b, c, n, o); }
void be_run_probe_gamma_kappa(const ArtisPacket *p, int n, double *o) { k_probe_gamma_kappa<<<PROBE_GRID(n)>>>(p, n, o); }
void be_run_probe_scnu(const int *a, const int *b, const int *c2, const int *d, const float *T, const double *z, int n, double *o) {
  k_probe_scnu<<<PROBE_GRID(n)>>>(a, b, c2, d, T, z, n, o);
}
void be_run_probe_levelpop(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  k_probe_levelpop<<<PROBE_GRID(n)>>>(a, b, c, d, n, o);
}

// ---- NCCL (loaded at run time so that single-GPU use needs no libnccl) ---------------------------------------------------
struct Nccl {
  void *h = nullptr;
  void *comm = nullptr;
  int nranks = 1;
  int (*GetUniqueId)(void *) = nullptr;
  int (*CommInitRank)(void **, int, char[128], int) = nullptr;  // ncclUniqueId is a 128-byte struct passed by value
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
} g_nccl;
struct NcclId {
  char b[128];
};
bool nccl_load(std::string *err) {
  if (g_nccl.h) return true;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char *nm : names) {
    g_nccl.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.h) break;
  }
  if (!g_nccl.h) {
    *err = "cannot dlopen libnccl.so.2";
    return false;
  }
  g_nccl.GetUniqueId = (int (*)(void *))dlsym(g_nccl.h, "ncclGetUniqueId");
  g_nccl.AllReduce = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t))dlsym(g_nccl.h, "ncclAllReduce");
  g_nccl.GetErrorString = (const char *(*)(int))dlsym(g_nccl.h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.AllReduce || !dlsym(g_nccl.h, "ncclCommInitRank")) {
    *err = "libnccl lacks the expected symbols";
    return false;
  }
  return true;
}
int be_comm_unique_id(char *id_out, std::string *err) {
  if (!nccl_load(err)) return ARTIS_ERR_NCCL;
  const int rc = g_nccl.GetUniqueId(id_out);
  if (rc != 0) {
    *err = std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(rc);
    return ARTIS_ERR_NCCL;
  }
  return ARTIS_OK;
}
int be_comm_init(const char *id, int rank, int nranks, std::string *err) {
  if (!nccl_load(err)) return ARTIS_ERR_NCCL;
  NcclId nid;
  memcpy(nid.b, id, 128);
  auto init = (int (*)(void **, int, NcclId, int))dlsym(g_nccl.h, "ncclCommInitRank");
  const int rc = init(&g_nccl.comm, nranks, nid, rank);
  if (rc != 0) {
    *err = std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(rc);
    return ARTIS_ERR_NCCL;
  }
  g_nccl.nranks = nranks;
  return ARTIS_OK;
}
// the exchange step of a timestep: sum all-reduces over the FP64 estimator block, the int64 counters and the int32 line
// counters as ONE NCCL group (a single launch), issued on the transport stream right behind the last transport kernel;
// its device time is kept for artis_gpu_last_timings()[5]
double g_allreduce_ms = 0.;
int be_comm_allreduce(double *d, size_t nd, long long *l, size_t nl, int *i, size_t ni, std::string *err) {
  g_allreduce_ms = 0.;
  if (g_nccl.comm == nullptr || g_nccl.nranks <= 1) return ARTIS_OK;  // single shard: nothing to reduce
  const int ncclSum = 0, ncclInt32 = 2, ncclInt64 = 4, ncclFloat64 = 8;
  auto group_start = (int (*)())dlsym(g_nccl.h, "ncclGroupStart");
  auto group_end = (int (*)())dlsym(g_nccl.h, "ncclGroupEnd");
  note(cudaEventRecord(g_ev0, g_stream), "event record");
  int rc = group_start ? group_start() : 0;
  if (rc == 0) rc = g_nccl.AllReduce(d, d, nd, ncclFloat64, ncclSum, g_nccl.comm, g_stream);
  if (rc == 0 && nl > 0) rc = g_nccl.AllReduce(l, l, nl, ncclInt64, ncclSum, g_nccl.comm, g_stream);
  if (rc == 0 && ni > 0) rc = g_nccl.AllReduce(i, i, ni, ncclInt32, ncclSum, g_nccl.comm, g_stream);
  if (group_end) {
    const int rc2 = group_end();
    if (rc == 0) rc = rc2;
  }
  if (rc != 0) {
    *err = std::string("ncclAllReduce: ") + g_nccl.GetErrorString(rc);
    return ARTIS_ERR_NCCL;
  }
  float ms = 0.f;
  note(cudaEventRecord(g_ev1, g_stream), "event record");
  note(cudaEventSynchronize(g_ev1), "allreduce sync");
  note(cudaEventElapsedTime(&ms, g_ev0, g_ev1), "event elapsed");
  g_allreduce_ms = ms;
  return ARTIS_OK;
}
double be_last_allreduce_ms() { return g_allreduce_ms; }

}  // namespace

#include "engine.inc"

// ---- escaped-packet compaction without any allocation -------------------------------------------------------------------
// artis_gpu_pack_escaped_noalloc (include/artis_gpu.h): for callers that must not reach cudaMalloc (a CUDA graph being
// captured, a serving loop, a harness that forbids it). Same three steps as artis_gpu_packets_download_escaped -- per-chunk
// counts, one scan, cooperative 304-byte copies -- but on the caller's own device arrays and with the per-chunk counts in
// a file-scope device array of fixed size instead of a cudaMalloc'ed one. Every GPU that loads the library holds that
// array until the process exits; every call shares it, so calls that overlap on one GPU race on it.
constexpr int NOALLOC_MAX_PACKETS = 1 << 22;  // 4 194 304 packets per call
constexpr int NOALLOC_MAX_CHUNKS = NOALLOC_MAX_PACKETS / artis::ESC_CHUNK;
static __device__ __align__(256) int g_noalloc_counts[NOALLOC_MAX_CHUNKS + 2];  // 64 KiB + 8 B per GPU

namespace {
// Host code may not pass g_noalloc_counts itself to a kernel (there the name is a host-side stand-in, not a device
// address): the address is asked for once per device, with that device current.
constexpr int NOALLOC_MAX_DEVICES = 64;
int *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

__global__ void __launch_bounds__(artis::ESC_CHUNK) k_na_count(const ArtisPacket *pkts, int n, int *counts) {
  const int i = blockIdx.x * artis::ESC_CHUNK + threadIdx.x;
  const int sel = i < n && pkts[i].type == ARTIS_TYPE_ESCAPE;
  const int c = __syncthreads_count(sel);
  if (threadIdx.x == 0) counts[blockIdx.x] = c;
}
// one block: exclusive scan of counts[0..nchunks) in place, counts[nchunks] = total, which is also written to *n_out
__global__ void __launch_bounds__(1024) k_na_scan(int *counts, int nchunks, int *n_out) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int base = 0; base < nchunks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < nchunks ? counts[i] : 0;
    int x = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) warp_tot[w] = x;
    __syncthreads();
    if (w == 0) {
      int t = warp_tot[lane];
      for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, t, d);
        if (lane >= d) t += y;
      }
      warp_tot[lane] = t;
    }
    __syncthreads();
    const int before = carry + (w > 0 ? warp_tot[w - 1] : 0) + x - v;
    if (i < nchunks) counts[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    counts[nchunks] = carry;
    *n_out = carry;
  }
}
__global__ void __launch_bounds__(artis::ESC_CHUNK) k_na_scatter(const ArtisPacket *pkts, int n, const int *offsets, ArtisPacket *out,
                                                                int capacity) {
  __shared__ int warp_base[artis::ESC_CHUNK / 32];
  const int i = blockIdx.x * artis::ESC_CHUNK + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const bool sel = i < n && pkts[i].type == ARTIS_TYPE_ESCAPE;
  const unsigned m = __ballot_sync(0xffffffffu, sel);
  if (lane == 0) warp_base[w] = __popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = offsets[blockIdx.x];
    for (int k = 0; k < artis::ESC_CHUNK / 32; k++) {
      const int c = warp_base[k];
      warp_base[k] = run;
      run += c;
    }
  }
  __syncthreads();
  const int warp_first = i - lane;
  int rank = 0;
  for (unsigned rest = m; rest; rest &= rest - 1, rank++) {
    const int src_lane = __ffs(rest) - 1;
    const int at = warp_base[w] + rank;
    if (at >= capacity) break;  // the caller's buffer is full: *n_out still reports how many there are
    const uint4 *src = reinterpret_cast<const uint4 *>(pkts + warp_first + src_lane);
    uint4 *dst = reinterpret_cast<uint4 *>(out + at);
    if (lane < (int)(sizeof(ArtisPacket) / 16)) dst[lane] = src[lane];
  }
}
}  // namespace

extern "C" int artis_gpu_pack_escaped_noalloc(const ArtisPacket *dev_packets, int npkts, ArtisPacket *dev_out, int capacity,
                                              int *dev_n_out, void *cuda_stream) {
  if (npkts < 0 || capacity < 0 || dev_n_out == nullptr || (npkts > 0 && dev_packets == nullptr) ||
      (capacity > 0 && dev_out == nullptr))
    return fail(ARTIS_ERR_ABI, "artis_gpu_pack_escaped_noalloc: bad arguments");
  if (npkts > NOALLOC_MAX_PACKETS)
    return fail(ARTIS_ERR_ABI, "artis_gpu_pack_escaped_noalloc: " + std::to_string(npkts) + " packets exceed the fixed scratch of " +
                                   std::to_string(NOALLOC_MAX_PACKETS) + " packets per call (split the array, or use "
                                   "artis_gpu_packets_download_escaped, which allocates)");
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NOALLOC_MAX_DEVICES)
    return fail(ARTIS_ERR_CUDA, "artis_gpu_pack_escaped_noalloc: no current CUDA device");
  if (g_noalloc_ptr[dev] == nullptr) {  // set-up, once per device
    int *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_counts);
    if (e != cudaSuccess || ptr == nullptr)
      return fail(ARTIS_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_counts): ") + cudaGetErrorString(e));
    g_noalloc_ptr[dev] = ptr;
  }
  int *counts = g_noalloc_ptr[dev];
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  const int nchunks = (npkts + artis::ESC_CHUNK - 1) / artis::ESC_CHUNK;
  // nothing in the scratch has to start at zero: counts[0..nchunks] are all written before they are read, whatever the
  // previous call left there
  if (nchunks > 0) k_na_count<<<nchunks, artis::ESC_CHUNK, 0, stream>>>(dev_packets, npkts, counts);
  k_na_scan<<<1, 1024, 0, stream>>>(counts, nchunks, dev_n_out);
  if (nchunks > 0 && capacity > 0) k_na_scatter<<<nchunks, artis::ESC_CHUNK, 0, stream>>>(dev_packets, npkts, counts, dev_out, capacity);
  E.launches += (nchunks > 0 ? 2 : 0) + 1;
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(ARTIS_ERR_CUDA, std::string("artis_gpu_pack_escaped_noalloc launch: ") + cudaGetErrorString(e));
  return ARTIS_OK;
}
