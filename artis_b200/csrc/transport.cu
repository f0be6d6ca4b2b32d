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

#include <string>

#include "kernels.cuh"

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
__global__ void k_ma(long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_ma(d_ctx, i);
}
__global__ void k_cool(long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    artis::body_cool(d_ctx, i);
}
// one kernel per event kind, each compiled with only that kind's physics
template <int KIND>
__global__ void __launch_bounds__(128) k_wavefront(ArtisPacket *pkts, unsigned long long *rngctr, artis::QueueView qv,
                                                   long long max_events) {
  artis::est_private_begin(nullptr, 0u);
  artis::body_wavefront<KIND>(d_ctx, pkts, rngctr, qv, max_events);
}
// ---- barrier-free scheduler ------------------------------------------------------------------------------------------
// One persistent kernel, one ring of packet indices per event kind, warps specialised by kind. A warp that has no
// packet left picks the kind with the longest backlog and keeps that role until it runs dry; its lanes claim ring
// entries (warp-aggregated), advance their packet by events of that kind only and, when the packet changes kind, write
// it back and push its index on the ring of the new kind. Nothing waits for the slowest packet of a round: the step
// ends when the longest single packet history ends, and a warp never executes two kinds' code in one iteration.
//
// Ring protocol (capacity >= number of packets, so at most `cap` entries are ever outstanding): a slot holds -1 when
// empty. Producers take a ticket from tail (atomicAdd) and CAS their index into slot ticket % cap; consumers CAS head
// forward only over tickets below the tail they observed and then exchange the slot back to -1, spinning for the few
// cycles a producer may still need to fill it. Packet state written before the push is made visible with
// __threadfence() and read with ld.global.cg on the consumer side (L1 is not coherent across SMs).
struct PersistView {
  int *ring[artis::NKIND];
  unsigned long long *head;  // [NKIND]
  unsigned long long *tail;  // [NKIND]
  int *n_active;             // packets that have not finished the timestep
  int n_initial;             // its value at launch
  unsigned long long *t_start;  // globaltimer at launch (written by thread 0 of block 0)
  int cap;
  double *kcache;            // [cap][kstride] continuum-opacity cache of a packet handed over in mid-visit
  int *kflag;                // [cap] 1 when kcache holds that packet's cache
  int kstride;
  double *est_block;         // global estimator block and the length of its block-private copy (0 = none)
  unsigned est_private_doubles;
};
// A warp whose ring has run dry and that is left with at most this many busy lanes hands those packets back to the ring
// (to be absorbed by the idle lanes of other warps of the same kind) and takes up a kind that has work for all 32 lanes.
#ifndef ARTIS_CONSOLIDATE_MAX_BUSY
#define ARTIS_CONSOLIDATE_MAX_BUSY -1  /* off: measured no gain on the bench workload (profiles/r1_notes.md) */
#endif
constexpr int CONSOLIDATE_MAX_BUSY = ARTIS_CONSOLIDATE_MAX_BUSY;
constexpr long long CONSOLIDATE_MIN_BACKLOG = 64;

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p) {
  return *reinterpret_cast<const volatile unsigned long long *>(p);
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
static_assert(sizeof(ArtisPacket) % 8 == 0, "packet copied as 8-byte words");
__device__ __forceinline__ void packet_load_cg(ArtisPacket &p, const ArtisPacket *src) {
  const unsigned long long *s = reinterpret_cast<const unsigned long long *>(src);
  unsigned long long *d = reinterpret_cast<unsigned long long *>(&p);
#pragma unroll
  for (int i = 0; i < (int)(sizeof(ArtisPacket) / 8); i++) d[i] = __ldcg(s + i);
}
__device__ __forceinline__ void packet_store_cg(ArtisPacket *dst, const ArtisPacket &p) {
  unsigned long long *d = reinterpret_cast<unsigned long long *>(dst);
  const unsigned long long *s = reinterpret_cast<const unsigned long long *>(&p);
#pragma unroll
  for (int i = 0; i < (int)(sizeof(ArtisPacket) / 8); i++) __stcg(d + i, s[i]);
}

// Warp-cooperative continuum opacity (K2): every lane that is about to need kappa_cont for its own packet and misses its
// cache ("owner") gets its bound-free sum evaluated by all 32 lanes, a tile of 32 consecutive continua at a time --
// coalesced reads of the per-cell table rows, no dependence between tiles, and the step's latency no longer grows with
// the number of continua below the packet's frequency (up to all 3204 for a UV packet, on one lane before). The result
// is left in the owner's cache exactly as calculate_kappa_rpkt_cont would leave it, so the unchanged get_event that
// follows finds a hit. Summation order = rpkt.cuh BF_TILE trees, the same as the host restatement.
static_assert(artis::BF_TILE == 32, "one continuum per lane");
constexpr int COOP_UNROLL = 4;
#ifndef ARTIS_COOP_MAX_OWNERS
#define ARTIS_COOP_MAX_OWNERS 8
#endif
constexpr int COOP_MAX_OWNERS = ARTIS_COOP_MAX_OWNERS;
__device__ __forceinline__ void coop_kappa_warm(const artis::Ctx &c, const bool wants, const int mgi, const double nu_cmf,
                                                artis::KappaCache &k, artis::Tally &tl) {
  using namespace artis;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const bool need = wants && c.st.do_r_lc && c.st.opacity_case == 4 && !kappa_cache_hit(k, mgi, nu_cmf);
  unsigned owners = __ballot_sync(full, need);
  // With many owners the per-lane evaluation inside get_event has the better throughput (32 sums advance at once);
  // the cooperative form is for sparsely filled warps, where it cuts the latency of the step (the tail of a timestep).
  if (owners == 0u || __popc(owners) > COOP_MAX_OWNERS) return;
  double sigma = 0., kappa_ff = 0., kappa_bf = 0.;
  if (need) {
    tl.kappa_evals++;
    sigma = SIGMA_T * c.cs.nne[mgi];
    kappa_ff = calculate_kappa_ff(c, mgi, nu_cmf);
    const int ng = c.st.nbfcontinua_ground;
    for (int g = 0; g < ng; g++) k.gamma_contr[g] = 0.;
  }
  const int nbf = c.st.nbfcontinua;
  while (owners != 0u) {
    const int owner = __ffs(owners) - 1;
    owners &= owners - 1u;
    const double nu = __shfl_sync(full, nu_cmf, owner);
    const int omgi = __shfl_sync(full, mgi, owner);
    const int tc = tabcell(c, omgi);
    const double expfac = exp(-HOVERKB * nu / (double)c.cs.Te[omgi]);
    const int i0 = bf_stop_index(c, nu);
    const int ilo = bf_start_index(c, nu, i0);
    double sum = 0.;
    for (int base = ilo; base < i0; base += BF_TILE * COOP_UNROLL) {
      // COOP_UNROLL tiles in flight per lane: their loads overlap, their trees interleave; tile sums are still added
      // in tile order
      double t[COOP_UNROLL], gcontr[COOP_UNROLL];
      int slot[COOP_UNROLL];
#pragma unroll
      for (int u = 0; u < COOP_UNROLL; u++) {
        const int i = base + u * BF_TILE + lane;
        t[u] = 0.;
        gcontr[u] = 0.;
        slot[u] = -1;
        if (i < i0) {
          int sl;
          t[u] = kappa_bf_term(c, tc, i, nu, expfac, &gcontr[u], &sl);
          if (gcontr[u] != 0.) slot[u] = sl;
        }
      }
#pragma unroll
      for (int off = 1; off < BF_TILE; off <<= 1) {  // adjacent pairs first: the tree of rpkt.cuh PairTree
#pragma unroll
        for (int u = 0; u < COOP_UNROLL; u++) t[u] += __shfl_down_sync(full, t[u], off);
      }
#pragma unroll
      for (int u = 0; u < COOP_UNROLL; u++) {
        sum += t[u];  // lane 0 holds each tile sum in tree order; only lane 0's running sum is used
        unsigned g = __ballot_sync(full, slot[u] >= 0);
        while (g != 0u) {  // ground-continuum contributions, in index order like the serial loop
          const int src = __ffs(g) - 1;
          g &= g - 1u;
          const double v = __shfl_sync(full, gcontr[u], src);
          const int sl = __shfl_sync(full, slot[u], src);
          if (lane == owner) k.gamma_contr[sl] += v;
        }
      }
    }
    sum = __shfl_sync(full, sum, 0);
    if (lane == owner) {
      kappa_bf = sum;
      tl.bf_evaluated += i0 - ilo;
    }
  }
  (void)nbf;
  if (need) kappa_store(c, k, mgi, nu_cmf, sigma, kappa_ff, kappa_bf, kappa_ff, tl);
}

#ifndef ARTIS_ROLE_AFFINITY
#define ARTIS_ROLE_AFFINITY 0
#endif
#ifndef ARTIS_PERSIST_MINBLOCKS
#define ARTIS_PERSIST_MINBLOCKS 2
#endif
#ifndef ARTIS_EST_PRIVATE_MAX_KB
#define ARTIS_EST_PRIVATE_MAX_KB 0  /* off: the shared-memory copy costs more L1 than the atomics it saves (profiles/r1_notes.md) */
#endif
// about 250 registers per thread: 8 warps per SM, as two blocks of 4 or (with role affinity) one block of 8
constexpr int PERSIST_THREADS = ARTIS_ROLE_AFFINITY ? 256 : 128;
__global__ void __launch_bounds__(PERSIST_THREADS, ARTIS_ROLE_AFFINITY ? 1 : ARTIS_PERSIST_MINBLOCKS) k_transport_persistent(ArtisPacket *pkts, unsigned long long *rngctr, PersistView pv,
                                                                 long long max_events, unsigned long long idle_timeout_ns) {
  using namespace artis;
  const Ctx &c = d_ctx;
  const double t2 = c.t2;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  Tally tl = {};
  KappaCache k;
  k.valid = false;
  bool have = false;
  int idx = 0, role = 0;
  long long events = 0;
  ArtisPacket p;
  Rng rng;
  unsigned long long idle_since = 0ull;
  est_private_begin(pv.est_block, pv.est_private_doubles);
  if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long *>(pv.t_start) = global_ns();
#if ARTIS_ROLE_AFFINITY
  __shared__ int s_role;
  if (threadIdx.x == 0) s_role = -1;
  __syncthreads();
#endif
  long long fill_iters = 0, fill_lanes = 0;  // iterations run in kind `lane`, and the busy lanes summed over them
  while (true) {
    __syncwarp(full);
#ifdef ARTIS_ROLE_CYCLES
    const long long tr0 = clock64();
#endif
    const unsigned need = __ballot_sync(full, !have);
    if (need != 0u) {
      if (need == full) {  // the whole warp is out of work: choose the kind with the longest backlog
        long long backlog = 0;
        if (lane < NKIND) {
          const unsigned long long h = ld_volatile_u64(pv.head + lane), t = ld_volatile_u64(pv.tail + lane);
          backlog = t > h ? (long long)(t - h) : 0;
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
        // Soft role affinity: the warps of a block (one block per SM) prefer the kind their neighbours are running as
        // long as its ring has at least a warp's worth of work, so that an SM's instruction caches hold one kind's code.
#if ARTIS_ROLE_AFFINITY
        {
          const int pref = *reinterpret_cast<volatile int *>(&s_role);
          const long long pref_backlog = __shfl_sync(full, backlog, pref >= 0 ? pref : 0);
          if (pref >= 0 && pref_backlog >= 32 && best > 0) {
            best_kind = pref;
            best = pref_backlog;
          } else if (best > 0 && lane == 0) {
            *reinterpret_cast<volatile int *>(&s_role) = best_kind;
          }
        }
#endif
        if (best <= 0) {
          int left = 0;
          if (lane == 0) left = *reinterpret_cast<volatile int *>(pv.n_active);
          left = __shfl_sync(full, left, 0);
          if (left <= 0) break;  // every packet has reached the end of the timestep
          unsigned long long now = 0ull;
          if (lane == 0) now = global_ns();
          now = __shfl_sync(full, now, 0);
          if (idle_since == 0ull) idle_since = now;
          if (now - idle_since > idle_timeout_ns) {  // a lost packet would otherwise hang the device
            if (lane == 0 && c.error_flag) atomicCAS(c.error_flag, 0, 900);
            break;
          }
#ifdef ARTIS_ROLE_CYCLES
          const long long ti0 = clock64();
#endif
          __nanosleep(1000);
#ifdef ARTIS_ROLE_CYCLES
          (void)ti0;
#endif
          continue;
        }
        role = best_kind;
        idle_since = 0ull;
      }
      int take = 0;
      unsigned long long base = 0ull;
      if (lane == 0) {
        const int want = __popc(need);
        for (int tries = 0; tries < 4; tries++) {
          const unsigned long long h = ld_volatile_u64(pv.head + role), t = ld_volatile_u64(pv.tail + role);
          const long long avail = t > h ? (long long)(t - h) : 0;
          take = avail < want ? (int)avail : want;
          if (take == 0) break;
          if (atomicCAS(pv.head + role, h, h + (unsigned long long)take) == h) {
            base = h;
            break;
          }
          take = 0;
        }
      }
      take = __shfl_sync(full, take, 0);
      base = __shfl_sync(full, base, 0);
      const int rank = __popc(need & lt_mask);
      if (!have && rank < take) {
        int *slot = pv.ring[role] + (size_t)((base + (unsigned long long)rank) % (unsigned long long)pv.cap);
        int v;
        while ((v = atomicExch(slot, -1)) < 0) {
        }
        __threadfence();
        idx = v;
        packet_load_cg(p, pkts + idx);
        rng = rng_init(c, p.number, __ldcg(rngctr + idx));
        k.valid = false;
        if (CONSOLIDATE_MAX_BUSY >= 0 && role == KIND_RPKT_THIN && __ldcg(pv.kflag + idx) != 0) {  // handed over in mid-visit: its cache comes along
          const double *kc = pv.kcache + (size_t)idx * pv.kstride;
          k.nu = __ldcg(kc + 0);
          k.total = __ldcg(kc + 1);
          k.es = __ldcg(kc + 2);
          k.ff = __ldcg(kc + 3);
          k.bf = __ldcg(kc + 4);
          k.ffheating = __ldcg(kc + 5);
          k.mgi = (int)__ldcg(kc + 6);
          k.valid = true;
          const int ng = c.st.nbfcontinua_ground;
          for (int g = 0; g < ng; g++) k.gamma_contr[g] = __ldcg(kc + 8 + g);
          __stcg(pv.kflag + idx, 0);
        }
        events = 0;
        have = true;
      }
      if (need == full && take == 0) continue;  // lost the race for the last entries
      if (CONSOLIDATE_MAX_BUSY >= 0 && take == 0 && __popc(~need) <= CONSOLIDATE_MAX_BUSY) {
        long long other = 0;
        if (lane < NKIND && lane != role) {
          const unsigned long long h = ld_volatile_u64(pv.head + lane), t = ld_volatile_u64(pv.tail + lane);
          other = t > h ? (long long)(t - h) : 0;
        }
#pragma unroll
        for (int off = 4; off >= 1; off >>= 1) {
          const long long o = __shfl_down_sync(full, other, off);
          other = o > other ? o : other;
        }
        other = __shfl_sync(full, other, 0);
        if (other >= CONSOLIDATE_MIN_BACKLOG) {
          if (have) {
            __stcg(rngctr + idx, (unsigned long long)rng.ctr);
            packet_store_cg(pkts + idx, p);
            if (role == KIND_RPKT_THIN && k.valid) {
              double *kc = pv.kcache + (size_t)idx * pv.kstride;
              __stcg(kc + 0, k.nu);
              __stcg(kc + 1, k.total);
              __stcg(kc + 2, k.es);
              __stcg(kc + 3, k.ff);
              __stcg(kc + 4, k.bf);
              __stcg(kc + 5, k.ffheating);
              __stcg(kc + 6, (double)k.mgi);
              const int ng = c.st.nbfcontinua_ground;
              for (int g = 0; g < ng; g++) __stcg(kc + 8 + g, k.gamma_contr[g]);
              __stcg(pv.kflag + idx, 1);
            }
            __threadfence();
          }
          const unsigned m = __ballot_sync(full, have);
          unsigned long long b = 0ull;
          if (lane == 0) b = atomicAdd(pv.tail + role, (unsigned long long)__popc(m));
          b = __shfl_sync(full, b, 0);
          if (have) {
            int *slot = pv.ring[role] + (size_t)((b + (unsigned long long)__popc(m & lt_mask)) % (unsigned long long)pv.cap);
            while (atomicCAS(slot, -1, idx) != -1) {
            }
            have = false;
            tl.handoffs++;
          }
          continue;  // the warp is idle now and picks the kind with the longest backlog
        }
      }
    }
    __syncwarp(full);
    bool done = false;
    int k2 = -1;
    {
      const unsigned busy = __ballot_sync(full, have);
      if (lane == role) {  // lane k keeps the occupancy statistics of kind k
        fill_iters++;
#ifndef ARTIS_ROLE_CYCLES
        fill_lanes += __popc(busy);
#endif
      }
    }
#ifdef ARTIS_ROLE_CYCLES
    const long long tc0 = clock64();
#endif
#if defined(ARTIS_ROLE_CYCLES) && !defined(ARTIS_ROLE_CYCLES2)
    if (lane == 3) {  // experiment: kind-3 slots hold the scheduling overhead (refill + previous push)
      fill_lanes += tc0 - tr0;
      fill_iters++;
    }
#endif
#ifndef ARTIS_COOP_KAPPA
#define ARTIS_COOP_KAPPA 1
#endif
    if (ARTIS_COOP_KAPPA && role == KIND_RPKT_THIN) {  // warp-uniform: the step in two halves around the cooperative opacity evaluation
      RpktStepState st;
      st.mgi = 0;
      bool proceed = false;
      if (have) {
        tl.packet_updates++;
        proceed = rpkt_step_begin<1>(c, p, rng, tl, st);
      }
#ifdef ARTIS_ROLE_CYCLES2
      const long long tb1 = clock64();
#endif
      coop_kappa_warm(c, proceed, st.mgi, p.nu_cmf, k, tl);
#ifdef ARTIS_ROLE_CYCLES2
      const long long tb2 = clock64();
      if (proceed) calculate_kappa_rpkt_cont(c, p, st.mgi, k, tl);  // what get_event would do first
      __syncwarp(full);
      const long long tb3 = clock64();
      if (lane == 2) {  // kind-2 slots: begin + coop; kind-3 slots: per-lane kappa
        fill_lanes += tb2 - tc0;
        fill_iters++;
      }
      if (lane == 3) {
        fill_lanes += tb3 - tb2;
        fill_iters++;
      }
      (void)tb1;
#endif
      if (proceed) rpkt_step_finish<1>(c, p, t2, k, rng, tl, st);
      if (have && p.type == ARTIS_TYPE_ESCAPE) est_add(&c.es.ts_scalars[ARTIS_TS_CMF_LUM], p.e_cmf);
    } else if (have) {
      switch (role) {  // warp-uniform
        case 0: do_packet_kind<0>(c, p, t2, k, rng, tl); break;
        case 1: do_packet_kind<1>(c, p, t2, k, rng, tl); break;
        case 2: do_packet_kind<2>(c, p, t2, k, rng, tl); break;
        case 4: do_packet_kind<4>(c, p, t2, k, rng, tl); break;
        default: do_packet_kind<3>(c, p, t2, k, rng, tl); break;
      }
    }
    if (have) {
      events++;
      const bool active = (p.type != ARTIS_TYPE_ESCAPE) && (p.prop_time < t2);
      if (!(active && kind_of(c, p) == role && events < max_events)) {
        __stcg(rngctr + idx, (unsigned long long)rng.ctr);
        packet_store_cg(pkts + idx, p);
        __threadfence();
        done = true;
        k2 = active ? kind_of(c, p) : -1;
        have = false;
      }
    }
    __syncwarp(full);
#ifdef ARTIS_ROLE_CYCLES
    if (lane == role) fill_lanes += clock64() - tc0 - 27;  // experiment: slots 56.. hold cycles instead of busy lanes
    const long long tp0 = clock64();
#endif
    if (__ballot_sync(full, done) != 0u) {
#pragma unroll
      for (int kk = 0; kk < NKIND; kk++) {
        const unsigned m = __ballot_sync(full, done && k2 == kk);
        if (m != 0u) {
          const int leader = __ffs(m) - 1;
          unsigned long long b = 0ull;
          if (lane == leader) b = atomicAdd(pv.tail + kk, (unsigned long long)__popc(m));
          b = __shfl_sync(full, b, leader);
          if (done && k2 == kk) {
            int *slot = pv.ring[kk] + (size_t)((b + (unsigned long long)__popc(m & lt_mask)) % (unsigned long long)pv.cap);
            while (atomicCAS(slot, -1, idx) != -1) {
            }
          }
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
#if defined(ARTIS_ROLE_CYCLES) && !defined(ARTIS_ROLE_CYCLES2)
    if (lane == 2) {  // experiment: kind-2 slots hold the push section
      fill_lanes += clock64() - tp0;
      fill_iters++;
    }
#endif
  }
  tally_flush(c, tl);
  est_private_flush();
  if (lane < NKIND && fill_iters > 0) {
    ctr_add(&c.es.counters[ARTIS_CTR_WARP_ITERS0 + lane], fill_iters);
    ctr_add(&c.es.counters[ARTIS_CTR_BUSY_LANES0 + lane], fill_lanes);
  }
}

__global__ void k_bucket_hist(ArtisPacket *pkts, unsigned long long *rngctr, int n, int *counts) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    artis::body_bucket_hist(d_ctx, pkts, rngctr, i, counts);
}
__global__ void k_bucket_scan(int *counts, int *kind_sizes) { artis::body_bucket_scan(d_ctx, counts, kind_sizes); }
__global__ void k_bucket_scatter(const ArtisPacket *pkts, int n, int *offsets, int *const *queues) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    artis::body_bucket_scatter(d_ctx, pkts, i, offsets, queues);
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
__global__ void k_probe_levelpop(const int *a, const int *b, const int *c, const int *d, int n, double *o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) artis::body_probe_levelpop(d_ctx, a, b, c, d, i, o);
}

cudaStream_t g_stream = nullptr;
cudaEvent_t g_ev0 = nullptr, g_ev1 = nullptr;
int g_sms = SM_COUNT_B200;
int g_transport_blocks_per_sm = 4;
int g_persistent_blocks_per_sm = 2;
int g_nbfcontinua_ground = 0;
double *g_est_block = nullptr;
size_t g_est_ndoubles = 0, g_persistent_smem_bytes = 0;
constexpr size_t g_est_private_max_bytes = (size_t)ARTIS_EST_PRIVATE_MAX_KB * 1024;
double *g_kcache = nullptr;
int *g_kflag = nullptr;
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
bool be_select_device(int device, std::string *why) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    *why = std::string("cudaSetDevice: ") + cudaGetErrorString(e);
    return false;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) g_sms = prop.multiProcessorCount;
  int nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_wavefront<1>, 128, 0) == cudaSuccess && nb > 0)
    g_transport_blocks_per_sm = nb;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_transport_persistent, PERSIST_THREADS, 0) == cudaSuccess && nb > 0)
    g_persistent_blocks_per_sm = nb;  // the whole grid must be resident: idle warps wait for the others' packets
  if (g_stream == nullptr) {
    note(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking), "cudaStreamCreate");
    note(cudaEventCreate(&g_ev0), "cudaEventCreate");
    note(cudaEventCreate(&g_ev1), "cudaEventCreate");
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
void be_run_ma(long long n) { if (n > 0) k_ma<<<grid_for(n, 128, 8), 128, 0, g_stream>>>(n); }
void be_run_cool(long long n) { if (n > 0) k_cool<<<grid_for(n, 64, 8), 64, 0, g_stream>>>(n); }
void be_run_bucket(ArtisPacket *p, unsigned long long *r, int n, int *buckets, int *kind_sizes, int *const *queues) {
  k_bucket_hist<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(p, r, n, buckets);
  k_bucket_scan<<<1, 1, 0, g_stream>>>(buckets, kind_sizes);
  k_bucket_scatter<<<grid_for(n, 256, 8), 256, 0, g_stream>>>(p, n, buckets, queues);
}
void be_run_wavefront(ArtisPacket *p, unsigned long long *r, const artis::QueueView &qv, int kind, long long maxev, int block) {
  if (block > 128) block = 128;  // __launch_bounds__
  // persistent grid: every SM gets its full complement of resident blocks; the queue size is read on the device
  const int grid = g_sms * g_transport_blocks_per_sm;
  switch (kind) {
    case 0: k_wavefront<0><<<grid, block, 0, g_stream>>>(p, r, qv, maxev); break;
    case 1: k_wavefront<1><<<grid, block, 0, g_stream>>>(p, r, qv, maxev); break;
    case 2: k_wavefront<2><<<grid, block, 0, g_stream>>>(p, r, qv, maxev); break;
    case 4: k_wavefront<4><<<grid, block, 0, g_stream>>>(p, r, qv, maxev); break;
    default: k_wavefront<3><<<grid, block, 0, g_stream>>>(p, r, qv, maxev); break;
  }
}
// barrier-free scheduler: rings = NKIND rings of `cap` ints (initial entries written by be_run_bucket, -1 elsewhere),
// pmeta = head[NKIND], tail[NKIND], n_active (int)
bool be_run_persistent(ArtisPacket *p, unsigned long long *r, int *rings, int cap, unsigned long long *pmeta, int n_initial,
                       long long maxev) {
  PersistView pv;
  for (int k = 0; k < artis::NKIND; k++) pv.ring[k] = rings + (size_t)k * cap;
  pv.head = pmeta;
  pv.tail = pmeta + artis::NKIND;
  pv.n_active = reinterpret_cast<int *>(pmeta + 2 * artis::NKIND);
  pv.n_initial = n_initial;
  pv.t_start = pmeta + 2 * artis::NKIND + 1;
  pv.cap = cap;
  const int kstride = 8 + g_nbfcontinua_ground;
  if (CONSOLIDATE_MAX_BUSY >= 0 && (size_t)cap * kstride > g_kcache_doubles) {
    if (g_kcache) cudaFree(g_kcache);
    if (g_kflag) cudaFree(g_kflag);
    g_kcache = nullptr;
    g_kflag = nullptr;
    g_kcache_doubles = 0;
    if (cudaMalloc(&g_kcache, (size_t)cap * kstride * sizeof(double)) != cudaSuccess ||
        cudaMalloc(&g_kflag, (size_t)cap * sizeof(int)) != cudaSuccess) {
      note(cudaGetLastError(), "cudaMalloc (opacity cache side buffer)");
      return true;  // be_sync reports the error
    }
    g_kcache_doubles = (size_t)cap * kstride;
  }
  if (g_kflag) note(cudaMemsetAsync(g_kflag, 0, (size_t)cap * sizeof(int), g_stream), "memset");
  pv.kcache = g_kcache;
  pv.kflag = g_kflag;
  pv.kstride = kstride;
  // block-private estimators when the block fits next to a second resident CTA (227 KB per SM, 1 KB reserved each)
  size_t est_bytes = g_est_ndoubles * sizeof(double);
  if (est_bytes > g_est_private_max_bytes) est_bytes = 0;
  pv.est_block = g_est_block;
  pv.est_private_doubles = (unsigned)(est_bytes / sizeof(double));
  if (est_bytes != g_persistent_smem_bytes) {
    note(cudaFuncSetAttribute(k_transport_persistent, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)est_bytes),
         "cudaFuncSetAttribute");
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_transport_persistent, PERSIST_THREADS, est_bytes) ==
            cudaSuccess && nb > 0)
      g_persistent_blocks_per_sm = nb;
    g_persistent_smem_bytes = est_bytes;
  }
  const int grid = g_sms * g_persistent_blocks_per_sm;
  k_transport_persistent<<<grid, PERSIST_THREADS, est_bytes, g_stream>>>(p, r, pv, maxev, 20ull * 1000000000ull);
  return true;
}
void be_fill_bytes(void *p, int byte, size_t n) { note(cudaMemsetAsync(p, byte, n, g_stream), "memset"); }
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
// one sum all-reduce over the FP64 estimator block, one over the int64 counters and one over the int32 line
// counters, issued on the transport stream right behind the last transport kernel
int be_comm_allreduce(double *d, size_t nd, long long *l, size_t nl, int *i, size_t ni, std::string *err) {
  if (g_nccl.comm == nullptr || g_nccl.nranks <= 1) return ARTIS_OK;  // single shard: nothing to reduce
  const int ncclSum = 0, ncclInt32 = 2, ncclInt64 = 4, ncclFloat64 = 8;
  int rc = g_nccl.AllReduce(d, d, nd, ncclFloat64, ncclSum, g_nccl.comm, g_stream);
  if (rc == 0) rc = g_nccl.AllReduce(l, l, nl, ncclInt64, ncclSum, g_nccl.comm, g_stream);
  if (rc == 0 && ni > 0) rc = g_nccl.AllReduce(i, i, ni, ncclInt32, ncclSum, g_nccl.comm, g_stream);
  if (rc != 0) {
    *err = std::string("ncclAllReduce: ") + g_nccl.GetErrorString(rc);
    return ARTIS_ERR_NCCL;
  }
  note(cudaStreamSynchronize(g_stream), "allreduce sync");
  return ARTIS_OK;
}

}  // namespace

#include "engine.inc"
