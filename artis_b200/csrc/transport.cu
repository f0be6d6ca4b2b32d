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
  int cap;
};

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

#ifndef ARTIS_PERSIST_MINBLOCKS
#define ARTIS_PERSIST_MINBLOCKS 2
#endif
__global__ void __launch_bounds__(128, ARTIS_PERSIST_MINBLOCKS) k_transport_persistent(ArtisPacket *pkts, unsigned long long *rngctr, PersistView pv,
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
  while (true) {
    __syncwarp(full);
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
          __nanosleep(1000);
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
        events = 0;
        have = true;
      }
      if (need == full && take == 0) continue;  // lost the race for the last entries
    }
    __syncwarp(full);
    bool done = false;
    int k2 = -1;
    if (have) {
      switch (role) {  // warp-uniform
        case 0: do_packet_kind<0>(c, p, t2, k, rng, tl); break;
        case 1: do_packet_kind<1>(c, p, t2, k, rng, tl); break;
        case 2: do_packet_kind<2>(c, p, t2, k, rng, tl); break;
        case 4: do_packet_kind<4>(c, p, t2, k, rng, tl); break;
        default: do_packet_kind<3>(c, p, t2, k, rng, tl); break;
      }
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
      if (fin != 0u && lane == __ffs(fin) - 1) atomicSub(pv.n_active, __popc(fin));
    }
  }
  tally_flush(c, tl);
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
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_transport_persistent, 128, 0) == cudaSuccess && nb > 0)
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
bool be_run_persistent(ArtisPacket *p, unsigned long long *r, int *rings, int cap, unsigned long long *pmeta, long long maxev) {
  PersistView pv;
  for (int k = 0; k < artis::NKIND; k++) pv.ring[k] = rings + (size_t)k * cap;
  pv.head = pmeta;
  pv.tail = pmeta + artis::NKIND;
  pv.n_active = reinterpret_cast<int *>(pmeta + 2 * artis::NKIND);
  pv.cap = cap;
  const int grid = g_sms * g_persistent_blocks_per_sm;
  k_transport_persistent<<<grid, 128, 0, g_stream>>>(p, r, pv, maxev, 20ull * 1000000000ull);
  return true;
}
void be_fill_bytes(void *p, int byte, size_t n) { note(cudaMemsetAsync(p, byte, n, g_stream), "memset"); }
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
