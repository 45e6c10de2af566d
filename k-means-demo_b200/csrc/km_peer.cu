// km_peer.cu — one-shot all-reduce of a small int64 table over peer memory (NVLink / NVSwitch).
//
// The small-d path ends every pass with a table of a few hundred int64 (per-cluster delta sums,
// counts, changed / rescored counters) that has to be summed over the ranks before the next pass
// may start.  A library all-reduce of 1.5 KB costs 20-30 us of pure latency between two dependent
// kernels — 7 % of a 0.4 ms step at 8 GPUs, two thirds of a configs[1] step.  Here instead every
// rank keeps its three rotating tables inside one region that its peers have mapped (CUDA IPC, set
// up once in km_comm_init), and ONE tiny kernel per pass
//   1. publishes "my table of epoch e is complete" by writing e into a flag slot in every peer's
//      region (a system-scope fence, then the store: the pass's RED atomics are visible first),
//   2. waits until all ranks' flags in its OWN region have reached e (ld.acquire.sys, bounded spin),
//   3. sums the ranks' tables with plain peer loads (ld.relaxed.sys: never from a stale L1 line)
//      into a local buffer the next pass reads.
// Integer sums: the order of the ranks does not matter, every rank computes the same bits.
// Re-use of a table is safe without a second handshake: a rank overwrites table q again only two
// passes later, after it has seen every peer's flag of the pass in between — which a peer raises
// only after its own read of table q has completed (stream order).
// All loads here are ordinary ld/st on mapped peer addresses; no TMA is involved.
#include "km_internal.h"

__device__ __forceinline__ void fence_then_st_sys(unsigned long long *p, unsigned long long v) {
  // one system-scope fence, then a relaxed store: release semantics for everything this GPU wrote before
  asm volatile("fence.acq_rel.sys;\n\tst.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ long long ld_relaxed_sys(const long long *p) {
  long long v;
  asm volatile("ld.relaxed.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

struct KmPeerArgs {
  unsigned char *base[KM_PEER_MAX_WORLD];   // every rank's region as mapped HERE (base[rank] = own)
  int world, rank;
  size_t table_off;                          // byte offset of this pass's table inside a region
  size_t flags_off;                          // ... and of the epoch flags this sum uses
  int64_t len;                               // int64 entries
  unsigned long long epoch;
  long long *out;                            // local: the summed table
  int *err;                                  // host-mapped: set to 1 if a peer never showed up
};
#define KM_PEER_THREADS 1024

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Block 0 publishes; every block waits for all flags by itself (no grid-wide sync needed), then
// sums its slice.  A peer that has not arrived after KM_PEER_TIMEOUT_NS is reported through the
// host-mapped error word (the host turns it into KM_ERR_NCCL) instead of spinning for ever.
#define KM_PEER_TIMEOUT_NS 120000000000ULL
__global__ void __launch_bounds__(KM_PEER_THREADS) km_peer_allreduce_kernel(const KmPeerArgs a) {
  if (blockIdx.x == 0 && (int)threadIdx.x < a.world)   // thread r tells rank r
    fence_then_st_sys(reinterpret_cast<unsigned long long *>(a.base[threadIdx.x] + a.flags_off) + a.rank, a.epoch);
  if ((int)threadIdx.x < a.world) {
    const unsigned long long *f = reinterpret_cast<const unsigned long long *>(a.base[a.rank] + a.flags_off) + threadIdx.x;
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(f) < a.epoch) {
      __nanosleep(32);
      if (global_ns() - t0 > KM_PEER_TIMEOUT_NS) { *reinterpret_cast<volatile int *>(a.err) = 1; break; }
    }
  }
  __syncthreads();
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < a.len; e += (int64_t)gridDim.x * blockDim.x) {
    long long s = 0;
#pragma unroll 4
    for (int r = 0; r < a.world; r++) s += ld_relaxed_sys(reinterpret_cast<const long long *>(a.base[r] + a.table_off) + e);
    a.out[e] = s;
  }
}

// len == 0 is a barrier: every rank has finished reading every table of the epochs before.
cudaError_t km_launch_peer_allreduce(void *const *bases, int world, int rank, size_t table_off, size_t flags_off, int64_t len,
                                     unsigned long long epoch, long long *out, int *err, int sm_count, cudaStream_t s, int64_t *launches) {
  KmPeerArgs a;
  for (int r = 0; r < KM_PEER_MAX_WORLD; r++) a.base[r] = r < world ? static_cast<unsigned char *>(bases[r]) : nullptr;
  a.world = world; a.rank = rank; a.table_off = table_off; a.flags_off = flags_off; a.len = len; a.epoch = epoch; a.out = out; a.err = err;
  int blocks = (int)((len + KM_PEER_THREADS - 1) / KM_PEER_THREADS);
  const int cap = sm_count > 8 ? sm_count / 2 : 8;   // every block polls the flags by itself: all of them must be resident
  blocks = blocks < 1 ? 1 : (blocks > cap ? cap : blocks);
  km_peer_allreduce_kernel<<<blocks, KM_PEER_THREADS, 0, s>>>(a);
  (*launches)++;
  return cudaGetLastError();
}
