// km_resident.cu — km_run for a job that fits on the chip: ONE cooperative launch runs the whole
// Lloyd loop (A5: assign; it++; stop if nothing changed or it == max_iter; update) with the points
// and their labels resident in shared memory (configs[1]: 1e6 x d2 = 12 MB over 148 SMs).
//
// Launch by launch, an iteration of such a job is 16 MB of traffic — 2.4 us at the HBM roofline,
// less than a kernel launch costs — so the stepwise path is latency-bound at ~18 us per iteration
// however fast its kernel is.  Here an iteration costs its arithmetic plus one grid-wide barrier:
//   every block   argmin over all k for its points in the oracle's own binary64 arithmetic
//                 (km_dist2_f64: dims ascending, separate multiply and add; strict <, lowest
//                 index first — no filter, nothing to re-score: k * d is small by construction),
//                 moves of changed points into a block table of order-free int64 sums,
//                 block table -> global delta table (integer RED);
//   grid.sync()
//   every block   adds the delta table to its own copy of the merged sums, reads `changed`, and
//                 — unless the loop ends — derives the next centroids from the sums with the same
//                 operations as the stepwise prologue (km_assign_small.cu: km_small_centroid), so
//                 all blocks hold identical centroids without exchanging them.
// Three rotating delta tables make the single barrier enough: block 0 clears the table of
// iteration t+2 after barrier t, when every block has long finished reading it (after barrier t-1).
// Bits out — labels, sums, centroids, iteration count, inertia (from the sums) — are
// those of the stepwise path and of the oracle; tests/test_gpu_resident.py compares all three.
#include <cooperative_groups.h>
#include <cstdio>

#include "km_commit.cuh"
#include "km_internal.h"

namespace cg = cooperative_groups;

#define KR_THREADS 512
#define KR_PPR 4            // points per thread per round of the assignment loop

struct KmResidentParams {
  const float *x;          // [n][D] in store order
  int32_t *labels;         // [n] in/out (all -1 on entry: km_run resets them)
  int64_t n;
  int k;
  int per_block;           // points per block (the last blocks may hold fewer, or none)
  int max_iter;
  const double *C_in;      // [k][D] starting centroids
  double *C_out;           // [k][D] centroids the last assignment used
  long long *S_out;        // merged sums + counts of the final labels (table layout, L.scal entries)
  long long *Dt[3];        // rotating delta tables, zero on entry; left zero on exit
  KmStatSlot *ring;        // pinned host ring: slot [it % KM_RING] gets the final iteration's stats
  int32_t *state_it;
  double scale_a, scale_b;
  float tau2, rho2;        // the binary32 filter's absolute and relative error terms (hi-only centroids)
  int QB, qa;
  long long Ta, Tb;
  int qt, QBt, qg, QBg;
};

template <int D>
__global__ void __launch_bounds__(KR_THREADS, 1) km_resident_kernel(const KmResidentParams p) {
  cg::grid_group grid = cg::this_grid();
  int it = 0;
#ifdef KR_TIMING
  __shared__ unsigned long long kr_ts[64];
  int kr_nts = 0;
#define KR_STAMP() do { if (blockIdx.x == 0 && threadIdx.x == 0 && kr_nts < 64) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); kr_ts[kr_nts++] = t_; } } while (0)
#else
#define KR_STAMP() do { } while (0)
#endif
  KR_STAMP();
  extern __shared__ __align__(16) unsigned char kr_smem[];
  const int k = p.k, pb = p.per_block;
  const KmAccLayout L = km_acc_layout(k, D);
  double *cd = reinterpret_cast<double *>(kr_smem);                 // [k][D] current centroids
  long long *Sloc = reinterpret_cast<long long *>(cd + k * D);     // this block's copy of the merged sums + counts
  long long *tab = Sloc + L.scal;                                   // [warps][L.len] deltas of the running iteration, a table per warp
  float *xs = reinterpret_cast<float *>(tab + L.len * (KR_THREADS / 32));   // [D][pb] points, one dimension after the other
  int *lab = reinterpret_cast<int *>(xs + (size_t)D * pb);          // [pb]
  __shared__ long long s_red[KR_THREADS / 32][2];
  const int lane = threadIdx.x & 31;
  const int64_t i0 = (int64_t)blockIdx.x * pb;
  const int cnt = (int)max((int64_t)0, min((int64_t)pb, p.n - i0));

  for (int e = threadIdx.x; e < cnt * D; e += KR_THREADS) {
    const int pi = e / D, t = e - pi * D;
    xs[(size_t)t * pb + pi] = __ldg(p.x + i0 * D + e);
  }
  for (int pi = threadIdx.x; pi < cnt; pi += KR_THREADS) lab[pi] = -1;   // km_run starts from unlabelled points
  for (int e = threadIdx.x; e < k * D; e += KR_THREADS) cd[e] = p.C_in[e];
  for (int e = threadIdx.x; e < (int)L.scal; e += KR_THREADS) Sloc[e] = 0;
  KR_STAMP();   // (after the loads are issued; they complete at the first barrier)
  long long changed = 0, rescored = 0;
  long long *wtab = tab + (size_t)(threadIdx.x >> 5) * L.len;      // this warp's own delta table: no atomics between warps
  KmBlockTab T;
  T.L = L;
  T.tab = wtab;
  const int rounds = (cnt + KR_THREADS * KR_PPR - 1) / (KR_THREADS * KR_PPR);
  const int rounds1 = (cnt + KR_THREADS - 1) / KR_THREADS;
  float *cf = reinterpret_cast<float *>(lab + pb);                  // [k][D] binary32 copies of the centroids (the filter's operands)
  int *wq = reinterpret_cast<int *>(cf + k * D) + (threadIdx.x >> 5) * (64 * 3);   // this warp's queue of moves {point, from, to}
  while (true) {
    KR_STAMP();   // loop top
    km_tab_zero(tab, L.len * (KR_THREADS / 32));
    KR_STAMP();   // zeroed
    for (int e = threadIdx.x; e < k * D; e += KR_THREADS) cf[e] = __double2float_rn(cd[e]);
#ifdef KR_TIMING
    __shared__ unsigned long long kr_arr[KR_THREADS / 32];
    if (blockIdx.x == 0 && lane == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); kr_arr[threadIdx.x >> 5] = t_; }
#endif
    __syncthreads();
#ifdef KR_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 0 && it == 3) {
      printf("arrivals at loop top (ns since start):");
      for (int w = 0; w < KR_THREADS / 32; w++) printf(" %llu", kr_arr[w] - kr_ts[0]);
      printf("\n");
    }
#endif
    // ---- a3: assignment of this block's points.  Binary32 differences first; the oracle's binary64
    // arithmetic only where the two best are closer than the filter's error bound.
    KR_STAMP();   // loop top synced
    long long mine = 0, resc = 0;
    int q_count = 0;                                 // warp-uniform: moves waiting in this warp's queue
    auto drain = [&](int base, int m) {              // lane l carries out queued move base + l
      __syncwarp();
      if (lane < m) {
        const int *e = wq + (base + lane) * 3;
        const int pi = e[0];
        float xm[D];
#pragma unroll
        for (int t = 0; t < D; t++) xm[t] = xs[(size_t)t * pb + pi];
        km_move_point(T, xm, D, e[2], e[1], p.scale_a, p.scale_b);
      }
      __syncwarp();
    };
    for (int r = 0; r < rounds; r++) {
      // KR_PPR points per thread at a time: independent chains, each centroid read once for all of them
      float xv[KR_PPR][D], b1[KR_PPR], b2[KR_PPR];
      int bj[KR_PPR];
#pragma unroll
      for (int q = 0; q < KR_PPR; q++) {
        const int pi = (r * KR_PPR + q) * KR_THREADS + threadIdx.x;
#pragma unroll
        for (int t = 0; t < D; t++) xv[q][t] = pi < cnt ? xs[(size_t)t * pb + pi] : 0.f;
        b1[q] = __int_as_float(0x7f800000); b2[q] = b1[q]; bj[q] = 0;
      }
      for (int j = 0; j < k; j++) {
        float cj[D];
#pragma unroll
        for (int t = 0; t < D; t++) cj[t] = cf[j * D + t];
#pragma unroll
        for (int q = 0; q < KR_PPR; q++) {
          float part = 0.f;
#pragma unroll
          for (int t = 0; t < D; t++) { const float df = xv[q][t] - cj[t]; part = fmaf(df, df, part); }
          // best and second best without a branch (a NaN score leaves both alone: the gap test then fails)
          bj[q] = part < b1[q] ? j : bj[q];
          b2[q] = fminf(b2[q], fmaxf(b1[q], part));
          b1[q] = fminf(b1[q], part);
        }
      }
#pragma unroll
      for (int q = 0; q < KR_PPR; q++) {
        const int pi = (r * KR_PPR + q) * KR_THREADS + threadIdx.x;
        const bool in = pi < cnt;
        int nw = bj[q];
        if (!((b2[q] - b1[q]) > fmaf(p.rho2, b2[q], p.tau2))) {      // too close to call in binary32 (or not finite)
          double best = __longlong_as_double(0x7FF0000000000000LL);
          nw = 0;
          if (in) resc++;
          for (int j = 0; j < k; j++) {
            const double dj = km_dist2_f64(xv[q], cd + j * D, D);
            if (dj < best) { best = dj; nw = j; }
          }
        }
        const int old = in ? lab[pi] : 0;
        const bool moved = in && nw != old;
        if (moved) { mine++; lab[pi] = nw; }
        // the first iteration moves every point: its sums are taken cluster by cluster below instead.
        // Later ones queue their moves per warp and carry them out 32 at a time with every lane busy.
        if (it > 0) {
          const unsigned bal = __ballot_sync(0xffffffffu, moved);
          if (bal) {
            if (moved) {
              int *e = wq + (q_count + __popc(bal & ((1u << lane) - 1u))) * 3;
              e[0] = pi; e[1] = old; e[2] = nw;
            }
            q_count += __popc(bal);
            if (q_count >= 32) { drain(q_count - 32, 32); q_count -= 32; }
          }
        }
      }
    }
    KR_STAMP();   // scored
    if (q_count > 0) drain(0, q_count);
    if (it == 0) {
      // every point is new to its cluster: per cluster, each thread adds up its own points in registers,
      // the warp adds up its threads with shuffles, one lane writes them into the warp's table
      for (int j = 0; j < k; j++) {
        long long sa[D], sb[D], c = 0;
#pragma unroll
        for (int t = 0; t < D; t++) { sa[t] = 0; sb[t] = 0; }
        for (int r = 0; r < rounds1; r++) {
          const int pi = r * KR_THREADS + threadIdx.x;
          if (pi < cnt && lab[pi] == j) {
            c++;
#pragma unroll
            for (int t = 0; t < D; t++) {
              long long a, b;
              km_fixed_split((double)xs[(size_t)t * pb + pi], p.scale_a, p.scale_b, a, b);
              sa[t] += a; sb[t] += b;
            }
          }
        }
        c = km_warp_sum_ll(c);
        if (c == 0) continue;                        // warp-uniform
#pragma unroll
        for (int t = 0; t < D; t++) {
          const long long ra = km_warp_sum_ll(sa[t]), rb = km_warp_sum_ll(sb[t]);
          if (lane == 0) { wtab[L.sa + j * D + t] = ra; wtab[L.sb + j * D + t] = rb; }
        }
        if (lane == 0) wtab[L.cnt + j] = c;
      }
    }
    KR_STAMP();   // assign done (thread 0)
    mine = km_warp_sum_ll(mine);
    if (lane == 0 && mine != 0) wtab[L.scal + KM_ACC_CHANGED] += mine;
    resc = km_warp_sum_ll(resc);
    if (lane == 0 && resc != 0) wtab[L.scal + KM_ACC_RESCORED] += resc;
    __syncthreads();
    long long *Dw = p.Dt[it % 3];
    // the warps' tables -> one RED per non-zero entry; 8 lanes share an entry (two tables each)
    static_assert(KR_THREADS / 32 == 16, "the flush pairs 16 warp tables");
    for (int e8 = threadIdx.x; e8 < (((int)L.len * 8 + 31) & ~31); e8 += KR_THREADS) {   // whole warps: the shuffles need all lanes
      const int e = e8 >> 3, w0 = e8 & 7;
      long long v = e < (int)L.len ? tab[(size_t)w0 * L.len + e] + tab[(size_t)(w0 + 8) * L.len + e] : 0;
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      if (w0 == 0 && v != 0 && e < (int)L.len) km_atomic_add_ll(Dw + e, v);
    }
    KR_STAMP();   // flushed
    grid.sync();                                   // every block's deltas of this iteration are in Dw
    KR_STAMP();   // barrier passed
    it++;
    for (int e = threadIdx.x; e < (int)L.scal; e += KR_THREADS) Sloc[e] += __ldcg(Dw + e);
    changed = __ldcg(Dw + L.scal + KM_ACC_CHANGED);
    rescored = __ldcg(Dw + L.scal + KM_ACC_RESCORED);
    if (blockIdx.x == 0) {                         // the table of the iteration after the next one
      long long *Dz = p.Dt[(it + 1) % 3];
      for (int e = threadIdx.x; e < (int)L.len; e += KR_THREADS) Dz[e] = 0;
    }
    __syncthreads();
    KR_STAMP();   // merged
    if (changed == 0 || it == p.max_iter) break;   // A5: the update after the last assignment never happens
    // ---- a4: centroids from the sums; an empty cluster keeps its centroid (A4)
    for (int e = threadIdx.x; e < k * D; e += KR_THREADS) {
      const long long cn = Sloc[L.cnt + e / D];
      if (cn > 0) cd[e] = __ddiv_rn(km_fixed_to_double(Sloc[L.sa + e], Sloc[L.sb + e], p.QB, p.qa + p.QB), (double)cn);
    }
  }

  for (int pi = threadIdx.x; pi < cnt; pi += KR_THREADS) p.labels[i0 + pi] = lab[pi];
  KR_STAMP();   // labels written
  grid.sync();                                     // every block has read the last table: leave all three empty
  if (blockIdx.x == 0)
    for (int q = 0; q < 3; q++)
      for (int e = threadIdx.x; e < (int)L.len; e += KR_THREADS) p.Dt[q][e] = 0;
#ifdef KR_TIMING
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    printf("resident stamps (ns since start):");
    for (int q = 1; q < kr_nts; q++) printf(" %llu", kr_ts[q] - kr_ts[0]);
    printf("\n");
  }
#endif
  if (blockIdx.x == 0) {
    // inertia of the last assignment from the sums: T - sum c (2 S - n c), c = the centroids it used
    const double scale_g = scalbn(1.0, p.qg), scale_gb = scalbn(1.0, p.QBg);
    long long ga = 0, gb = 0;
    for (int e = threadIdx.x; e < k * D; e += KR_THREADS) {
      long long ta, tb;
      km_inertia_term(cd[e], Sloc[L.sa + e], Sloc[L.sb + e], Sloc[L.cnt + e / D], p.QB, p.qa + p.QB, scale_g, scale_gb, ta, tb);
      ga += ta; gb += tb;
      p.C_out[e] = cd[e];
    }
    for (int e = threadIdx.x; e < (int)L.scal; e += KR_THREADS) p.S_out[e] = Sloc[e];
    ga = km_warp_sum_ll(ga);
    gb = km_warp_sum_ll(gb);
    if (lane == 0) { s_red[threadIdx.x >> 5][0] = ga; s_red[threadIdx.x >> 5][1] = gb; }
    __syncthreads();
    if (threadIdx.x == 0) {
      ga = 0; gb = 0;
      for (int w = 0; w < KR_THREADS / 32; w++) { ga += s_red[w][0]; gb += s_red[w][1]; }
      KmStatSlot sl;
      sl.it = it;
      sl.changed = changed;
      sl.inertia = km_inertia_from_limbs(p.Ta, p.Tb, p.QBt, p.qt, ga, gb, p.QBg, p.qg);
      sl.rescored = rescored;
      sl.updated = 0;
      sl.pad = 0;
      p.ring[0] = sl;
      *p.state_it = it;
    }
  }
}

// ------------------------------------------------------------------ host side
// Fits when one block per SM can hold its share of the points, their labels and the three small tables.
bool km_resident_plan(int64_t n, int32_t d, int32_t k, int dtype, int sm_count, int max_smem_optin, int *per_block, int *grid,
                      size_t *smem) {
  if (dtype != KM_DTYPE_F32 || d < 1 || d > 8 || k < 1 || (int64_t)k * d > KM_RESIDENT_MAX_KD || n < 1 || sm_count < 1) return false;
  const KmAccLayout L = km_acc_layout(k, d);
  int64_t pb = (n + sm_count - 1) / sm_count;
  pb = (pb + 31) & ~(int64_t)31;
  const size_t fixed = (size_t)k * d * sizeof(double) + (size_t)(L.scal + L.len * (KR_THREADS / 32)) * sizeof(long long);
  const size_t need = fixed + (size_t)pb * (d + 1) * 4 + (size_t)k * d * sizeof(float) + (size_t)(KR_THREADS / 32) * 64 * 3 * sizeof(int);
  if (need + 2048 > (size_t)max_smem_optin) return false;
  *per_block = (int)pb;
  *grid = (int)((n + pb - 1) / pb);
  *smem = need;
  return true;
}

template <int D>
static cudaError_t kr_launch(const KmResidentParams &p, int grid, size_t smem, cudaStream_t stream) {
  static bool attr_set = false;   // per process; every device here is the same part
  cudaError_t e;
  if (!attr_set) {
    e = cudaFuncSetAttribute(km_resident_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  void *args[] = {const_cast<KmResidentParams *>(&p)};
  return cudaLaunchCooperativeKernel((const void *)km_resident_kernel<D>, dim3(grid), dim3(KR_THREADS), args, smem, stream);
}

cudaError_t km_launch_resident(const KmResidentArgs &a, cudaStream_t stream, int64_t *launches) {
  KmResidentParams p;
  p.x = a.x; p.labels = a.labels; p.n = a.n; p.k = a.k; p.per_block = a.per_block; p.max_iter = a.max_iter;
  p.C_in = a.C_in; p.C_out = a.C_out; p.S_out = a.S_out;
  for (int q = 0; q < 3; q++) p.Dt[q] = a.Dt[q];
  p.ring = a.ring; p.state_it = a.state_it;
  p.scale_a = a.scale_a; p.scale_b = a.scale_b; p.tau2 = a.tau2; p.rho2 = a.rho2; p.QB = a.QB; p.qa = a.qa;
  p.Ta = a.Ta; p.Tb = a.Tb; p.qt = a.qt; p.QBt = a.QBt; p.qg = a.qg; p.QBg = a.QBg;
  cudaError_t e;
  switch (a.d) {
    case 1: e = kr_launch<1>(p, a.grid, a.smem, stream); break;
    case 2: e = kr_launch<2>(p, a.grid, a.smem, stream); break;
    case 3: e = kr_launch<3>(p, a.grid, a.smem, stream); break;
    case 4: e = kr_launch<4>(p, a.grid, a.smem, stream); break;
    case 5: e = kr_launch<5>(p, a.grid, a.smem, stream); break;
    case 6: e = kr_launch<6>(p, a.grid, a.smem, stream); break;
    case 7: e = kr_launch<7>(p, a.grid, a.smem, stream); break;
    case 8: e = kr_launch<8>(p, a.grid, a.smem, stream); break;
    default: return cudaErrorInvalidValue;
  }
  if (e != cudaSuccess) return e;
  (*launches)++;
  return cudaGetLastError();
}
