// km_assign_small.cu — KM_PATH_FILTER_F32 for small d (1..8): the bandwidth-bound path
// (BASELINE.json configs[1] and [4]: d=2, k=8 / k=64).
//
// Shape of the kernel:
//   * persistent grid: (SMs x resident blocks) fat blocks stride over tiles of points, so the
//     per-block prologue (centroids -> shared memory) and epilogue (table flush, scalar
//     reduction, optional fused finalize) are paid once per SM, not once per tile;
//   * each lane moves 16 bytes per load when d is 1, 2 or 4 (float4), labels as int2/int4;
//   * distances in binary32 two dimensions at a time with packed f32x2 add/mul/fma (sm_100a
//     add.rn.f32x2 / fma.rn.f32x2): a point's (x0,x1) arrives from the vector load already
//     as an aligned register pair, centroids are stored pre-negated so a centroid is one
//     broadcast LDS and (x - c) is one packed add;
//   * top-2 per point on integer keys: (distance bits & ~mask) | centroid index, three
//     integer min/max per (point, centroid), no selects;
//   * a point is re-scored in binary64 (the oracle's exact operation order) only when the gap
//     between its two best keys is inside the rigorous error bound of DESIGN.md §5, so the
//     labels are the oracle's bit for bit at binary32 cost;
//   * label compare, move of re-labelled points between the order-free per-cluster sums,
//     change count and inertia are fused into the same pass: x is read once per iteration;
//   * the centroid update is folded into the NEXT pass's prologue: every block rebuilds the
//     centroids from the order-free sums (k*d divisions — nothing next to its share of the
//     points), so a Lloyd iteration is ONE kernel launch with no grid-wide hand-off, no
//     last-block tail and, across GPUs, one all-reduce and nothing else.
#include <type_traits>

#include "km_commit.cuh"
#include "km_finalize.cuh"
#include "km_internal.h"

#define KM_SMALL_THREADS 512

// Packed binary32 pairs held as one 64-bit register each, so the register allocator keeps
// a pair together across the centroid loop instead of re-packing it at every use.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// fire-and-forget 8-byte global -> shared copy (LDGSTS): no register holds the value in flight
__device__ __forceinline__ void km_cp_async8(void *smem_dst, const void *gsrc) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void km_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void km_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int D>
struct KmSmallCfg {
  static constexpr int PPV = (D == 1) ? 4 : (D == 2) ? 2 : 1;   // points per 16-byte vector (else 1 row per "vector")
  static constexpr int NV = (D == 1) ? 1 : (D <= 4) ? 2 : 1;    // vectors / rows per thread per round
  static constexpr int PPT = PPV * NV;                          // points per thread per tile
  static constexpr int NP = D / 2;                              // packed dimension pairs per point
  static constexpr bool TAIL = (D & 1) != 0;                    // odd last dimension handled as a scalar
};

// Offset, within a round, of the first point of this thread's v-th vector.  A warp's NV vectors sit next to
// each other, so the points of one warp-round are ONE contiguous range of 32 * PPT points (on the
// cell-sorted store: usually a single cell).
// WQ: rounds are per warp (32 * PPT points, handed out warp by warp), not per block.
template <int NV, int PPV, bool WQ>
__device__ __forceinline__ int km_slot(int v) {
  return WQ ? (v * 32 + ((int)threadIdx.x & 31)) * PPV
            : ((((int)threadIdx.x >> 5) * NV + v) * 32 + ((int)threadIdx.x & 31)) * PPV;
}

#define KM_PRUNE_MAGIC 12582912.0f
#define KM_PRUNE_CAP 7   // candidates kept per cell (one 64-bit word: count byte + 7 indices)

struct KmSmallParams {
  const float *x;
  int64_t n;
  int32_t k;
  int32_t *labels;
  int incremental;
  int exact_inertia;
  int tab_in_smem;
  int ntab;            // tables in shared memory: 1 per block, or one per warp (passes that accumulate every point)
  unsigned key_mask;   // low bits of the distance key that hold the centroid index
  float rho2;          // relative part of the ambiguity threshold (carries the 2.5x margin)
  double scale_a, scale_b;  // fixed-point scales of the sums (constant per upload)
  int64_t chunk;       // points per chunk (multiple of 4); block b walks chunks b, b + grid, ...
  int64_t nchunks;
  int64_t n_big;       // cell-sorted store: units [0, n_big) have `chunk` points, the rest `chunk_small` (the tail of the pass, for balance)
  int64_t chunk_small;
  float tau2;          // absolute error term of the filter for this instantiation (hi or hi/lo)
  float cboundf;       // the bound on max|c| itself
  KmPass ps;           // sums / centroids / deltas this pass reads and writes
  KmPrune2 pr;         // d == 2 candidate grid (used by the PRUNE instantiation)
  int sorted;          // PRUNE: the points are stored in ascending cell order of exactly this grid
  const unsigned long long *cells_g;   // ... and the candidate words of this pass were built by km_prune_grid_kernel
  unsigned *work_counter;              // ... and the next unit of work is taken from here
};

// What one thread holds for one round: PPT points as packed dimension pairs (+ the odd last
// coordinate), their previous labels, and where they sit.
template <int D>
struct KmBatch {
  using Cfg = KmSmallCfg<D>;
  f32x2 xq[Cfg::PPT][Cfg::NP > 0 ? Cfg::NP : 1];
  float xs[Cfg::PPT];
  int old[Cfg::PPT];
};

// Load the thread's NV vectors of round `r0` (first point index of the round) within the
// block's chunk [.., c1).  A vector wholly inside the chunk is one 16-byte load (d = 1, 2, 4)
// or a few 8/16-byte loads (other even d); the ragged end is read point by point.
template <int D, bool LABELS = true, bool WQ = false>
__device__ __forceinline__ void km_load_batch(KmBatch<D> &b, const float *__restrict__ xb,
                                              const int32_t *__restrict__ lb, int r, int cn) {
  using Cfg = KmSmallCfg<D>;
  constexpr int PPV = Cfg::PPV, NV = Cfg::NV, NP = Cfg::NP;
  constexpr bool TAIL = Cfg::TAIL;
#pragma unroll
  for (int v = 0; v < NV; v++) {
    const int i0 = r + km_slot<NV, PPV, WQ>(v);   // offset within the chunk
    if (i0 + PPV <= cn) {
      const float *row = xb + i0 * D;
      if constexpr (D == 1) {
        const float4 q = __ldg(reinterpret_cast<const float4 *>(row));
        b.xs[v * 4 + 0] = q.x; b.xs[v * 4 + 1] = q.y; b.xs[v * 4 + 2] = q.z; b.xs[v * 4 + 3] = q.w;
        if (LABELS) {
          const int4 o = *reinterpret_cast<const int4 *>(lb + i0);
          b.old[v * 4 + 0] = o.x; b.old[v * 4 + 1] = o.y; b.old[v * 4 + 2] = o.z; b.old[v * 4 + 3] = o.w;
        }
      } else if constexpr (D == 2) {
        const ulonglong2 q = __ldg(reinterpret_cast<const ulonglong2 *>(row));
        b.xq[v * 2 + 0][0] = q.x; b.xq[v * 2 + 1][0] = q.y;
        if (LABELS) {
          const int2 o = *reinterpret_cast<const int2 *>(lb + i0);
          b.old[v * 2 + 0] = o.x; b.old[v * 2 + 1] = o.y;
        }
      } else {
        if constexpr (D % 4 == 0) {
#pragma unroll
          for (int m = 0; m < D / 4; m++) {
            const ulonglong2 q = __ldg(reinterpret_cast<const ulonglong2 *>(row) + m);
            b.xq[v][2 * m] = q.x; b.xq[v][2 * m + 1] = q.y;
          }
        } else if constexpr (D % 2 == 0) {
#pragma unroll
          for (int m = 0; m < NP; m++) b.xq[v][m] = __ldg(reinterpret_cast<const unsigned long long *>(row) + m);
        } else {
#pragma unroll
          for (int m = 0; m < NP; m++) b.xq[v][m] = pk2(__ldg(row + 2 * m), __ldg(row + 2 * m + 1));
          b.xs[v] = __ldg(row + D - 1);
        }
        if (LABELS) b.old[v] = lb[i0];
      }
    } else {
#pragma unroll
      for (int u = 0; u < PPV; u++) {
        const int i = i0 + u;
        const bool live = i < cn;
        const int q = v * PPV + u;
        if (LABELS) b.old[q] = live ? lb[i] : 0;
#pragma unroll
        for (int m = 0; m < NP; m++)
          b.xq[q][m] = live ? pk2(xb[i * D + 2 * m], xb[i * D + 2 * m + 1]) : pk2(0.f, 0.f);
        if (D == 1 || TAIL) b.xs[q] = live ? xb[i * D + D - 1] : 0.f;
      }
    }
  }
}

// The previous labels of the thread's points alone (same indexing as km_load_batch).  The pruned
// kernel fetches them at the top of the round they are used in rather than a round ahead: they
// are not needed before the tail, and four fewer live registers keep the prefetched labels from
// being spilled (a spill store waits for the load it spills — the prefetch would be lost).
template <int D, bool WQ = false>
__device__ __forceinline__ void km_load_labels(KmBatch<D> &b, const int32_t *__restrict__ lb, int r, int cn) {
  using Cfg = KmSmallCfg<D>;
  constexpr int PPV = Cfg::PPV, NV = Cfg::NV;
#pragma unroll
  for (int v = 0; v < NV; v++) {
    const int i0 = r + km_slot<NV, PPV, WQ>(v);
    if (i0 + PPV <= cn) {
      if constexpr (PPV == 4) {
        const int4 o = *reinterpret_cast<const int4 *>(lb + i0);
        b.old[v * 4 + 0] = o.x; b.old[v * 4 + 1] = o.y; b.old[v * 4 + 2] = o.z; b.old[v * 4 + 3] = o.w;
      } else if constexpr (PPV == 2) {
        const int2 o = *reinterpret_cast<const int2 *>(lb + i0);
        b.old[v * 2 + 0] = o.x; b.old[v * 2 + 1] = o.y;
      } else {
        b.old[v] = lb[i0];
      }
    } else {
#pragma unroll
      for (int u = 0; u < PPV; u++) b.old[v * PPV + u] = (i0 + u < cn) ? lb[i0 + u] : 0;
    }
  }
}

// One warp builds the candidate word of one cell, lanes over the centroids (ncf: negated binary32
// centroids, `crow` floats per row).  Same arithmetic, same words as the thread-per-cell build inside
// the assignment kernel; the criterion is explained there.  The word is returned in every lane.
__device__ __forceinline__ unsigned long long km_prune_cell_word(const float *ncf, int crow, int k, int cell,
                                                                 const KmPrune2 &pr, float margin, int lane) {
  const float hx = pr.inv_sx, hy = pr.inv_sy;
  const int iy = cell / pr.gp, ix = cell - iy * pr.gp;
  const float mx = ((float)ix - pr.kx) * pr.inv_sx, my = ((float)iy - pr.ky) * pr.inv_sy;
  unsigned long long bestkey = ~0ULL;   // (distance bits, index): min = nearest, lowest index on ties
  for (int j = lane; j < k; j += 32) {
    const float dx = mx + ncf[j * crow], dy = my + ncf[j * crow + 1];
    const unsigned long long key = ((unsigned long long)__float_as_uint(fmaf(dx, dx, dy * dy)) << 32) | (unsigned)j;
    bestkey = min(bestkey, key);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bestkey = min(bestkey, __shfl_xor_sync(0xffffffffu, bestkey, o));
  const int bj = (int)(bestkey & 0xFFFFFFFFu);
  const float dbest = __uint_as_float((unsigned)(bestkey >> 32));
  const float bx = ncf[bj * crow], by = ncf[bj * crow + 1];
  unsigned wlo = 0, whi = 0;
  int cnt = 0;
  for (int j0 = 0; j0 < k; j0 += 32) {
    const int j = j0 + lane;
    bool keep = false;
    if (j < k) {
      const float ncx = ncf[j * crow], ncy = ncf[j * crow + 1];
      const float dx = mx + ncx, dy = my + ncy;
      const float fm = fmaf(dx, dx, dy * dy) - dbest;
      const float swing = fabsf(ncx - bx) * hx + fabsf(ncy - by) * hy;
      keep = fm - swing <= margin;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const int rank = cnt + __popc(bal & ((1u << lane) - 1u));
      if (rank < KM_PRUNE_CAP) {
        const unsigned long long w = (unsigned long long)j << (8 * (rank + 1));
        wlo |= (unsigned)w;
        whi |= (unsigned)(w >> 32);
      }
    }
    cnt += __popc(bal);
  }
  wlo = __reduce_or_sync(0xffffffffu, wlo);
  whi = __reduce_or_sync(0xffffffffu, whi);
  return ((unsigned long long)whi << 32) | wlo | (unsigned long long)(cnt > KM_PRUNE_CAP ? 0xFF : cnt);
}

// the centroid coordinate the pass's prologue derives (same operations, same bits, wherever it is evaluated)
__device__ __forceinline__ double km_small_centroid(const KmPass &ps, const KmAccLayout &L, int e, int D, bool pend, bool upd) {
  const int jj = e / D;
  long long sa = ps.S_in[L.sa + e], sb = ps.S_in[L.sb + e], cn = ps.S_in[L.cnt + jj];
  if (pend) {
    const long long da = ps.D_in[L.sa + e], db = ps.D_in[L.sb + e], dc = ps.D_in[L.cnt + jj];
    if (ps.full_mode) { sa = da; sb = db; cn = dc; } else { sa += da; sb += db; cn += dc; }
  }
  double c = ps.C_in[e];
  if (upd && cn > 0) c = __ddiv_rn(km_fixed_to_double(sa, sb, ps.QB, ps.qa + ps.QB), (double)cn);
  return c;
}

// Cell-sorted d = 2 store: the candidate words of ALL cells, once per pass, by a small kernel in front of
// the assignment kernel (16 cells per block, a warp per cell).  Each block first derives the pass's
// centroids exactly as the assignment kernel's prologue will.  With the words in global memory no block
// of the assignment kernel ever builds a grid or waits at a barrier for one.
#define KM_GRID_THREADS 512
__global__ void __launch_bounds__(KM_GRID_THREADS) km_prune_grid_kernel(const KmPass ps, int k, KmPrune2 pr, float cboundf,
                                                                        unsigned long long *__restrict__ cells_g, unsigned *work_counter) {
  extern __shared__ float gk_ncf[];   // [k][2] negated binary32 centroids
  const KmAccLayout L = km_acc_layout(k, 2);
  const bool pend = ps.D_in != nullptr;
  const long long changed_prev = pend ? ps.D_in[L.scal + KM_ACC_CHANGED] : 0;
  const bool upd = ps.apply_update && (!ps.needs_change || changed_prev > 0);
  for (int e = threadIdx.x; e < k * 2; e += KM_GRID_THREADS)
    gk_ncf[e] = -__double2float_rn(km_small_centroid(ps, L, e, 2, pend, upd));
  __syncthreads();
  const float prune_M = cboundf + pr.absmax;
  const float margin = 1e-5f * prune_M * prune_M;
  const int lane = threadIdx.x & 31;
  // one resident wave of blocks: every warp walks its share of the cells (the prologue above — dependent global
  // loads and a division — is paid once per block, not once per 16 cells)
  const int nwarps = gridDim.x * (KM_GRID_THREADS / 32);
  for (int cell = blockIdx.x * (KM_GRID_THREADS / 32) + (threadIdx.x >> 5); cell < pr.gp * pr.gp; cell += nwarps) {
    const unsigned long long w = km_prune_cell_word(gk_ncf, 2, k, cell, pr, margin, lane);
    if (lane == 0) cells_g[cell] = w;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *work_counter = 0;   // the assignment kernel's warps start counting here
}

// The rare paths of the per-point tail live out of line: they would otherwise be replicated for every
// point slot of a thread (the kernel's code outgrew the instruction cache) and hold registers hostage.
template <int D>
struct KmRow { float v[D]; };
// binary64 argmin over all k in the oracle's operation order (A1/A2): strict <, lowest index wins
template <int D>
__device__ __noinline__ int km_small_rescore(KmRow<D> x, const double *cd, int k) {
  double best = __longlong_as_double(0x7FF0000000000000LL);
  int lab = 0;
  for (int j = 0; j < k; j++) {
    const double dj = km_dist2_f64(x.v, cd + (size_t)j * D, D);
    if (dj < best) { best = dj; lab = j; }
  }
  return lab;
}
// one queued move (x, from, to) per lane into the block's table
template <int D>
__device__ __noinline__ void km_small_drain(long long *tab, int k, const unsigned *e, double scale_a, double scale_b) {
  KmBlockTab T;
  T.L = km_acc_layout(k, D);
  T.tab = tab;
  float xm[D];
#pragma unroll
  for (int t = 0; t < D; t++) xm[t] = __uint_as_float(e[t]);
  km_move_point(T, xm, D, (int)e[D + 1], (int)e[D], scale_a, scale_b);
}

// The same for a full batch of 32 queued moves, called by the whole warp.  On the cell-sorted store (and on
// any first or full-update pass over spatially coherent points) most of a batch shares one (from, to) pair:
// those lanes add their limbs up among themselves — each 64-bit limb as three 21/21/22-bit pieces through the
// 32-bit warp reduction, sums modulo 2^64 being exact whatever the split — and lane 0 alone touches the table.
// Thirty-two lanes hammering one table entry with 64-bit shared-memory atomics (compare-and-swap loops) is
// the slowest thing this kernel can do; this is what made a pass that accumulates every point 10x a
// steady-state one.  Lanes with another pair move their point by themselves as before.
template <int D>
__device__ __noinline__ void km_small_drain_warp(long long *tab, int k, const unsigned *e, double scale_a, double scale_b) {
  KmBlockTab T;
  T.L = km_acc_layout(k, D);
  T.tab = tab;
  float xm[D];
#pragma unroll
  for (int t = 0; t < D; t++) xm[t] = __uint_as_float(e[t]);
  const int from = (int)e[D], to = (int)e[D + 1];
  const int from0 = __shfl_sync(0xffffffffu, from, 0), to0 = __shfl_sync(0xffffffffu, to, 0);
  const bool same = from == from0 && to == to0;
  const unsigned grp = __ballot_sync(0xffffffffu, same);
  if (__popc(grp) < 8) {                       // warp-uniform: not worth the reductions
    km_move_point(T, xm, D, to, from, scale_a, scale_b);
    return;
  }
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int t = 0; t < D; t++) {
    long long a = 0, b = 0;
    if (same) km_fixed_split((double)xm[t], scale_a, scale_b, a, b);
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const unsigned long long v = (unsigned long long)(h == 0 ? a : b);
      const unsigned r0 = __reduce_add_sync(0xffffffffu, (unsigned)(v & 0x1FFFFFu));
      const unsigned r1 = __reduce_add_sync(0xffffffffu, (unsigned)((v >> 21) & 0x1FFFFFu));
      const unsigned r2 = __reduce_add_sync(0xffffffffu, (unsigned)(v >> 42));
      if (lane == 0) {
        const long long tot = (long long)((unsigned long long)r0 + ((unsigned long long)r1 << 21) + ((unsigned long long)r2 << 42));
        const int64_t off = (h == 0 ? T.L.sa : T.L.sb) + t;
        km_atomic_add_ll(tab + off + (int64_t)to0 * D, tot);
        if (from0 >= 0) km_atomic_add_ll(tab + off + (int64_t)from0 * D, -tot);
      }
    }
  }
  if (lane == 0) {
    km_atomic_add_ll(tab + T.L.cnt + to0, (long long)__popc(grp));
    if (from0 >= 0) km_atomic_add_ll(tab + T.L.cnt + from0, -(long long)__popc(grp));
  }
  if (!same) km_move_point(T, xm, D, to, from, scale_a, scale_b);
}

// one move from registers (the odd point of a round whose points otherwise share one target)
template <int D>
__device__ __noinline__ void km_small_move_one(long long *tab, int k, KmRow<D> x, int from, int to, double scale_a, double scale_b) {
  KmBlockTab T;
  T.L = km_acc_layout(k, D);
  T.tab = tab;
  km_move_point(T, x.v, D, to, from, scale_a, scale_b);
}
// per-lane partial sums of points that all go from `from0` to `to0` -> the warp's total -> lane 0 adds it to the table
template <int D>
struct KmLimbs { long long a[D], b[D]; };
template <int D>
__device__ __noinline__ void km_small_warp_add(long long *tab, int k, KmLimbs<D> s, int c, int from0, int to0) {
  const KmAccLayout L = km_acc_layout(k, D);
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int t = 0; t < D; t++) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const unsigned long long v = (unsigned long long)(h == 0 ? s.a[t] : s.b[t]);
      const unsigned r0 = __reduce_add_sync(0xffffffffu, (unsigned)(v & 0x1FFFFFu));
      const unsigned r1 = __reduce_add_sync(0xffffffffu, (unsigned)((v >> 21) & 0x1FFFFFu));
      const unsigned r2 = __reduce_add_sync(0xffffffffu, (unsigned)(v >> 42));
      if (lane == 0) {
        const long long tot = (long long)((unsigned long long)r0 + ((unsigned long long)r1 << 21) + ((unsigned long long)r2 << 42));
        const int64_t off = (h == 0 ? L.sa : L.sb) + t;
        km_atomic_add_ll(tab + off + (int64_t)to0 * D, tot);
        if (from0 >= 0) km_atomic_add_ll(tab + off + (int64_t)from0 * D, -tot);
      }
    }
  }
  const int ctot = __reduce_add_sync(0xffffffffu, c);
  if (lane == 0 && ctot != 0) {
    km_atomic_add_ll(tab + L.cnt + to0, (long long)ctot);
    if (from0 >= 0) km_atomic_add_ll(tab + L.cnt + from0, -(long long)ctot);
  }
}

// Block 0, prologue: the inertia of the pending pass from the sums (km_common.cuh):
// this warp's share of the terms c (2S - n c) over the merged sums and the centroids that pass
// used (C_in), left in sg[warp][0..1].  Kept out of line so its double-double arithmetic costs
// the main loop no registers.
__device__ __noinline__ void km_small_inertia_terms(const long long *S_in, const long long *D_in, const double *C_in,
                                                    int full_mode, int k, int D, int QB, int qa, int qg, int QBg,
                                                    long long (*sg)[2]) {
  const KmAccLayout L = km_acc_layout(k, D);
  const double scale_g = scalbn(1.0, qg), scale_gb = scalbn(1.0, QBg);
  long long ga = 0, gb = 0;
  for (int e = threadIdx.x; e < k * D; e += KM_SMALL_THREADS) {
    const int jj = e / D;
    long long sa = S_in[L.sa + e], sb = S_in[L.sb + e], cn = S_in[L.cnt + jj];
    const long long da = D_in[L.sa + e], db = D_in[L.sb + e], dc = D_in[L.cnt + jj];
    if (full_mode) { sa = da; sb = db; cn = dc; } else { sa += da; sb += db; cn += dc; }
    long long ta, tb;
    km_inertia_term(C_in[e], sa, sb, cn, QB, qa + QB, scale_g, scale_gb, ta, tb);
    ga += ta; gb += tb;
  }
  ga = km_warp_sum_ll(ga);
  gb = km_warp_sum_ll(gb);
  if ((threadIdx.x & 31) == 0) { sg[threadIdx.x >> 5][0] = ga; sg[threadIdx.x >> 5][1] = gb; }
}

template <int D, bool HILO, bool PRUNE, bool SORTED, bool ALLPUSH>
__global__ void __launch_bounds__(KM_SMALL_THREADS, 2) km_assign_small_kernel(const KmSmallParams p) {
  static_assert(!SORTED || PRUNE, "the cell-sorted store belongs to the pruned path");
  static_assert(!PRUNE || D == 2, "candidate pruning is built for d == 2");
  using Cfg = KmSmallCfg<D>;
  constexpr int PPV = Cfg::PPV, NV = Cfg::NV, PPT = Cfg::PPT, NP = Cfg::NP;
  constexpr bool TAIL = Cfg::TAIL;
  constexpr int CW = HILO ? 2 : 1;   // binary32 words per centroid coordinate (hi [, lo])
  constexpr int CROW = D * CW;       // floats per centroid row: -hi[0..D) then -lo[0..D)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int k = p.k;
  float *ncf = reinterpret_cast<float *>(smem_raw);                                  // [k][CROW], negated
  double *cd = reinterpret_cast<double *>(ncf + (((size_t)k * CROW + 3) & ~(size_t)3));  // [k][D] binary64 masters
  long long *stab = reinterpret_cast<long long *>(cd + (size_t)k * D);               // optional per-block table
  const KmAccLayout L = km_acc_layout(k, D);
  // PRUNE: candidate word per cell — in shared memory, built by this block (points in the caller's order),
  // or in global memory, built once per pass by km_prune_grid_kernel (cell-sorted store: a warp stays in
  // one cell for thousands of points and reads its word through L1)
  unsigned long long *cells_sm = reinterpret_cast<unsigned long long *>(stab + (p.tab_in_smem ? L.scal * p.ntab : 0));
  constexpr bool cells_in_smem = PRUNE && !SORTED;
  const unsigned long long *cells = cells_in_smem ? cells_sm : p.cells_g;
  // per-warp queue of points whose sums must move (x, from, to): filled by ballot, drained 32
  // at a time with every lane busy, instead of running the atomics under divergence
  constexpr int QCAP = 64, QW = D + 2;
  unsigned *qbuf = reinterpret_cast<unsigned *>(cells_sm + (cells_in_smem ? p.pr.gp * p.pr.gp : 0)) +
                   (threadIdx.x >> 5) * (QCAP * QW);
  int q_head = 0, q_count = 0;   // warp-uniform
  const KmPass &ps = p.ps;

  // this block's contiguous chunk of points, walked in rounds of THREADS*PPT; everything
  // inside the chunk is addressed with 32-bit offsets from the chunk's base
  // Cell-sorted store: work per point depends on where it lies (how many candidates its cell has), so equal
  // static shares are not equal work.  There the shard is cut into units of a few rounds which the WARPS
  // of the grid take one after another from a global counter (zeroed by km_prune_grid_kernel).
  constexpr bool WQ = SORTED;
  constexpr int per_round = (WQ ? 32 : KM_SMALL_THREADS) * PPT;
  const int lane = threadIdx.x & 31;
  auto grab = [&]() -> int64_t {
    unsigned u = 0;
    if (lane == 0) u = atomicAdd(p.work_counter, 1u);
    return (int64_t)__shfl_sync(0xffffffffu, u, 0);
  };
  int64_t ch = blockIdx.x, ch_next = (int64_t)blockIdx.x + gridDim.x;
  if (WQ) { ch = grab(); ch_next = grab(); }
  auto unit_start = [&](int64_t c) -> int64_t { return (WQ && c >= p.n_big) ? p.n_big * p.chunk + (c - p.n_big) * p.chunk_small : c * p.chunk; };
  auto chunk_len = [&](int64_t c) {
    const int64_t len = (WQ && c >= p.n_big) ? p.chunk_small : p.chunk;
    return (int)max((int64_t)0, min(p.n - unit_start(c), len));
  };
  // block 0 owes the host the pending pass's inertia: its terms first, while no register is live
  __shared__ long long s_g[KM_SMALL_THREADS / 32][2];   // per-warp shares of those terms
  if (blockIdx.x == 0 && ps.D_in != nullptr && ps.slot != nullptr)
    km_small_inertia_terms(ps.S_in, ps.D_in, ps.C_in, ps.full_mode, k, D, ps.QB, ps.qa, ps.qg, ps.QBg, s_g);

  // first round's loads go out before anything else so they fly during the prologue
  KmBatch<D> cur;
  constexpr bool EARLY_LABELS = true;     // labels travel with their points, a round ahead
  km_load_batch<D, EARLY_LABELS, WQ>(cur, p.x + unit_start(ch) * D, p.labels + unit_start(ch), 0, chunk_len(ch));

  // ---- prologue = the previous pass's centroid update (a4), redone by every block:
  //   sums      S = S_in (+ D_in when a pass is pending; = D_in in full mode)
  //   centroid  c = S / count where the update applies and count > 0, else C_in (A4: an
  //             empty cluster keeps its centroid, /root/reference/app.mjs:322,484-488)
  // Block 0 also writes S and c back (S_out, C_out), publishes the pending pass's stats
  // and clears the delta table of the pass after this one.
  const bool pend = ps.D_in != nullptr;
  const long long changed_prev = pend ? ps.D_in[L.scal + KM_ACC_CHANGED] : 0;
  const bool upd = ps.apply_update && (!ps.needs_change || changed_prev > 0);
  for (int e = threadIdx.x; e < k * D; e += KM_SMALL_THREADS) {
    const int jj = e / D, t = e - jj * D;
    long long sa = ps.S_in[L.sa + e], sb = ps.S_in[L.sb + e], cn = ps.S_in[L.cnt + jj];
    if (pend) {
      const long long da = ps.D_in[L.sa + e], db = ps.D_in[L.sb + e], dc = ps.D_in[L.cnt + jj];
      if (ps.full_mode) { sa = da; sb = db; cn = dc; } else { sa += da; sb += db; cn += dc; }
    }
    const double c = km_small_centroid(ps, L, e, D, pend, upd);
    const float h = __double2float_rn(c);
    ncf[jj * CROW + t] = -h;
    if (HILO) ncf[jj * CROW + D + t] = -__double2float_rn(__dsub_rn(c, (double)h));
    cd[e] = c;
    if (blockIdx.x == 0) {
      ps.S_out[L.sa + e] = sa;
      ps.S_out[L.sb + e] = sb;
      if (t == 0) ps.S_out[L.cnt + jj] = cn;
      ps.C_out[e] = c;
    }
  }
  __shared__ long long s_scal[KM_ACC_SCALARS];   // block totals: changed, rescored
  if (threadIdx.x < KM_ACC_SCALARS) s_scal[threadIdx.x] = 0;
  const bool publish = blockIdx.x == 0 && pend && ps.slot != nullptr;
  KmBlockTab T;
  T.L = L;
  // a table per warp where the launcher found room (ALLPUSH passes on an unsorted store: every lane moves a point to
  // its own cluster, and a shared table would take 512 threads' compare-and-swap loops on a few hundred entries)
  T.tab = p.tab_in_smem ? stab + (size_t)(p.ntab > 1 ? (threadIdx.x >> 5) : 0) * L.scal : ps.D_out;
  if (p.tab_in_smem) km_tab_zero(stab, L.scal * p.ntab);
  if (blockIdx.x == 0) {
    if (ps.D_zero)
      for (int e = threadIdx.x; e < (int)L.len; e += KM_SMALL_THREADS) ps.D_zero[e] = 0;
  }
  __syncthreads();
  if (publish && threadIdx.x == 0) {
    long long ga = 0, gb = 0;
    for (int w = 0; w < KM_SMALL_THREADS / 32; w++) { ga += s_g[w][0]; gb += s_g[w][1]; }
    KmStatSlot sl;
    sl.it = ps.it_pending;
    sl.changed = changed_prev;
    sl.inertia = km_inertia_from_limbs(ps.Ta, ps.Tb, ps.QBt, ps.qt, ga, gb, ps.QBg, ps.qg);
    sl.rescored = ps.D_in[L.scal + KM_ACC_RESCORED];
    sl.updated = upd ? 1 : 0;
    sl.pad = 0;
    *ps.slot = sl;   // pinned host memory; visible once this kernel's event completes
  }
  // The error term of the filter and the inertia scale need max|c|.  Every centroid the
  // update produces is a mean of points, so max(max|x|, max|c0|) — known on the host when the
  // centroids were set — bounds it for every later pass: no reduction, no second barrier.
  const double scale_a = p.scale_a, scale_b = p.scale_b;
  const float tau2 = p.tau2;   // absolute part of the ambiguity threshold (2.5x margin included)
  const float rho2 = p.rho2;
  const unsigned mask = p.key_mask;

  // Candidate list of a cell (a box with centre m and half sides hx/2, hy/2).  Let b be the
  // centroid nearest to m.  For any other centroid j,
  //     f_j(x) = |x - c_j|^2 - |x - c_b|^2 = a . x + const,   a = -2 (c_j - c_b)
  // is LINEAR in x, so its minimum over the box is f_j(m) - (|a_x| hx + |a_y| hy) / 2.  If that
  // minimum is positive, b beats j at every point of the box and j can neither win nor tie
  // there: j is dropped.  Everything else (b included) is a candidate.  The test is exact
  // geometry; `margin` (1e-5 of the squared coordinate scale, ~7x the binary32 error of
  // the expressions below incl. the rounding of m) only ever keeps extra candidates.
  const float prune_M = p.cboundf + p.pr.absmax;
  const float margin = 1e-5f * prune_M * prune_M;
  auto cell_of = [&](float2 v) {
    return min(__float_as_int(fmaf(v.y, p.pr.sy, p.pr.kym)) & 0x3FFFFF, p.pr.gp - 1) * p.pr.gp +
           min(__float_as_int(fmaf(v.x, p.pr.sx, p.pr.kxm)) & 0x3FFFFF, p.pr.gp - 1);
  };
  // one thread per cell: for building the whole grid (points in the caller's order)
  auto build_cells_all = [&]() {
    const KmPrune2 &pr = p.pr;
    const float hx = pr.inv_sx, hy = pr.inv_sy;
    for (int cell = threadIdx.x; cell < pr.gp * pr.gp; cell += KM_SMALL_THREADS) {
      const int iy = cell / pr.gp, ix = cell - iy * pr.gp;
      const float mx = ((float)ix - pr.kx) * pr.inv_sx, my = ((float)iy - pr.ky) * pr.inv_sy;
      float dbest = __int_as_float(0x7f800000), bx = 0.f, by = 0.f;
      for (int j = 0; j < k; j++) {
        const float ncx = ncf[j * CROW], ncy = ncf[j * CROW + 1];   // negated centroid (hi part)
        const float dx = mx + ncx, dy = my + ncy;
        const float d2 = fmaf(dx, dx, dy * dy);
        if (d2 < dbest) { dbest = d2; bx = ncx; by = ncy; }
      }
      unsigned long long word = 0;
      int cnt = 0;
      for (int j = 0; j < k; j++) {
        const float ncx = ncf[j * CROW], ncy = ncf[j * CROW + 1];
        const float dx = mx + ncx, dy = my + ncy;
        const float fm = fmaf(dx, dx, dy * dy) - dbest;                                   // f_j(m)
        const float swing = fabsf(ncx - bx) * hx + fabsf(ncy - by) * hy;                  // (|a_x| hx + |a_y| hy) / 2
        if (fm - swing <= margin) {
          if (cnt < KM_PRUNE_CAP) word |= (unsigned long long)j << (8 * (cnt + 1));
          cnt++;
        }
      }
      cells_sm[cell] = word | (unsigned long long)(cnt > KM_PRUNE_CAP ? 0xFF : cnt);
    }
  };
  if constexpr (PRUNE) {
    if (!SORTED) {
      build_cells_all();
      __syncthreads();
    }
  }

  int changed = 0, rescored = 0;

  while (ch < p.nchunks) {
  const int cn = chunk_len(ch);
  const float *__restrict__ xb = p.x + unit_start(ch) * D;
  int32_t *__restrict__ lb = p.labels + unit_start(ch);
  for (int r0 = 0; r0 < cn; r0 += per_round) {
    // software pipeline: the next round's loads (the next chunk's first, at the end of this
    // one) are in flight during this round's arithmetic
    KmBatch<D> nxt;
    if (!EARLY_LABELS) km_load_labels<D, WQ>(cur, lb, r0, cn);
    if (r0 + per_round < cn) {
      km_load_batch<D, EARLY_LABELS, WQ>(nxt, xb, lb, r0 + per_round, cn);
    } else if (ch_next < p.nchunks) {
      km_load_batch<D, EARLY_LABELS, WQ>(nxt, p.x + unit_start(ch_next) * D, p.labels + unit_start(ch_next), 0, chunk_len(ch_next));
    }

    // ---- binary32 filter, top-2 on integer keys
    unsigned m1[PPT], m2[PPT];
#pragma unroll
    for (int q = 0; q < PPT; q++) { m1[q] = 0xFFFFFFFFu; m2[q] = 0xFFFFFFFFu; }
    bool one_candidate = false;   // warp-uniform: every point of the round has the same single candidate

    if constexpr (PRUNE) {
      // each point looks only at its cell's candidates; a cell that overflowed its word
      // (count byte 0xFF) sends its points to the binary64 scan of all k below
      const KmPrune2 &pr = p.pr;
      auto cell_of_q = [&](int q) {
        float x0, x1;
        unpk2(cur.xq[q][0], x0, x1);
        // every real point maps inside the grid; the clamp is for the zero-filled lanes past
        // the end of a ragged round (a negative index shows up as a huge masked value)
        const int ix = min(__float_as_int(fmaf(x0, pr.sx, pr.kxm)) & 0x3FFFFF, pr.gp - 1);
        const int iy = min(__float_as_int(fmaf(x1, pr.sy, pr.kym)) & 0x3FFFFF, pr.gp - 1);
        return iy * pr.gp + ix;
      };
      // Do all 32 x PPT points of the warp share one candidate word?  With the points stored in
      // cell order nearly every warp does: the candidates are then warp-uniform (one broadcast
      // LDS each, no predication), and a single-candidate word needs no filter at all.
      unsigned long long w[PPT], w0 = 0;
      bool uniform = false;
      if (SORTED && r0 + per_round <= cn) {
        // cell ids ascend along the store and every slot of a whole round is a real point: if
        // the warp's first and last point share a cell, so does everything between them
        const int ca = __shfl_sync(0xffffffffu, cell_of_q(0), 0);
        const int cz = __shfl_sync(0xffffffffu, cell_of_q(PPT - 1), 31);
        if (ca == cz) {
          uniform = true;
          w0 = cells[ca];
#pragma unroll
          for (int q = 0; q < PPT; q++) w[q] = w0;
        }
      }
      if (!uniform) {
#pragma unroll
        for (int q = 0; q < PPT; q++) w[q] = cells[cell_of_q(q)];
        w0 = __shfl_sync(0xffffffffu, w[0], 0);
        bool same = true;
#pragma unroll
        for (int q = 0; q < PPT; q++) same &= (w[q] == w0);
        uniform = __all_sync(0xffffffffu, same);
      }
      const int n0 = (int)(w0 & 0xFF);
      if (uniform && n0 != 0xFF) {
        if (n0 == 1) {
          // the only candidate wins: key = its index under distance bits 0 ("clear" below)
#pragma unroll
          for (int q = 0; q < PPT; q++) m1[q] = (unsigned)((w0 >> 8) & 0xFF);
          one_candidate = true;
        } else {
          for (int c = 0; c < n0; c++) {
            const int j = (int)((w0 >> (8 * (c + 1))) & 0xFF);
            const float *crow = ncf + j * CROW;
            const f32x2 nh = *reinterpret_cast<const f32x2 *>(crow);
            f32x2 nl = 0;
            if (HILO) nl = *reinterpret_cast<const f32x2 *>(crow + 2);
#pragma unroll
            for (int q = 0; q < PPT; q++) {
              f32x2 df = add2(cur.xq[q][0], nh);
              if (HILO) df = add2(df, nl);
              float lo, hi;
              unpk2(mul2(df, df), lo, hi);
              const unsigned key = (__float_as_uint(lo + hi) & ~mask) | (unsigned)j;
              m2[q] = min(m2[q], max(m1[q], key));
              m1[q] = min(m1[q], key);
            }
          }
        }
      } else {
        int nn[PPT], nmax = 0;
#pragma unroll
        for (int q = 0; q < PPT; q++) {
          nn[q] = (int)(w[q] & 0xFF);
          if (nn[q] == 0xFF) nn[q] = 0;   // overflow: leave m1 = m2 = max -> "ambiguous" -> full scan
          nmax = max(nmax, nn[q]);
        }
#pragma unroll
        for (int c = 0; c < KM_PRUNE_CAP; c++) {
          if (c >= nmax) break;
#pragma unroll
          for (int q = 0; q < PPT; q++) {
            if (c < nn[q]) {
              const int j = (int)((w[q] >> (8 * (c + 1))) & 0xFF);
              const float *crow = ncf + j * CROW;
              f32x2 df = add2(cur.xq[q][0], *reinterpret_cast<const f32x2 *>(crow));
              if (HILO) df = add2(df, *reinterpret_cast<const f32x2 *>(crow + 2));
              float lo, hi;
              unpk2(mul2(df, df), lo, hi);
              const unsigned key = (__float_as_uint(lo + hi) & ~mask) | (unsigned)j;
              m2[q] = min(m2[q], max(m1[q], key));
              m1[q] = min(m1[q], key);
            }
          }
        }
      }
    } else {
#pragma unroll 4
    for (int j = 0; j < k; j++) {
      const float *crow = ncf + j * CROW;
      f32x2 nh[NP > 0 ? NP : 1], nl[NP > 0 ? NP : 1];
      float th = 0.f, tl = 0.f;
      if constexpr (D == 2 && !HILO) {
        nh[0] = *reinterpret_cast<const f32x2 *>(crow);
      } else if constexpr (D == 2 && HILO) {
        const ulonglong2 c = *reinterpret_cast<const ulonglong2 *>(crow);
        nh[0] = c.x; nl[0] = c.y;
      } else {
#pragma unroll
        for (int m = 0; m < NP; m++) {
          nh[m] = pk2(crow[2 * m], crow[2 * m + 1]);
          if (HILO) nl[m] = pk2(crow[D + 2 * m], crow[D + 2 * m + 1]);
        }
        if (D == 1 || TAIL) { th = crow[D - 1]; if (HILO) tl = crow[2 * D - 1]; }
      }
#pragma unroll
      for (int q = 0; q < PPT; q++) {
        float dist;
        if constexpr (NP > 0) {
          f32x2 acc2;
#pragma unroll
          for (int m = 0; m < NP; m++) {
            f32x2 df = add2(cur.xq[q][m], nh[m]);
            if (HILO) df = add2(df, nl[m]);
            acc2 = (m == 0) ? mul2(df, df) : fma2(df, df, acc2);
          }
          float lo, hi;
          unpk2(acc2, lo, hi);
          dist = lo + hi;
          if constexpr (TAIL) {
            float dt = cur.xs[q] + th;
            if (HILO) dt += tl;
            dist = fmaf(dt, dt, dist);
          }
        } else {
          float dt = cur.xs[q] + th;
          if (HILO) dt += tl;
          dist = dt * dt;
        }
        const unsigned key = (__float_as_uint(dist) & ~mask) | (unsigned)j;
        m2[q] = min(m2[q], max(m1[q], key));
        m1[q] = min(m1[q], key);
      }
    }
    }

    // ---- per point: ambiguity test, rare binary64 re-score, inertia, commit.  A round that
    // lies wholly inside the chunk takes the copy with no bounds checks.
    auto tail = [&](auto full_tag) {
      constexpr bool FULL = decltype(full_tag)::value;
      int newlab[PPT];
      bool anypush = false, allpush = true;
      if (PRUNE && FULL && one_candidate) {
        // whole warp, one candidate: nothing to compare, nothing to measure (the inertia comes
        // from the sums) — only a look at the old labels
        const int lab = (int)m1[0];
#pragma unroll
        for (int q = 0; q < PPT; q++) {
          newlab[q] = lab;
          const bool moved = lab != cur.old[q];
          changed += moved ? 1 : 0;
          anypush |= p.incremental ? moved : true;
          allpush &= p.incremental ? moved : true;
        }
      } else {
#pragma unroll
        for (int q = 0; q < PPT; q++) {
          bool live = true;
          if (!FULL) live = r0 + km_slot<NV, PPV, WQ>(q / PPV) + (q % PPV) < cn;
          int lab = (int)(m1[q] & mask);
          float xr[D];
#pragma unroll
          for (int m = 0; m < NP; m++) unpk2(cur.xq[q][m], xr[2 * m], xr[2 * m + 1]);
          if (D == 1 || TAIL) xr[D - 1] = cur.xs[q];
          const float b1 = __uint_as_float(m1[q] & ~mask), b2 = __uint_as_float(m2[q] & ~mask);
          // unambiguous iff the gap clears the error bound; NaN/Inf fall through to the re-score.
          // No runner-up at all (k == 1, or a pruned cell with a single candidate) is unambiguous
          // by construction; no winner at all (an overflowed cell) is not.
          const bool clear = (m2[q] == 0xFFFFFFFFu) ? (m1[q] != 0xFFFFFFFFu) : ((b2 - b1) > fmaf(rho2, b2, tau2));
          if (live && !clear) {
            rescored++;
            KmRow<D> row;
#pragma unroll
            for (int t = 0; t < D; t++) row.v[t] = xr[t];
            lab = km_small_rescore<D>(row, cd, k);
          }
          if (!live) lab = cur.old[q];
          newlab[q] = lab;
          const bool moved = lab != cur.old[q];
          changed += moved ? 1 : 0;
          anypush |= live && (p.incremental ? moved : true);
          allpush &= live && (p.incremental ? moved : true);
        }
      }
      // ---- commit.  Most rounds of a converging run re-label nobody: one vote settles that.
      if (!__any_sync(0xffffffffu, anypush)) return;
      if (ALLPUSH && __all_sync(0xffffffffu, allpush)) {
        // Every point of the round moves (a first pass, the full-update mode; this branch exists only in the
        // instantiation the host picks for such passes, so the steady-state kernel pays no registers for it).
        // Nearly all of them go the same way as lane 0's first point (the store is cell-sorted; a first pass is
        // spatially coherent anyhow): each lane adds those up in registers, the warp adds the lanes up with
        // reductions, lane 0 touches the table; the odd ones out move by themselves.
        const int to0 = __shfl_sync(0xffffffffu, newlab[0], 0);
        const int from0 = p.incremental ? __shfl_sync(0xffffffffu, cur.old[0], 0) : -1;
        KmLimbs<D> acc;
#pragma unroll
        for (int t = 0; t < D; t++) { acc.a[t] = 0; acc.b[t] = 0; }
        int c = 0;
#pragma unroll
        for (int q = 0; q < PPT; q++) {
          KmRow<D> row;
#pragma unroll
          for (int m = 0; m < NP; m++) unpk2(cur.xq[q][m], row.v[2 * m], row.v[2 * m + 1]);
          if (D == 1 || TAIL) row.v[D - 1] = cur.xs[q];
          const int fr = p.incremental ? cur.old[q] : -1;
          if (newlab[q] == to0 && fr == from0) {
#pragma unroll
            for (int t = 0; t < D; t++) {
              long long a, b;
              km_fixed_split((double)row.v[t], scale_a, scale_b, a, b);
              acc.a[t] += a; acc.b[t] += b;
            }
            c++;
          } else {
            km_small_move_one<D>(T.tab, k, row, fr, newlab[q], scale_a, scale_b);
          }
        }
        km_small_warp_add<D>(T.tab, k, acc, c, from0, to0);
      } else
#pragma unroll
      for (int q = 0; q < PPT; q++) {
        bool live = true;
        if (!FULL) live = r0 + km_slot<NV, PPV, WQ>(q / PPV) + (q % PPV) < cn;
        const int lab = newlab[q];
        // queue the move (incremental: re-labelled points only; full: every live point)
        const bool push = live && (p.incremental ? (lab != cur.old[q]) : true);
        const unsigned bal = __ballot_sync(0xffffffffu, push);
        if (bal) {
          if (push) {
            unsigned *e = qbuf + ((q_head + q_count + __popc(bal & ((1u << lane) - 1u))) & (QCAP - 1)) * QW;
            float xr[D];
#pragma unroll
            for (int m = 0; m < NP; m++) unpk2(cur.xq[q][m], xr[2 * m], xr[2 * m + 1]);
            if (D == 1 || TAIL) xr[D - 1] = cur.xs[q];
#pragma unroll
            for (int t = 0; t < D; t++) e[t] = __float_as_uint(xr[t]);
            e[D] = (unsigned)(p.incremental ? cur.old[q] : -1);
            e[D + 1] = (unsigned)lab;
          }
          q_count += __popc(bal);
          __syncwarp();
          if (q_count >= 32) {
            km_small_drain_warp<D>(T.tab, k, qbuf + ((q_head + lane) & (QCAP - 1)) * QW, scale_a, scale_b);
            q_head = (q_head + 32) & (QCAP - 1);
            q_count -= 32;
            __syncwarp();
          }
        }
      }
#pragma unroll
      for (int v = 0; v < NV; v++) {
        bool any = false;
#pragma unroll
        for (int u = 0; u < PPV; u++) any |= newlab[v * PPV + u] != cur.old[v * PPV + u];
        if (!any) continue;
        const int i0 = r0 + km_slot<NV, PPV, WQ>(v);
        if (FULL || i0 + PPV <= cn) {
          if constexpr (PPV == 4) {
            *reinterpret_cast<int4 *>(lb + i0) =
                make_int4(newlab[v * 4], newlab[v * 4 + 1], newlab[v * 4 + 2], newlab[v * 4 + 3]);
          } else if constexpr (PPV == 2) {
            *reinterpret_cast<int2 *>(lb + i0) = make_int2(newlab[v * 2], newlab[v * 2 + 1]);
          } else {
            lb[i0] = newlab[v];
          }
        } else {
#pragma unroll
          for (int u = 0; u < PPV; u++)
            if (i0 + u < cn) lb[i0 + u] = newlab[v * PPV + u];
        }
      }
    };
    if (r0 + per_round <= cn) tail(std::true_type{}); else tail(std::false_type{});
    cur = nxt;
  }
  ch = ch_next;
  ch_next = WQ ? grab() : ch_next + gridDim.x;
  }

  // whatever is still queued (fewer than 32 entries per warp)
  if (lane < q_count) km_small_drain<D>(T.tab, k, qbuf + ((q_head + lane) & (QCAP - 1)) * QW, scale_a, scale_b);

  // ---- block epilogue.  Counts go through one REDUX per warp.
  {
    const int wc = __reduce_add_sync(0xffffffffu, changed), wr = __reduce_add_sync(0xffffffffu, rescored);
    if (lane == 0) {
      if (wc) km_atomic_add_ll(&s_scal[KM_ACC_CHANGED], wc);
      if (wr) km_atomic_add_ll(&s_scal[KM_ACC_RESCORED], wr);
    }
  }
  __syncthreads();
  if (p.tab_in_smem) {
    if (p.ntab == 1) km_tab_flush(stab, ps.D_out, L.scal);
    else
      for (int64_t e = threadIdx.x; e < L.scal; e += KM_SMALL_THREADS) {
        long long v = 0;
        for (int w = 0; w < p.ntab; w++) v += stab[(size_t)w * L.scal + e];
        if (v != 0) km_atomic_add_ll(ps.D_out + e, v);
      }
  }
  if (threadIdx.x < KM_ACC_SCALARS && s_scal[threadIdx.x] != 0)
    km_atomic_add_ll(ps.D_out + L.scal + threadIdx.x, s_scal[threadIdx.x]);
}

static size_t small_centroid_smem(int32_t d, int32_t k, bool hilo) {
  size_t floats = ((size_t)k * d * (hilo ? 2 : 1) + 3) & ~(size_t)3;
  return floats * sizeof(float) + (size_t)k * d * sizeof(double);
}

bool km_small_path_supported(int32_t d, int32_t k, int dtype, int max_smem_optin) {
  if (dtype != KM_DTYPE_F32 || d < 1 || d > 8 || k > 4096) return false;
  if ((int64_t)k * d > 8192) return false;   // every block rebuilds the centroids in its prologue
  return small_centroid_smem(d, k, true) + 4096 <= (size_t)max_smem_optin / 2;
}

template <int D, bool HILO, bool PRUNE, bool SORTED, bool ALLPUSH>
static cudaError_t launch_small_ap(const KmAssignArgs &a, int64_t *launches) {
  const KmAccLayout L = km_acc_layout(a.k, D);
  size_t smem = small_centroid_smem(D, a.k, HILO);
  KmSmallParams p;
  p.x = (const float *)a.x; p.n = a.n; p.k = a.k; p.labels = a.labels;
  p.incremental = a.incremental; p.exact_inertia = a.exact_inertia;
  p.ps = a.ps;
  p.tab_in_smem = 0;
  const size_t tab_bytes = (size_t)L.scal * sizeof(long long);
  if (tab_bytes <= 48 * 1024 && smem + tab_bytes + 4096 <= (size_t)a.max_smem_optin / 2) {
    p.tab_in_smem = 1;
    smem += tab_bytes;
  }
  int bits = 0;
  while ((1 << bits) < a.k) bits++;
  p.key_mask = (1u << bits) - 1u;
  // (D+6) u for the arithmetic, 2^-(23-bits) for the key truncation, x2.5 margin
  p.rho2 = 2.5f * ((float)(D + 6) * 0x1.0p-24f + ldexpf(1.0f, bits - 23));
  p.sorted = 0;
  p.cells_g = nullptr;
  if (PRUNE) {
    p.pr = a.pr;
    p.sorted = SORTED ? 1 : 0;
    p.cells_g = a.cells_g;
    if (!p.sorted) smem += (size_t)a.pr.gp * a.pr.gp * sizeof(unsigned long long);   // the block builds its own copy
  }
  smem += (size_t)(KM_SMALL_THREADS / 32) * 64 * (D + 2) * sizeof(unsigned);   // per-warp move queues
  p.ntab = 1;
  if (ALLPUSH && p.tab_in_smem && smem + (KM_SMALL_THREADS / 32 - 1) * tab_bytes + 4096 <= (size_t)a.max_smem_optin / 2) {
    p.ntab = KM_SMALL_THREADS / 32;   // a table per warp
    smem += (KM_SMALL_THREADS / 32 - 1) * tab_bytes;
  }
  auto kern = km_assign_small_kernel<D, HILO, PRUNE, SORTED, ALLPUSH>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, KM_SMALL_THREADS, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  // balanced contiguous chunks: one resident wave of blocks, every block the same share
  const int64_t per_round = (int64_t)KM_SMALL_THREADS * KmSmallCfg<D>::PPT;
  const int64_t cap = (int64_t)a.sm_count * per_sm;
  int64_t want = (a.n + per_round - 1) / per_round;
  if (want < 1) want = 1;
  int64_t grid64 = want < cap ? want : cap;
  int64_t chunk = (a.n + grid64 - 1) / grid64;
  chunk = (chunk + 3) & ~(int64_t)3;
  if (chunk < 4) chunk = 4;
  const int64_t max_chunk = ((int64_t)1 << 30) / D;   // 32-bit offsets inside a chunk
  if (chunk > max_chunk) chunk = max_chunk & ~(int64_t)3;
  grid64 = (a.n + chunk - 1) / chunk;
  if (grid64 < 1) grid64 = 1;   // n == 0: one block still runs the prologue (publish / update)
  int grid = (int)grid64;
  p.chunk = chunk;
  p.nchunks = grid64;
  p.n_big = grid64;
  p.chunk_small = chunk;
  p.work_counter = a.work_counter;
  if (SORTED) {
    // units of 16 warp-rounds (2048 points at d = 2), taken dynamically by the warps of one resident wave
    const int64_t unit = (int64_t)32 * KmSmallCfg<D>::PPT * 16;
    p.chunk = unit;
    const int64_t nunits = (a.n + unit - 1) / unit;
    const int64_t warps_wanted = nunits < 1 ? 1 : nunits;
    const int64_t blocks_wanted = (warps_wanted + KM_SMALL_THREADS / 32 - 1) / (KM_SMALL_THREADS / 32);
    grid = (int)(blocks_wanted < cap ? blocks_wanted : cap);
    // guided: the last unit's worth of points per warp goes out in units of two warp-rounds, so that the warps
    // finish within a few hundred points of each other (at 13 units per warp a whole unit of slack is 8 % of the pass)
    const int64_t warps = (int64_t)grid * (KM_SMALL_THREADS / 32);
    int64_t tail_units = warps < nunits / 2 ? warps : nunits / 2;
    p.n_big = nunits - tail_units;
    p.chunk_small = unit / 8;
    const int64_t rest = a.n - p.n_big * unit;
    p.nchunks = p.n_big + (rest > 0 ? (rest + p.chunk_small - 1) / p.chunk_small : 0);
  }
  p.scale_a = a.scale_a;
  p.scale_b = a.scale_b;
  p.tau2 = HILO ? a.h_tau_hilo : a.h_tau_hi;
  p.cboundf = (float)a.h_cbound * 1.0000002f;
  if (PRUNE && p.sorted && a.n > 0) {
    // this pass's candidate words, for every cell, in front of the assignment kernel
    const int cellsn = a.pr.gp * a.pr.gp;
    const int gblocks = (cellsn + KM_GRID_THREADS / 32 - 1) / (KM_GRID_THREADS / 32);
    km_prune_grid_kernel<<<gblocks < 2 * a.sm_count ? gblocks : 2 * a.sm_count, KM_GRID_THREADS,
                           (size_t)a.k * 2 * sizeof(float), a.stream>>>(a.ps, a.k, a.pr, p.cboundf, a.cells_g, p.work_counter);
    (*launches)++;
  }
  kern<<<grid, KM_SMALL_THREADS, smem, a.stream>>>(p);
  (*launches)++;
  return cudaGetLastError();
}

// a pass in which every point is accumulated (labels all -1, or the full-update mode) runs the instantiation with
// the warp-aggregated commit; the steady-state instantiation does not carry that code
template <int D, bool HILO, bool PRUNE, bool SORTED>
static cudaError_t launch_small(const KmAssignArgs &a, int64_t *launches) {
  return a.all_push ? launch_small_ap<D, HILO, PRUNE, SORTED, true>(a, launches) : launch_small_ap<D, HILO, PRUNE, SORTED, false>(a, launches);
}

template <bool HILO>
static cudaError_t dispatch_small(const KmAssignArgs &a, int64_t *launches) {
  switch (a.d) {
    case 1: return launch_small<1, HILO, false, false>(a, launches);
    case 2:
      if (!a.prune) return launch_small<2, HILO, false, false>(a, launches);
      return (a.sorted && a.cells_g && a.work_counter) ? launch_small<2, HILO, true, true>(a, launches) : launch_small<2, HILO, true, false>(a, launches);
    case 3: return launch_small<3, HILO, false, false>(a, launches);
    case 4: return launch_small<4, HILO, false, false>(a, launches);
    case 5: return launch_small<5, HILO, false, false>(a, launches);
    case 6: return launch_small<6, HILO, false, false>(a, launches);
    case 7: return launch_small<7, HILO, false, false>(a, launches);
    case 8: return launch_small<8, HILO, false, false>(a, launches);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t km_launch_assign_small(const KmAssignArgs &a, int64_t *launches) {
  if ((reinterpret_cast<uintptr_t>(a.x) & 15) || (reinterpret_cast<uintptr_t>(a.labels) & 15))
    return cudaErrorMisalignedAddress;
  return a.hilo ? dispatch_small<true>(a, launches) : dispatch_small<false>(a, launches);
}

// d == 2 candidate grid for this job's bounding box; false = do not prune this shape
bool km_small_prune_setup(int32_t d, int32_t k, int64_t n_local, int64_t min_points, int sm_count, const double bbox[4],
                          int fine, KmPrune2 *out) {
  if (d != 2 || k < 16 || k > 256) return false;
  // below ~32 points per thread the per-block grid build would not pay off
  const int64_t need = min_points > 0 ? min_points : (int64_t)sm_count * 2 * KM_SMALL_THREADS * 32;
  if (n_local < need) return false;
  int g = 16;
  while (g < 64 && g * g < 64 * k) g *= 2;   // ~8 cells across a Voronoi cell, capped by shared memory
  // Cell-sorted store (`fine`): the words live in global memory, so the grid may be finer — more cells see a
  // single candidate — as long as a cell's run of points stays long against a warp-round (>= 4096 on average).
  if (fine)
    while (g < 256 && (int64_t)(2 * g) * (2 * g) * 4096 <= n_local) g *= 2;
  const double wx = bbox[1] - bbox[0], wy = bbox[3] - bbox[2];
  if (!(wx > 0.0) || !(wy > 0.0)) return false;
  const double sx = g / (wx * (1.0 + 1e-6)), sy = g / (wy * (1.0 + 1e-6));
  const double kx = ceil(0.5 - bbox[0] * sx) + 1.0, ky = ceil(0.5 - bbox[2] * sy) + 1.0;
  if (fabs(kx) > 1e6 || fabs(ky) > 1e6) return false;   // box far from the origin relative to its size
  out->gp = g + 4;
  out->sx = (float)sx; out->sy = (float)sy;
  out->kx = (float)kx; out->ky = (float)ky;
  out->kxm = (float)kx + KM_PRUNE_MAGIC; out->kym = (float)ky + KM_PRUNE_MAGIC;
  out->inv_sx = 1.0f / out->sx; out->inv_sy = 1.0f / out->sy;
  out->r = 0.5f * sqrtf(out->inv_sx * out->inv_sx + out->inv_sy * out->inv_sy);
  out->absmax = (float)fmax(fmax(fabs(bbox[0]), fabs(bbox[1])), fmax(fabs(bbox[2]), fabs(bbox[3])));
  // every point must land inside [0, gp): check the corners with the kernel's own arithmetic
  const float xs[2] = {(float)bbox[0], (float)bbox[1]}, ys[2] = {(float)bbox[2], (float)bbox[3]};
  for (int q = 0; q < 2; q++) {
    const float fx = fmaf(xs[q], out->sx, out->kx), fy = fmaf(ys[q], out->sy, out->ky);
    if (!(fx >= 0.5f && fx < out->gp - 0.5f && fy >= 0.5f && fy < out->gp - 0.5f)) return false;
  }
  return true;
}
