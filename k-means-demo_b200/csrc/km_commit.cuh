// km_commit.cuh — what every assignment kernel does once a point's label is known:
// compare with the old label, move the point between the order-free per-cluster sums,
// add the winner distance to the inertia limbs.  (a3/a4 fused)
#pragma once
#include "km_common.cuh"

struct KmThreadScalars {
  long long changed = 0, ia = 0, ib = 0, rescored = 0;
};

// Per-cluster table the block accumulates into: shared memory when it fits, else the
// global table directly.  Integer atomics commute exactly, so either is deterministic.
struct KmBlockTab {
  long long *tab;   // generic pointer (shared or global)
  KmAccLayout L;
};

__device__ __forceinline__ void km_tab_zero(long long *stab, int64_t len) {
  for (int64_t i = threadIdx.x; i < len; i += blockDim.x) stab[i] = 0;
}

// add the non-zero entries of the block's shared table into the global one
__device__ __forceinline__ void km_tab_flush(const long long *stab, long long *gacc, int64_t len) {
  for (int64_t i = threadIdx.x; i < len; i += blockDim.x) {
    long long v = stab[i];
    if (v != 0) km_atomic_add_ll(gacc + i, v);
  }
}

// move one point: +x into cluster `to`, and -x out of cluster `from` when from >= 0
template <typename XT>
__device__ __forceinline__ void km_move_point(const KmBlockTab &T, const XT *xrow, int d, int to, int from,
                                              double scale_a, double scale_b) {
  for (int t = 0; t < d; t++) {
    long long a, b;
    km_fixed_split((double)xrow[t], scale_a, scale_b, a, b);
    km_atomic_add_ll(T.tab + T.L.sa + (int64_t)to * d + t, a);
    km_atomic_add_ll(T.tab + T.L.sb + (int64_t)to * d + t, b);
    if (from >= 0) {
      km_atomic_add_ll(T.tab + T.L.sa + (int64_t)from * d + t, -a);
      km_atomic_add_ll(T.tab + T.L.sb + (int64_t)from * d + t, -b);
    }
  }
  km_atomic_add_ll(T.tab + T.L.cnt + to, 1);
  if (from >= 0) km_atomic_add_ll(T.tab + T.L.cnt + from, -1);
}

template <typename XT>
__device__ __forceinline__ void km_commit_point(const KmBlockTab &T, KmThreadScalars &S, const XT *xrow, int d,
                                                int old, int lab, double best, int incremental, double scale_a,
                                                double scale_b, double inertia_scale) {
  if (lab != old) S.changed++;
  if (incremental) {
    if (lab != old) km_move_point(T, xrow, d, lab, old, scale_a, scale_b);
  } else {
    km_move_point(T, xrow, d, lab, -1, scale_a, scale_b);
  }
  long long a, b;
  km_fixed_split(best, inertia_scale, scale_b, a, b);
  S.ia += a;
  S.ib += b;
}

// the same without the per-point inertia (the context API takes the inertia from the sums,
// DESIGN.md §4.1): label compare and the move between the order-free sums
template <typename XT>
__device__ __forceinline__ void km_commit_point_sums(const KmBlockTab &T, KmThreadScalars &S, const XT *xrow, int d,
                                                     int old, int lab, int incremental, double scale_a, double scale_b) {
  if (lab != old) S.changed++;
  if (incremental) {
    if (lab != old) km_move_point(T, xrow, d, lab, old, scale_a, scale_b);
  } else {
    km_move_point(T, xrow, d, lab, -1, scale_a, scale_b);
  }
}

// block-wide sum of the per-thread scalars, then four global atomics per block
__device__ __forceinline__ void km_flush_scalars(const KmThreadScalars &S, long long *gacc, const KmAccLayout &L) {
  __shared__ long long red[4][32];
  long long v[4] = {S.changed, S.ia, S.ib, S.rescored};
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    long long w = km_warp_sum_ll(v[q]);
    if (lane == 0) red[q][warp] = w;
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
      long long w = lane < nwarps ? red[q][lane] : 0;
      w = km_warp_sum_ll(w);
      if (lane == 0 && w != 0) km_atomic_add_ll(gacc + L.scal + q, w);
    }
  }
}
