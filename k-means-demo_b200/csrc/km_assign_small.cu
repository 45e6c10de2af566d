// km_assign_small.cu — KM_PATH_FILTER_F32 for small d (1..8): the bandwidth-bound path
// (BASELINE.json configs[1] and [4]: d=2, k=8 / k=64).
//
// One thread per point (vectorised so each lane moves 16 bytes when d is 1, 2 or 4),
// centroids staged once per block in shared memory as binary32 hi/lo pairs plus the
// binary64 masters.  Every distance is first taken in binary32; the point is re-scored in
// binary64 (the oracle's exact operation order) only when the gap between its two best
// binary32 distances is inside the rigorous error bound of DESIGN.md §5.  Labels are
// therefore the oracle's, bit for bit, at binary32 cost.  The label compare, the move of
// re-labelled points between the order-free per-cluster sums, the change count and the
// inertia are fused into the same pass: x is read exactly once per iteration.
#include "km_commit.cuh"
#include "km_internal.h"

#define KM_SMALL_THREADS 256

template <int D>
struct KmSmallCfg {
  static constexpr int PPT = (D == 1) ? 4 : (D == 2) ? 2 : 1;  // points per thread per tile
};

template <int D>
__global__ void __launch_bounds__(KM_SMALL_THREADS)
km_assign_small_kernel(const float *__restrict__ x, int64_t n, int32_t k, int32_t *__restrict__ labels,
                       const double *__restrict__ c64, const float *__restrict__ chi, const float *__restrict__ clo,
                       long long *__restrict__ acc, const KmState *__restrict__ st, int incremental,
                       int exact_inertia, int tab_in_smem) {
  constexpr int PPT = KmSmallCfg<D>::PPT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float *cf = reinterpret_cast<float *>(smem_raw);                    // [k][2D]: hi[D], lo[D]
  double *cd = reinterpret_cast<double *>(cf + (size_t)k * 2 * D);    // [k][D]
  long long *stab = reinterpret_cast<long long *>(cd + (size_t)k * D);
  const KmAccLayout L = km_acc_layout(k, D);

  for (int e = threadIdx.x; e < k * D; e += KM_SMALL_THREADS) {
    const int j = e / D, t = e - j * D;
    cf[j * 2 * D + t] = chi[e];
    cf[j * 2 * D + D + t] = clo[e];
    cd[e] = c64[e];
  }
  KmBlockTab T;
  T.L = L;
  T.tab = tab_in_smem ? stab : acc;
  if (tab_in_smem) km_tab_zero(stab, L.scal);
  __syncthreads();

  const double scale_a = st->scale_a, scale_b = st->scale_b, inertia_scale = st->inertia_scale;
  const float tau2 = st->tau_hilo;                              // already carries the 2.5x margin
  constexpr float rho2 = 2.5f * (float)(D + 6) * 0x1.0p-24f;
  KmThreadScalars S;

  const int64_t per_tile = (int64_t)KM_SMALL_THREADS * PPT;
  const int64_t ntiles = (n + per_tile - 1) / per_tile;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = (tile * KM_SMALL_THREADS + threadIdx.x) * PPT;
    float xv[PPT][D];
    int old[PPT];
    bool live[PPT];
    if (base + PPT <= n) {
      // full group: 16-byte loads where the shape allows
      if constexpr (D == 1 || D == 2 || D == 4) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x + base * D));
        const float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int p = 0; p < PPT; p++)
#pragma unroll
          for (int t = 0; t < D; t++) xv[p][t] = f[p * D + t];
      } else {
#pragma unroll
        for (int t = 0; t < D; t++) xv[0][t] = __ldg(x + base * D + t);
      }
      if constexpr (PPT == 4) {
        const int4 o = *reinterpret_cast<const int4 *>(labels + base);
        old[0] = o.x; old[1] = o.y; old[2] = o.z; old[3] = o.w;
      } else if constexpr (PPT == 2) {
        const int2 o = *reinterpret_cast<const int2 *>(labels + base);
        old[0] = o.x; old[1] = o.y;
      } else {
        old[0] = labels[base];
      }
#pragma unroll
      for (int p = 0; p < PPT; p++) live[p] = true;
    } else {
#pragma unroll
      for (int p = 0; p < PPT; p++) {
        live[p] = base + p < n;
        old[p] = live[p] ? labels[base + p] : 0;
#pragma unroll
        for (int t = 0; t < D; t++) xv[p][t] = live[p] ? x[(base + p) * D + t] : 0.0f;
      }
    }

    float b1[PPT], b2[PPT];
    int i1[PPT];
#pragma unroll
    for (int p = 0; p < PPT; p++) { b1[p] = __int_as_float(0x7f800000); b2[p] = b1[p]; i1[p] = 0; }

#pragma unroll 4
    for (int j = 0; j < k; j++) {
      float hi[D], lo[D];
      if constexpr (D == 2) {
        const float4 c = *reinterpret_cast<const float4 *>(cf + j * 4);
        hi[0] = c.x; hi[1] = c.y; lo[0] = c.z; lo[1] = c.w;
      } else if constexpr (D == 4) {
        const float4 h = *reinterpret_cast<const float4 *>(cf + j * 8);
        const float4 l = *reinterpret_cast<const float4 *>(cf + j * 8 + 4);
        hi[0] = h.x; hi[1] = h.y; hi[2] = h.z; hi[3] = h.w;
        lo[0] = l.x; lo[1] = l.y; lo[2] = l.z; lo[3] = l.w;
      } else {
#pragma unroll
        for (int t = 0; t < D; t++) { hi[t] = cf[j * 2 * D + t]; lo[t] = cf[j * 2 * D + D + t]; }
      }
#pragma unroll
      for (int p = 0; p < PPT; p++) {
        float dj = 0.0f;
#pragma unroll
        for (int t = 0; t < D; t++) {
          const float a = __fsub_rn(xv[p][t], hi[t]);
          const float b = __fsub_rn(a, lo[t]);
          dj = fmaf(b, b, dj);
        }
        const bool lt = dj < b1[p];
        b2[p] = lt ? b1[p] : fminf(b2[p], dj);
        i1[p] = lt ? j : i1[p];
        b1[p] = lt ? dj : b1[p];
      }
    }

    int newlab[PPT];
    bool any_change = false;
#pragma unroll
    for (int p = 0; p < PPT; p++) {
      newlab[p] = old[p];
      if (!live[p]) continue;
      int lab = i1[p];
      double best;
      // unambiguous iff the gap clears the error bound; written so NaN/Inf fall to the re-score
      const bool clear = (b2[p] - b1[p]) > (rho2 * b2[p] + tau2);
      if (!clear) {
        S.rescored++;
        best = __longlong_as_double(0x7FF0000000000000LL);
        lab = 0;
        for (int j = 0; j < k; j++) {
          const double dj = km_dist2_f64(xv[p], cd + (size_t)j * D, D);
          if (dj < best) { best = dj; lab = j; }
        }
      } else {
        best = exact_inertia ? km_dist2_f64(xv[p], cd + (size_t)lab * D, D) : (double)b1[p];
      }
      newlab[p] = lab;
      any_change |= (lab != old[p]);
      km_commit_point(T, S, xv[p], D, old[p], lab, best, incremental, scale_a, scale_b, inertia_scale);
    }
    if (any_change) {
      if (base + PPT <= n) {
        if constexpr (PPT == 4) {
          *reinterpret_cast<int4 *>(labels + base) = make_int4(newlab[0], newlab[1], newlab[2], newlab[3]);
        } else if constexpr (PPT == 2) {
          *reinterpret_cast<int2 *>(labels + base) = make_int2(newlab[0], newlab[1]);
        } else {
          labels[base] = newlab[0];
        }
      } else {
#pragma unroll
        for (int p = 0; p < PPT; p++)
          if (live[p]) labels[base + p] = newlab[p];
      }
    }
  }
  __syncthreads();
  if (tab_in_smem) km_tab_flush(stab, acc, L.scal);
  km_flush_scalars(S, acc, L);
}

static size_t small_centroid_smem(int32_t d, int32_t k) { return (size_t)k * d * (2 * sizeof(float) + sizeof(double)); }

bool km_small_path_supported(int32_t d, int32_t k, int dtype, int max_smem_optin) {
  if (dtype != KM_DTYPE_F32 || d < 1 || d > 8) return false;
  return small_centroid_smem(d, k) + 4096 <= (size_t)max_smem_optin;
}

template <int D>
static cudaError_t launch_small(const KmAssignArgs &a, int64_t *launches) {
  const KmAccLayout L = km_acc_layout(a.k, D);
  size_t smem = small_centroid_smem(D, a.k);
  int tab_in_smem = 0;
  const size_t tab_bytes = (size_t)L.scal * sizeof(long long);
  if (tab_bytes <= 64 * 1024 && smem + tab_bytes + 4096 <= (size_t)a.max_smem_optin) {
    tab_in_smem = 1;
    smem += tab_bytes;
  }
  auto kern = km_assign_small_kernel<D>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, KM_SMALL_THREADS, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  const int64_t per_tile = (int64_t)KM_SMALL_THREADS * KmSmallCfg<D>::PPT;
  const int64_t ntiles = (a.n + per_tile - 1) / per_tile;
  const int64_t cap = (int64_t)a.sm_count * per_sm;   // one resident wave, grid-stride over tiles
  const int grid = (int)(ntiles < cap ? ntiles : cap);
  kern<<<grid, KM_SMALL_THREADS, smem, a.stream>>>((const float *)a.x, a.n, a.k, a.labels, a.c64, a.chi, a.clo, a.acc,
                                                   a.state, a.incremental, a.exact_inertia, tab_in_smem);
  (*launches)++;
  return cudaGetLastError();
}

cudaError_t km_launch_assign_small(const KmAssignArgs &a, int64_t *launches) {
  if (a.n == 0) return cudaSuccess;
  if ((reinterpret_cast<uintptr_t>(a.x) & 15) || (reinterpret_cast<uintptr_t>(a.labels) & 15))
    return cudaErrorMisalignedAddress;
  switch (a.d) {
    case 1: return launch_small<1>(a, launches);
    case 2: return launch_small<2>(a, launches);
    case 3: return launch_small<3>(a, launches);
    case 4: return launch_small<4>(a, launches);
    case 5: return launch_small<5>(a, launches);
    case 6: return launch_small<6>(a, launches);
    case 7: return launch_small<7>(a, launches);
    case 8: return launch_small<8>(a, launches);
    default: return cudaErrorInvalidValue;
  }
}
