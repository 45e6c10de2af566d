// km_kernels_misc.cu — store scan, blob generator, label fill, and the finalize kernel
// (order-free sums -> centroids, empty-cluster rule, per-iteration stats).
#include "km_finalize.cuh"
#include "km_internal.h"

// ------------------------------------------------------------------ a1: scan
// max|x| and a non-finite flag over the shard; one pass at upload time.
template <typename T>
__global__ void __launch_bounds__(256) km_scan_kernel(const T *__restrict__ x, int64_t count,
                                                      unsigned long long *maxabs_bits, int *nonfinite) {
  double m = 0.0;
  int bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * blockDim.x) {
    double v = fabs((double)x[i]);
    if (!(v <= 1.7976931348623157e308)) bad = 1;  // NaN or Inf
    else m = fmax(m, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    bad |= __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (m > 0.0) atomicMax(maxabs_bits, (unsigned long long)__double_as_longlong(m));
    if (bad) atomicOr(nonfinite, 1);
  }
}

cudaError_t km_launch_scan(const void *x, int dtype, int64_t count, double *maxabs_out, int *nonfinite_out,
                           int sm_count, cudaStream_t s, int64_t *launches) {
  cudaError_t e = cudaMemsetAsync(maxabs_out, 0, sizeof(double), s);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(nonfinite_out, 0, sizeof(int), s);
  if (e != cudaSuccess) return e;
  if (count == 0) return cudaSuccess;
  int64_t want = (count + 256 * 8 - 1) / (256 * 8);
  int grid = (int)(want < (int64_t)sm_count * 8 ? (want < 1 ? 1 : want) : (int64_t)sm_count * 8);
  if (dtype == KM_DTYPE_F32)
    km_scan_kernel<float><<<grid, 256, 0, s>>>((const float *)x, count, (unsigned long long *)maxabs_out, nonfinite_out);
  else
    km_scan_kernel<double><<<grid, 256, 0, s>>>((const double *)x, count, (unsigned long long *)maxabs_out, nonfinite_out);
  (*launches)++;
  return cudaGetLastError();
}

// T = sum_i |x_i|^2 as order-free limbs: every x^2 is put on the grid
// 2^-(qt+QBt) by itself (exact for binary32 points in the usual range, otherwise rounded once,
// deterministically), the limbs are summed as integers.  binary64 points: x^2 = p + e exactly
// (two-product), both parts go on the grid.
template <typename T>
__global__ void __launch_bounds__(256) km_sumsq_kernel(const T *__restrict__ x, int64_t count, double scale_t, double scale_tb,
                                                       unsigned long long *__restrict__ out2) {
  long long sa = 0, sb = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = (double)x[i];
    const double p = __dmul_rn(v, v);
    long long a, b;
    km_fixed_split(p, scale_t, scale_tb, a, b);
    sa += a; sb += b;
    if (sizeof(T) == 8) {
      km_fixed_split(__fma_rn(v, v, -p), scale_t, scale_tb, a, b);
      sa += a; sb += b;
    }
  }
  sa = km_warp_sum_ll(sa);
  sb = km_warp_sum_ll(sb);
  if ((threadIdx.x & 31) == 0) {
    if (sa) atomicAdd(out2, (unsigned long long)sa);
    if (sb) atomicAdd(out2 + 1, (unsigned long long)sb);
  }
}

cudaError_t km_launch_sumsq(const void *x, int dtype, int64_t count, double scale_t, double scale_tb, long long *out2,
                            int sm_count, cudaStream_t s, int64_t *launches) {
  cudaError_t e = cudaMemsetAsync(out2, 0, 2 * sizeof(long long), s);
  if (e != cudaSuccess || count == 0) return e;
  int64_t want = (count + 256 * 8 - 1) / (256 * 8);
  int grid = (int)(want < (int64_t)sm_count * 8 ? (want < 1 ? 1 : want) : (int64_t)sm_count * 8);
  if (dtype == KM_DTYPE_F32)
    km_sumsq_kernel<float><<<grid, 256, 0, s>>>((const float *)x, count, scale_t, scale_tb, (unsigned long long *)out2);
  else
    km_sumsq_kernel<double><<<grid, 256, 0, s>>>((const double *)x, count, scale_t, scale_tb, (unsigned long long *)out2);
  (*launches)++;
  return cudaGetLastError();
}

// bounding box of d == 2 float32 points (for the candidate grid): ordered-integer min/max
__device__ __forceinline__ unsigned km_f2ord(float f) {
  unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__global__ void __launch_bounds__(256) km_bbox2_kernel(const float2 *__restrict__ x, int64_t n, unsigned *out4) {
  unsigned lox = 0xFFFFFFFFu, hix = 0u, loy = 0xFFFFFFFFu, hiy = 0u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float2 v = x[i];
    const unsigned a = km_f2ord(v.x), b = km_f2ord(v.y);
    lox = min(lox, a); hix = max(hix, a); loy = min(loy, b); hiy = max(hiy, b);
  }
  lox = __reduce_min_sync(0xffffffffu, lox); hix = __reduce_max_sync(0xffffffffu, hix);
  loy = __reduce_min_sync(0xffffffffu, loy); hiy = __reduce_max_sync(0xffffffffu, hiy);
  if ((threadIdx.x & 31) == 0) {
    atomicMin(out4 + 0, lox); atomicMax(out4 + 1, hix); atomicMin(out4 + 2, loy); atomicMax(out4 + 3, hiy);
  }
}

// out4 (device, 4 unsigned): ordered encodings of xmin, xmax, ymin, ymax
cudaError_t km_launch_bbox2(const float *x, int64_t n, unsigned *out4, int sm_count, cudaStream_t s, int64_t *launches) {
  const unsigned init[4] = {0xFFFFFFFFu, 0u, 0xFFFFFFFFu, 0u};
  cudaError_t e = cudaMemcpyAsync(out4, init, sizeof(init), cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess || n == 0) return e;
  int64_t want = (n + 256 * 8 - 1) / (256 * 8);
  int grid = (int)(want < (int64_t)sm_count * 8 ? want : (int64_t)sm_count * 8);
  km_bbox2_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const float2 *>(x), n, out4);
  (*launches)++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ A8: generator
__global__ void __launch_bounds__(256) km_gen_blobs_kernel(float *__restrict__ x, int64_t n, int64_t i0, int32_t d,
                                                           int32_t k_true, uint64_t seed, double sigma) {
  int64_t total = n * (int64_t)d;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e / d;
    int32_t t = (int32_t)(e - r * d);
    int64_t i = i0 + r;
    int64_t b = i % k_true;
    uint64_t hc = km_hash3(seed ^ KM_CENTRE_TAG, (uint64_t)b, (uint64_t)t);
    double u = (double)(hc >> 11) * 0x1.0p-53;
    double mu = __dsub_rn(__dmul_rn(u, 20.0), 10.0);
    uint32_t sacc = 0;
#pragma unroll
    for (int q = 0; q < 3; q++) {
      uint64_t h = km_hash3(seed, (uint64_t)i, (uint64_t)(3 * (int64_t)t + q));
      sacc += (uint32_t)(h & 0xFFFF) + (uint32_t)((h >> 16) & 0xFFFF) + (uint32_t)((h >> 32) & 0xFFFF) +
              (uint32_t)(h >> 48);
    }
    double z = __dsub_rn((double)sacc * 0x1.0p-16, 6.0);
    double v = __dadd_rn(mu, __dmul_rn(sigma, z));
    x[e] = __double2float_rn(v);
  }
}

cudaError_t km_launch_gen_blobs(float *x, int64_t n, int64_t i0, int32_t d, int32_t k_true, uint64_t seed,
                                double sigma, int sm_count, cudaStream_t s, int64_t *launches) {
  int64_t total = n * (int64_t)d;
  if (total == 0) return cudaSuccess;
  int64_t want = (total + 255) / 256;
  int64_t cap = (int64_t)sm_count * 16;
  int grid = (int)(want < cap ? want : cap);
  km_gen_blobs_kernel<<<grid, 256, 0, s>>>(x, n, i0, d, k_true, seed, sigma);
  (*launches)++;
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) km_fill_i32_kernel(int32_t *p, int64_t n, int32_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

cudaError_t km_launch_fill_i32(int32_t *p, int64_t n, int32_t v, int sm_count, cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  int64_t want = (n + 255) / 256;
  int64_t cap = (int64_t)sm_count * 16;
  int grid = (int)(want < cap ? want : cap);
  km_fill_i32_kernel<<<grid, 256, 0, s>>>(p, n, v);
  (*launches)++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ a4: finalize
__global__ void __launch_bounds__(1024) km_finalize_kernel(const KmFinalizeArgs f) { km_finalize_block(f); }
__global__ void __launch_bounds__(256) km_finalize_grid_kernel(const KmFinalizeArgs f) { km_finalize_grid(f); }

cudaError_t km_launch_finalize(const KmFinalizeArgs &f, cudaStream_t s, int64_t *launches) {
  int64_t kd = (int64_t)f.k * f.d;
  if (kd > 1024) {   // more than one block's worth: spread the divisions (128-bit fixed point -> binary64, a double-double
                     // inertia term each: ~300 instructions per element) over many SMs — one block took 27 us at k*d = 8192
    int grid = (int)((kd + 255) / 256);
    if (grid > 148) grid = 148;
    km_finalize_grid_kernel<<<grid, 256, 0, s>>>(f);
  } else {
    int threads = kd >= 1024 ? 1024 : (int)(((kd + 31) / 32) * 32);
    if (threads < 32) threads = 32;
    km_finalize_kernel<<<1, threads, 0, s>>>(f);
  }
  (*launches)++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ a2 / f1: k-means++ passes
// One pass per new centre: D_i = min(D_i, dist2(x_i, c)) in the oracle's binary64 operation
// order, q_i = trunc(D_i * 2^qd) and the exact integer sum of q over each tile of KM_PP_TILE
// consecutive points.  Integer sums are what make the draw independent of blocking and of the
// number of GPUs (oracle: ko_kmeanspp_indices).
template <typename T>
__global__ void __launch_bounds__(256) km_pp_update_kernel(const T *__restrict__ x, int64_t n, int d,
                                                           const double *__restrict__ c, double *__restrict__ D,
                                                           double scale, int first, long long *__restrict__ partial) {
  extern __shared__ double pp_smem[];
  const double *pp_c = c;   // very wide rows: read the centre through the cache instead
  if (d <= 4096) {
    for (int t = threadIdx.x; t < d; t += blockDim.x) pp_smem[t] = c[t];
    __syncthreads();
    pp_c = pp_smem;
  }
  const int64_t start = (int64_t)blockIdx.x * KM_PP_TILE;
  const int64_t end = min(n, start + KM_PP_TILE);
  long long sum = 0;
  for (int64_t i = start + threadIdx.x; i < end; i += blockDim.x) {
    const double di = km_dist2_f64(x + i * d, pp_c, d);
    double cur = first ? __longlong_as_double(0x7FF0000000000000LL) : D[i];
    if (di < cur) { cur = di; D[i] = cur; }
    sum += __double2ll_rz(cur * scale);
  }
  __shared__ long long red[8];
  sum = km_warp_sum_ll(sum);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long tot = 0;
    for (int w = 0; w < 8; w++) tot += red[w];
    partial[blockIdx.x] = tot;
  }
}

// least i in [start, start + count) whose running sum of q exceeds `target` (one thread: a tile)
__global__ void km_pp_find_kernel(const double *__restrict__ D, int64_t start, int count, double scale,
                                  long long target, long long *out_idx) {
  long long run = 0, pick = start + count - 1;
  for (int i = 0; i < count; i++) {
    run += __double2ll_rz(D[start + i] * scale);
    if (run > target) { pick = start + i; break; }
  }
  *out_idx = pick;
}

cudaError_t km_launch_pp_update(const void *x, int dtype, int64_t n, int32_t d, const double *c_dev, double *D,
                                double scale, int first, long long *partial, cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  const int64_t tiles = (n + KM_PP_TILE - 1) / KM_PP_TILE;
  const size_t smem = d <= 4096 ? (size_t)d * sizeof(double) : 0;
  if (dtype == KM_DTYPE_F32)
    km_pp_update_kernel<float><<<(unsigned)tiles, 256, smem, s>>>((const float *)x, n, d, c_dev, D, scale, first, partial);
  else
    km_pp_update_kernel<double><<<(unsigned)tiles, 256, smem, s>>>((const double *)x, n, d, c_dev, D, scale, first, partial);
  (*launches)++;
  return cudaGetLastError();
}

cudaError_t km_launch_pp_find(const double *D, int64_t start, int count, double scale, long long target,
                              long long *out_idx, cudaStream_t s, int64_t *launches) {
  km_pp_find_kernel<<<1, 1, 0, s>>>(D, start, count, scale, target, out_idx);
  (*launches)++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ a6 / f2: per-cluster inertia
// sum over the members of cluster j of dist2(x_i, c_j), binary64 per point in the oracle's
// operation order, summed as order-free fixed-point limbs (table: [k] coarse, [k] fine).
template <typename T>
__global__ void __launch_bounds__(256) km_cluster_inertia_kernel(const T *__restrict__ x, const int32_t *__restrict__ labels,
                                                                 int64_t n, int d, int k, const double *__restrict__ c,
                                                                 double scale_i, double scale_b, long long *__restrict__ table,
                                                                 int use_smem) {
  extern __shared__ long long ci_tab[];
  if (use_smem) {
    for (int e = threadIdx.x; e < 2 * k; e += blockDim.x) ci_tab[e] = 0;
    __syncthreads();
  }
  long long *tab = use_smem ? ci_tab : table;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int lab = labels[i];
    if (lab < 0) continue;
    long long a, b;
    km_fixed_split(km_dist2_f64(x + i * d, c + (size_t)lab * d, d), scale_i, scale_b, a, b);
    km_atomic_add_ll(tab + lab, a);
    km_atomic_add_ll(tab + k + lab, b);
  }
  if (use_smem) {
    __syncthreads();
    for (int e = threadIdx.x; e < 2 * k; e += blockDim.x)
      if (ci_tab[e] != 0) km_atomic_add_ll(table + e, ci_tab[e]);
  }
}

cudaError_t km_launch_cluster_inertia(const void *x, int dtype, const int32_t *labels, int64_t n, int32_t d, int32_t k,
                                      const double *c_dev, double scale_i, double scale_b, long long *table /* 2k, zeroed here */,
                                      int sm_count, cudaStream_t s, int64_t *launches) {
  cudaError_t e = cudaMemsetAsync(table, 0, (size_t)2 * k * sizeof(long long), s);
  if (e != cudaSuccess || n == 0) return e;
  const int use_smem = (size_t)2 * k * sizeof(long long) <= 32 * 1024;
  const size_t smem = use_smem ? (size_t)2 * k * sizeof(long long) : 0;
  int64_t want = (n + 256 * 16 - 1) / (256 * 16);
  int grid = (int)(want < 1 ? 1 : (want < (int64_t)sm_count * 8 ? want : (int64_t)sm_count * 8));
  if (dtype == KM_DTYPE_F32)
    km_cluster_inertia_kernel<float><<<grid, 256, smem, s>>>((const float *)x, labels, n, d, k, c_dev, scale_i, scale_b, table, use_smem);
  else
    km_cluster_inertia_kernel<double><<<grid, 256, smem, s>>>((const double *)x, labels, n, d, k, c_dev, scale_i, scale_b, table, use_smem);
  (*launches)++;
  return cudaGetLastError();
}
