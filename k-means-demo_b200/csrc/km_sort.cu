// km_sort.cu — one-time spatial reordering of d = 2 points (a1: store layout).
//
// The pruned small-d kernel looks up, per point, the candidate list of the grid cell the
// point falls in.  If consecutive points fall in unrelated cells (the order they were
// uploaded in), the 32 lanes of a warp walk 32 different lists and the warp runs to its
// slowest lane.  Points never move during a run, so the store sorts them by cell ONCE (stable
// radix sort of (cell, original index) — deterministic), keeps the permutation, and
// un-permutes labels / rows whenever the caller asks for them.  Every per-cluster sum is
// order-free, so the results do not depend on this order.
#include <cub/device/device_radix_sort.cuh>

#include "km_internal.h"

__global__ void __launch_bounds__(256) km_cell_key_kernel(const float2 *__restrict__ x, int64_t n, KmPrune2 pr,
                                                          unsigned *__restrict__ key, unsigned *__restrict__ idx) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float2 v = x[i];
    const int ix = min(__float_as_int(fmaf(v.x, pr.sx, pr.kxm)) & 0x3FFFFF, pr.gp - 1);
    const int iy = min(__float_as_int(fmaf(v.y, pr.sy, pr.kym)) & 0x3FFFFF, pr.gp - 1);
    key[i] = (unsigned)(iy * pr.gp + ix);
    idx[i] = (unsigned)i;
  }
}

__global__ void __launch_bounds__(256) km_gather_kernel(const float2 *__restrict__ x, const int32_t *__restrict__ labels,
                                                        const unsigned *__restrict__ perm, int64_t n,
                                                        float2 *__restrict__ xs, int32_t *__restrict__ ls) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const unsigned j = perm[p];
    xs[p] = x[j];
    ls[p] = labels[j];
  }
}

// out[perm[p] - i0] = src[p] for the original rows [i0, i0 + cnt)
template <typename T>
__global__ void __launch_bounds__(256) km_unpermute_kernel(const T *__restrict__ src, const unsigned *__restrict__ perm,
                                                           int64_t n, int64_t i0, int64_t cnt, T *__restrict__ out) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = (int64_t)perm[p] - i0;
    if (j >= 0 && j < cnt) out[j] = src[p];
  }
}

static int grid_for(int64_t n, int sm_count) {
  int64_t want = (n + 256 * 4 - 1) / (256 * 4);
  int64_t cap = (int64_t)sm_count * 16;
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

// Sort the ctx-owned d = 2 points (and their labels) by grid cell.  On success *x and *labels
// point to new buffers (the old ones are freed) and *perm_out maps sorted position -> original row.
cudaError_t km_sort_points_by_cell(float **x, int32_t **labels, int64_t n, const KmPrune2 &pr, unsigned **perm_out,
                                   int sm_count, cudaStream_t s, int64_t *launches) {
  *perm_out = nullptr;
  if (n < 2 || n >= ((int64_t)1 << 31)) return cudaSuccess;
  unsigned *key_in = nullptr, *key_out = nullptr, *idx_in = nullptr, *idx_out = nullptr;
  void *temp = nullptr;
  float2 *xs = nullptr;
  int32_t *ls = nullptr;
  size_t temp_bytes = 0;
  cudaError_t e = cudaSuccess;
  auto cleanup = [&]() {
    if (key_in) cudaFree(key_in);
    if (key_out) cudaFree(key_out);
    if (idx_in) cudaFree(idx_in);
    if (temp) cudaFree(temp);
  };
  int bits = 1;
  while ((1 << bits) < pr.gp * pr.gp) bits++;
  if ((e = cudaMalloc(&key_in, n * sizeof(unsigned))) != cudaSuccess) goto fail;
  if ((e = cudaMalloc(&key_out, n * sizeof(unsigned))) != cudaSuccess) goto fail;
  if ((e = cudaMalloc(&idx_in, n * sizeof(unsigned))) != cudaSuccess) goto fail;
  if ((e = cudaMalloc(&idx_out, n * sizeof(unsigned))) != cudaSuccess) goto fail;
  km_cell_key_kernel<<<grid_for(n, sm_count), 256, 0, s>>>(reinterpret_cast<const float2 *>(*x), n, pr, key_in, idx_in);
  (*launches)++;
  if ((e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, key_in, key_out, idx_in, idx_out, (int)n, 0, bits, s)) != cudaSuccess) goto fail;
  if ((e = cudaMalloc(&temp, temp_bytes)) != cudaSuccess) goto fail;
  if ((e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, key_in, key_out, idx_in, idx_out, (int)n, 0, bits, s)) != cudaSuccess) goto fail;
  (*launches) += 4;
  if ((e = cudaMalloc(&xs, n * sizeof(float2))) != cudaSuccess) goto fail;
  if ((e = cudaMalloc(&ls, n * sizeof(int32_t))) != cudaSuccess) goto fail;
  km_gather_kernel<<<grid_for(n, sm_count), 256, 0, s>>>(reinterpret_cast<const float2 *>(*x), *labels, idx_out, n, xs, ls);
  (*launches)++;
  if ((e = cudaStreamSynchronize(s)) != cudaSuccess) goto fail;
  cudaFree(*x);
  cudaFree(*labels);
  *x = reinterpret_cast<float *>(xs);
  *labels = ls;
  *perm_out = idx_out;
  cleanup();
  return cudaSuccess;
fail:
  cleanup();
  if (idx_out) cudaFree(idx_out);
  if (xs) cudaFree(xs);
  if (ls) cudaFree(ls);
  return e;
}

cudaError_t km_unpermute_labels(const int32_t *labels_sorted, const unsigned *perm, int64_t n, int32_t *out_dev,
                                int sm_count, cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  km_unpermute_kernel<int32_t><<<grid_for(n, sm_count), 256, 0, s>>>(labels_sorted, perm, n, 0, n, out_dev);
  (*launches)++;
  return cudaGetLastError();
}

cudaError_t km_unpermute_rows2(const float *x_sorted, const unsigned *perm, int64_t n, int64_t i0, int64_t cnt,
                               float *out_dev, int sm_count, cudaStream_t s, int64_t *launches) {
  if (n == 0 || cnt == 0) return cudaSuccess;
  km_unpermute_kernel<float2><<<grid_for(n, sm_count), 256, 0, s>>>(reinterpret_cast<const float2 *>(x_sorted), perm, n, i0, cnt,
                                                                    reinterpret_cast<float2 *>(out_dev));
  (*launches)++;
  return cudaGetLastError();
}
