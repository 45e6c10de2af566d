// km_sort.cu — one-time spatial reordering of d = 2 points (a1: store layout).
//
// The pruned small-d kernel looks up, per point, the candidate list of the grid cell the
// point falls in.  If consecutive points fall in unrelated cells (the order they were
// uploaded in), the 32 lanes of a warp walk 32 different lists and the warp runs to its
// slowest lane.  Points never move during a run, so the store sorts them by cell ONCE (stable
// radix sort keyed by cell — deterministic), keeps the permutation, and
// un-permutes labels / rows whenever the caller asks for them.  Every per-cluster sum is
// order-free, so the results do not depend on this order.
#include <cub/device/device_radix_sort.cuh>

#include "km_internal.h"

// What travels through the radix sort with each cell key: the point itself, its original row and
// its label — 16 bytes, so the passes move every record with one coalesced 16-byte access and no
// random gather is needed afterwards (a gather of 8-byte points by a random permutation fetches
// a 32-byte sector per point: it cost more than the sort itself).
struct __align__(16) KmSortRec {
  float x, y;
  unsigned row;
  int label;
};

__global__ void __launch_bounds__(256) km_sort_pack_kernel(const float2 *__restrict__ x, const int32_t *__restrict__ labels,
                                                           int64_t n, KmPrune2 pr, unsigned *__restrict__ key,
                                                           KmSortRec *__restrict__ rec) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float2 v = x[i];
    const int ix = min(__float_as_int(fmaf(v.x, pr.sx, pr.kxm)) & 0x3FFFFF, pr.gp - 1);
    const int iy = min(__float_as_int(fmaf(v.y, pr.sy, pr.kym)) & 0x3FFFFF, pr.gp - 1);
    key[i] = (unsigned)(iy * pr.gp + ix);
    KmSortRec r;
    r.x = v.x; r.y = v.y; r.row = (unsigned)i; r.label = labels[i];
    rec[i] = r;
  }
}

__global__ void __launch_bounds__(256) km_sort_unpack_kernel(const KmSortRec *__restrict__ rec, int64_t n,
                                                             float2 *__restrict__ xs, int32_t *__restrict__ ls,
                                                             unsigned *__restrict__ perm) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const KmSortRec r = rec[p];
    xs[p] = make_float2(r.x, r.y);
    ls[p] = r.label;
    perm[p] = r.row;
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
// hold the sorted rows (same buffers, rewritten) and *perm_out maps sorted position -> original row.
cudaError_t km_sort_points_by_cell(float **x, int32_t **labels, int64_t n, const KmPrune2 &pr, unsigned **perm_out,
                                   int sm_count, cudaStream_t s, int64_t *launches) {
  *perm_out = nullptr;
  if (n < 2 || n >= ((int64_t)1 << 31)) return cudaSuccess;
  // one allocation for everything temporary: keys in/out, records in/out, CUB's scratch
  int bits = 1;
  while ((1 << bits) < pr.gp * pr.gp) bits++;
  size_t temp_bytes = 0;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, (const unsigned *)nullptr, (unsigned *)nullptr,
                                                  (const KmSortRec *)nullptr, (KmSortRec *)nullptr, (int)n, 0, bits, s);
  if (e != cudaSuccess) return e;
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const size_t kb = up((size_t)n * sizeof(unsigned)), rb = up((size_t)n * sizeof(KmSortRec));
  char *pool = nullptr;
  unsigned *perm = nullptr;
  if ((e = cudaMalloc(&pool, 2 * kb + 2 * rb + up(temp_bytes))) != cudaSuccess) return e;
  if ((e = cudaMalloc(&perm, (size_t)n * sizeof(unsigned))) != cudaSuccess) { cudaFree(pool); return e; }
  unsigned *key_in = reinterpret_cast<unsigned *>(pool), *key_out = reinterpret_cast<unsigned *>(pool + kb);
  KmSortRec *rec_in = reinterpret_cast<KmSortRec *>(pool + 2 * kb), *rec_out = reinterpret_cast<KmSortRec *>(pool + 2 * kb + rb);
  void *temp = pool + 2 * kb + 2 * rb;
  km_sort_pack_kernel<<<grid_for(n, sm_count), 256, 0, s>>>(reinterpret_cast<const float2 *>(*x), *labels, n, pr, key_in, rec_in);
  (*launches)++;
  e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, key_in, key_out, rec_in, rec_out, (int)n, 0, bits, s);   // stable
  (*launches) += 4;
  if (e == cudaSuccess) {
    km_sort_unpack_kernel<<<grid_for(n, sm_count), 256, 0, s>>>(rec_out, n, reinterpret_cast<float2 *>(*x), *labels, perm);
    (*launches)++;
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cudaFree(pool);
  if (e != cudaSuccess) { cudaFree(perm); return e; }
  *perm_out = perm;
  return cudaSuccess;
}

cudaError_t km_unpermute_labels(const int32_t *labels_sorted, const unsigned *perm, int64_t n, int32_t *out_dev,
                                int sm_count, cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  km_unpermute_kernel<int32_t><<<grid_for(n, sm_count), 256, 0, s>>>(labels_sorted, perm, n, 0, n, out_dev);
  (*launches)++;
  return cudaGetLastError();
}

// labels as bytes (k <= 255; 255 = never assigned), in the caller's row order: a quarter of the bytes to bring home
__global__ void __launch_bounds__(256) km_narrow_labels_kernel(const int32_t *__restrict__ labels, const unsigned *__restrict__ perm,
                                                               int64_t n, unsigned char *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (perm) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      const int l = labels[i];
      out[perm[i]] = (unsigned char)(l < 0 ? 255 : l);
    }
  } else {
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q * 4 < n; q += stride) {
      unsigned w = 0;
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const int64_t i = q * 4 + t;
        const int l = i < n ? labels[i] : 0;
        w |= (unsigned)(l < 0 ? 255 : (l & 255)) << (8 * t);
      }
      if (q * 4 + 4 <= n) reinterpret_cast<unsigned *>(out)[q] = w;
      else for (int t = 0; q * 4 + t < n; t++) out[q * 4 + t] = (unsigned char)(w >> (8 * t));
    }
  }
}

cudaError_t km_narrow_labels(const int32_t *labels, const unsigned *perm, int64_t n, unsigned char *out_dev, int sm_count,
                             cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  km_narrow_labels_kernel<<<grid_for(perm ? n : (n + 3) / 4, sm_count), 256, 0, s>>>(labels, perm, n, out_dev);
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
