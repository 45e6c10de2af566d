// km_assign_direct.cu — KM_PATH_DIRECT_F64: every distance in binary64 with the oracle's
// exact operation order.  Universal (any d, k, float32 or float64
// points) and slow; it is the on-device statement of the spec that the filtered paths
// must agree with bit for bit, and the path float64-stored points take.
#include "km_commit.cuh"
#include "km_internal.h"

#define KM_DIRECT_THREADS 128

template <typename XT>
__global__ void __launch_bounds__(KM_DIRECT_THREADS)
km_assign_direct_kernel(const XT *__restrict__ x, int64_t n, int32_t d, int32_t k, int32_t *__restrict__ labels,
                        const double *__restrict__ c64, long long *__restrict__ acc, const KmState *__restrict__ st,
                        int incremental, int chunk_rows, int tab_in_smem) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *cs = reinterpret_cast<double *>(smem_raw);                                 // [chunk_rows][d]
  long long *stab = reinterpret_cast<long long *>(cs + (size_t)chunk_rows * d);       // optional table
  const KmAccLayout L = km_acc_layout(k, d);
  KmBlockTab T;
  T.L = L;
  T.tab = tab_in_smem ? stab : acc;
  if (tab_in_smem) km_tab_zero(stab, L.scal);
  const double scale_a = st->scale_a, scale_b = st->scale_b;
  KmThreadScalars S;
  const int64_t ntiles = (n + KM_DIRECT_THREADS - 1) / KM_DIRECT_THREADS;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t i = tile * KM_DIRECT_THREADS + threadIdx.x;
    const bool live = i < n;
    const XT *xrow = x + (live ? i : 0) * (int64_t)d;
    double best = __longlong_as_double(0x7FF0000000000000LL);  // +Infinity
    int lab = 0;
    for (int j0 = 0; j0 < k; j0 += chunk_rows) {
      const int rows = min(chunk_rows, k - j0);
      __syncthreads();
      for (int e = threadIdx.x; e < rows * d; e += KM_DIRECT_THREADS) cs[e] = c64[(int64_t)j0 * d + e];
      __syncthreads();
      if (live) {
        for (int j = 0; j < rows; j++) {
          double dj = km_dist2_f64(xrow, cs + (size_t)j * d, d);
          if (dj < best) { best = dj; lab = j0 + j; }  // strict <: lowest index wins, NaN never wins
        }
      }
    }
    if (live) {
      const int old = labels[i];
      if (lab != old) labels[i] = lab;
      km_commit_point_sums(T, S, xrow, d, old, lab, incremental, scale_a, scale_b);
    }
  }
  __syncthreads();
  if (tab_in_smem) km_tab_flush(stab, acc, L.scal);
  km_flush_scalars(S, acc, L);
}

cudaError_t km_launch_assign_direct(const KmAssignArgs &a, int64_t *launches) {
  if (a.n == 0) return cudaSuccess;
  const KmAccLayout L = km_acc_layout(a.k, a.d);
  // centroid chunk: as many rows as fit in ~32 KB, at least one
  int chunk_rows = (int)(32768 / ((size_t)a.d * sizeof(double)));
  if (chunk_rows < 1) chunk_rows = 1;
  if (chunk_rows > a.k) chunk_rows = a.k;
  size_t smem = (size_t)chunk_rows * a.d * sizeof(double);
  int tab_in_smem = 0;
  if ((size_t)L.scal * sizeof(long long) <= 48 * 1024 && smem + (size_t)L.scal * 8 <= (size_t)a.max_smem_optin - 2048) {
    tab_in_smem = 1;
    smem += (size_t)L.scal * sizeof(long long);
  }
  if (smem > (size_t)a.max_smem_optin) return cudaErrorInvalidValue;  // d too large even for one row
  const int64_t ntiles = (a.n + KM_DIRECT_THREADS - 1) / KM_DIRECT_THREADS;
  const int64_t cap = (int64_t)a.sm_count * 8;
  const int grid = (int)(ntiles < cap ? ntiles : cap);
  cudaError_t e;
  if (a.dtype == KM_DTYPE_F32) {
    e = cudaFuncSetAttribute(km_assign_direct_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    km_assign_direct_kernel<float><<<grid, KM_DIRECT_THREADS, smem, a.stream>>>(
        (const float *)a.x, a.n, a.d, a.k, a.labels, a.c64, a.acc, a.state, a.incremental, chunk_rows, tab_in_smem);
  } else {
    e = cudaFuncSetAttribute(km_assign_direct_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    km_assign_direct_kernel<double><<<grid, KM_DIRECT_THREADS, smem, a.stream>>>(
        (const double *)a.x, a.n, a.d, a.k, a.labels, a.c64, a.acc, a.state, a.incremental, chunk_rows, tab_in_smem);
  }
  (*launches)++;
  return cudaGetLastError();
}
