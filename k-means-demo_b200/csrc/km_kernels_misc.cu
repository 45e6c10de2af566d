// km_kernels_misc.cu — store scan, blob generator, label fill, and the finalize kernel
// (order-free sums -> centroids, empty-cluster rule, per-iteration stats).
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
// Filter error terms and the inertia scale follow the current centroids; DESIGN.md §5
// derives the constants.
__device__ __forceinline__ int km_bitlen(int v) { return 32 - __clz(v); }

__device__ void km_refresh_scales(KmState *st, double cmax, double maxabs_x, int d) {
  double M = fmax(cmax, maxabs_x);
  int EM = (M > 0.0) ? (ilogb(M) + 1) : 0;
  int Ei = 2 * EM + 2 + km_bitlen(d);
  int qi = st->QB - Ei;
  st->qi = qi;
  st->inertia_scale = scalbn(1.0, qi);
  st->cmax = cmax;
  double c2 = cmax * cmax;
  double t_hilo = 2.5 * (4.2 * (double)d * 0x1.0p-72 * c2) + (double)d * 0x1.0p-140;
  double t_hi = 2.5 * (1.01 * (double)d * 0x1.0p-24 * c2) + (double)d * 0x1.0p-140;
  st->tau_hilo = __double2float_ru(t_hilo);
  st->tau_hi = __double2float_ru(t_hi);
}

// c_j = sum_j / count_j; an empty cluster keeps its centroid (A4; analog of a member-less
// centroid persisting, /root/reference/app.mjs:322,484-488).  Also refreshes the binary32
// working copies, and — in the last block to finish — publishes the iteration's stats to
// the pinned ring and clears the per-iteration scalars.
__global__ void __launch_bounds__(256)
km_finalize_kernel(const long long *__restrict__ acc_in, long long *__restrict__ acc_local, double *__restrict__ c64,
                   float *__restrict__ chi, float *__restrict__ clo, int32_t k, int32_t d, KmState *st,
                   KmStatSlot *slot, double maxabs_x, int do_stats, int do_update, int update_needs_change) {
  const KmAccLayout L = km_acc_layout(k, d);
  const long long changed = (do_stats || do_update) ? acc_in[L.scal + KM_ACC_CHANGED] : 0;
  const bool upd = do_update && (!update_needs_change || changed > 0);
  const int QB = st->QB, shift = st->qa + st->QB;
  const int64_t kd = (int64_t)k * d;
  double lmax = 0.0;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < kd;
       idx += (int64_t)gridDim.x * blockDim.x) {
    double c = c64[idx];
    if (upd) {
      long long cnt = acc_in[L.cnt + idx / d];
      if (cnt > 0) {
        double s = km_fixed_to_double(acc_in[L.sa + idx], acc_in[L.sb + idx], QB, shift);
        c = __ddiv_rn(s, (double)cnt);
        c64[idx] = c;
      }
    }
    float h = __double2float_rn(c);
    chi[idx] = h;
    clo[idx] = __double2float_rn(__dsub_rn(c, (double)h));
    lmax = fmax(lmax, fabs(c));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  if ((threadIdx.x & 31) == 0 && lmax > 0.0)
    atomicMax(&st->cmax_bits, (unsigned long long)__double_as_longlong(lmax));
  __threadfence();
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  __threadfence();
  unsigned long long bits = atomicExch(&st->cmax_bits, 0ULL);
  st->ticket = 0;
  const int qi_used = st->qi;  // the scale the assign kernel just used
  if (do_stats) {
    st->it += 1;
    slot->it = st->it;
    slot->changed = changed;
    slot->inertia = km_fixed_to_double(acc_in[L.scal + KM_ACC_INERTIA_A], acc_in[L.scal + KM_ACC_INERTIA_B], QB,
                                       qi_used + QB);
    slot->rescored = acc_in[L.scal + KM_ACC_RESCORED];
    slot->updated = upd ? 1 : 0;
    for (int q = 0; q < KM_ACC_SCALARS; q++) acc_local[L.scal + q] = 0;
    __threadfence_system();
  }
  km_refresh_scales(st, __longlong_as_double((long long)bits), maxabs_x, d);
}

cudaError_t km_launch_finalize(const long long *acc_in, long long *acc_local, double *c64, float *chi, float *clo,
                               int32_t k, int32_t d, KmState *state, KmStatSlot *slot, double maxabs_x,
                               int do_stats, int do_update, int update_needs_change, int /*zero_sums*/,
                               cudaStream_t s, int64_t *launches) {
  int64_t kd = (int64_t)k * d;
  int64_t want = (kd + 255) / 256;
  int grid = (int)(want < 1 ? 1 : (want > 128 ? 128 : want));
  km_finalize_kernel<<<grid, 256, 0, s>>>(acc_in, acc_local, c64, chi, clo, k, d, state, slot, maxabs_x, do_stats,
                                          do_update, update_needs_change);
  (*launches)++;
  return cudaGetLastError();
}

cudaError_t km_launch_prepare_centroids(const double *c64, float *chi, float *clo, int64_t kd, KmState *state,
                                        double maxabs_x, int32_t d, cudaStream_t s, int64_t *launches) {
  // finalize with nothing to consume: rewrites chi/clo from c64 and refreshes the scales
  int32_t k = (int32_t)(kd / d);
  return km_launch_finalize(nullptr, nullptr, const_cast<double *>(c64), chi, clo, k, d, state, nullptr, maxabs_x, 0,
                            0, 0, 0, s, launches);
}
