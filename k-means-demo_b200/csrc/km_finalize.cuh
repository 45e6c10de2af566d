// km_finalize.cuh — a4, the centroid-update step as ONE block's work:
//   c_j = sum_j / count_j from the order-free sums; an empty cluster keeps its centroid
//   (A4; the analog of a member-less centroid simply persisting in the reference,
//   /root/reference/app.mjs:322,484-488).  Also refreshes the binary32 working copies and the
//   filter / inertia scales that depend on the centroids, publishes the iteration's stats
//   to the pinned ring and clears the per-iteration scalars.
// Runs either as its own single-block kernel or inside the last block of an assignment
// kernel (one launch per Lloyd iteration).
#pragma once
#include "km_common.cuh"

struct KmFinalizeArgs {
  int fused;                  // 1: the assignment kernel's last block runs km_finalize_block
  const long long *acc_in;    // table to read (all-reduced copy when world > 1)
  long long *acc_local;       // this rank's table (per-iteration scalars are cleared here)
  double *c64;
  float *chi;
  float *clo;
  int32_t k, d;
  KmState *st;
  KmStatSlot *slot;           // pinned host slot (device alias) or nullptr
  double maxabs_x;
  int do_stats, do_update, update_needs_change;
};

#ifdef __CUDACC__
__device__ __forceinline__ int km_bitlen(int v) { return 32 - __clz(v); }

// Filter error terms and the inertia scale follow the current centroids.
struct KmScales {
  int qi;
  double inertia_scale;
  float tau_hilo, tau_hi;
};
__device__ __forceinline__ KmScales km_scales_from_cmax(double cmax, double maxabs_x, int d, int QB) {
  KmScales r;
  double M = fmax(cmax, maxabs_x);
  int EM = (M > 0.0) ? (ilogb(M) + 1) : 0;
  int Ei = 2 * EM + 2 + km_bitlen(d);
  r.qi = QB - Ei;
  r.inertia_scale = scalbn(1.0, r.qi);
  double c2 = cmax * cmax;
  double t_hilo = 2.5 * (4.2 * (double)d * 0x1.0p-72 * c2) + (double)d * 0x1.0p-140;
  double t_hi = 2.5 * (1.01 * (double)d * 0x1.0p-24 * c2) + (double)d * 0x1.0p-140;
  r.tau_hilo = __double2float_ru(t_hilo);
  r.tau_hi = __double2float_ru(t_hi);
  return r;
}
__device__ __forceinline__ void km_refresh_scales(KmState *st, double cmax, double maxabs_x, int d) {
  const KmScales r = km_scales_from_cmax(cmax, maxabs_x, d, st->QB);
  st->qi = r.qi;
  st->inertia_scale = r.inertia_scale;
  st->cmax = cmax;
  st->tau_hilo = r.tau_hilo;
  st->tau_hi = r.tau_hi;
}

// Part 1: the elements [first, kd) in steps of `stride`: c = S / count (or unchanged), the
// binary32 hi/lo copies; returns this thread's max |c|.
__device__ __forceinline__ double km_finalize_elements(const KmFinalizeArgs &f, int64_t first, int64_t stride,
                                                       long long changed) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool upd = f.do_update && (!f.update_needs_change || changed > 0);
  const int QB = f.st->QB, shift = f.st->qa + f.st->QB;
  const int64_t kd = (int64_t)f.k * f.d;
  double lmax = 0.0;
  for (int64_t idx = first; idx < kd; idx += stride) {
    double c = f.c64[idx];
    if (upd) {
      const long long cnt = __ldcg(f.acc_in + L.cnt + idx / f.d);
      if (cnt > 0) {
        const double s = km_fixed_to_double(__ldcg(f.acc_in + L.sa + idx), __ldcg(f.acc_in + L.sb + idx), QB, shift);
        c = __ddiv_rn(s, (double)cnt);
        f.c64[idx] = c;
      }
    }
    const float h = __double2float_rn(c);
    f.chi[idx] = h;
    f.clo[idx] = __double2float_rn(__dsub_rn(c, (double)h));
    lmax = fmax(lmax, fabs(c));
  }
  return lmax;
}

// Part 2 (one thread): publish the pass's stats, clear the per-pass scalars, refresh the scales.
__device__ __forceinline__ void km_finalize_publish(const KmFinalizeArgs &f, double cmax, long long changed) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool upd = f.do_update && (!f.update_needs_change || changed > 0);
  const int QB = f.st->QB;
  const int qi_used = f.st->qi;  // the scale the assignment just used
  if (f.do_stats) {
    f.st->it += 1;
    KmStatSlot s;
    s.it = f.st->it;
    s.changed = changed;
    s.inertia = km_fixed_to_double(__ldcg(f.acc_in + L.scal + KM_ACC_INERTIA_A),
                                   __ldcg(f.acc_in + L.scal + KM_ACC_INERTIA_B), QB, qi_used + QB);
    s.rescored = __ldcg(f.acc_in + L.scal + KM_ACC_RESCORED);
    s.updated = upd ? 1 : 0;
    s.pad = 0;
    for (int q = 0; q < KM_ACC_SCALARS; q++) f.acc_local[L.scal + q] = 0;
    if (f.slot) *f.slot = s;  // pinned host memory; visible to the host once the kernel's event completes
  }
  km_refresh_scales(f.st, cmax, f.maxabs_x, f.d);
}

// All threads of exactly one block call this.
__device__ __forceinline__ void km_finalize_block(const KmFinalizeArgs &f) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool reads = f.do_stats || f.do_update;
  const long long changed = reads ? __ldcg(f.acc_in + L.scal + KM_ACC_CHANGED) : 0;
  double lmax = km_finalize_elements(f, threadIdx.x, blockDim.x, changed);
  __shared__ double red[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = lmax;
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int nwarps = (blockDim.x + 31) >> 5;
  double cmax = 0.0;
  for (int w = 0; w < nwarps; w++) cmax = fmax(cmax, red[w]);
  km_finalize_publish(f, cmax, changed);
}

// Many blocks (large k*d): block maxima meet in st->cmax_bits, the last block to finish publishes.
__device__ __forceinline__ void km_finalize_grid(const KmFinalizeArgs &f) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool reads = f.do_stats || f.do_update;
  const long long changed = reads ? __ldcg(f.acc_in + L.scal + KM_ACC_CHANGED) : 0;
  double lmax = km_finalize_elements(f, (int64_t)blockIdx.x * blockDim.x + threadIdx.x, (int64_t)gridDim.x * blockDim.x, changed);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  if ((threadIdx.x & 31) == 0 && lmax > 0.0) atomicMax(&f.st->cmax_bits, (unsigned long long)__double_as_longlong(lmax));
  __shared__ int is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(&f.st->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  __threadfence();
  const unsigned long long bits = atomicExch(&f.st->cmax_bits, 0ULL);
  f.st->ticket = 0;
  km_finalize_publish(f, __longlong_as_double((long long)bits), changed);
}
#endif
