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
  int qg, QBg;                // grid of the inertia terms (host: from max(max|x|, max|c|))
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
                                                       long long changed, long long &ga, long long &gb) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool upd = f.do_update && (!f.update_needs_change || changed > 0);
  const int QB = f.st->QB, shift = f.st->qa + f.st->QB;
  const int64_t kd = (int64_t)f.k * f.d;
  const double scale_g = scalbn(1.0, f.qg), scale_gb = scalbn(1.0, f.QBg);
  double lmax = 0.0;
  ga = 0; gb = 0;
  for (int64_t idx = first; idx < kd; idx += stride) {
    double c = f.c64[idx];
    if (upd || f.do_stats) {
      const long long cnt = __ldcg(f.acc_in + L.cnt + idx / f.d);
      if (cnt > 0) {
        const long long sa = __ldcg(f.acc_in + L.sa + idx), sb = __ldcg(f.acc_in + L.sb + idx);
        if (f.do_stats) {   // the pass just made used c as it stands here (before any update)
          long long ta, tb;
          km_inertia_term(c, sa, sb, cnt, QB, shift, scale_g, scale_gb, ta, tb);
          ga += ta; gb += tb;
        }
        if (upd) {
          c = __ddiv_rn(km_fixed_to_double(sa, sb, QB, shift), (double)cnt);
          f.c64[idx] = c;
        }
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
__device__ __forceinline__ void km_finalize_publish(const KmFinalizeArgs &f, double cmax, long long changed, long long ga,
                                                    long long gb) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool upd = f.do_update && (!f.update_needs_change || changed > 0);
  if (f.do_stats) {
    f.st->it += 1;
    KmStatSlot s;
    s.it = f.st->it;
    s.changed = changed;
    s.inertia = km_inertia_from_limbs(f.st->Ta, f.st->Tb, f.st->QBt, f.st->qt, ga, gb, f.QBg, f.qg);
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
  long long ga, gb;
  double lmax = km_finalize_elements(f, threadIdx.x, blockDim.x, changed, ga, gb);
  __shared__ double red[32];
  __shared__ long long redg[2][32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  ga = km_warp_sum_ll(ga);
  gb = km_warp_sum_ll(gb);
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = lmax; redg[0][threadIdx.x >> 5] = ga; redg[1][threadIdx.x >> 5] = gb; }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int nwarps = (blockDim.x + 31) >> 5;
  double cmax = 0.0;
  ga = 0; gb = 0;
  for (int w = 0; w < nwarps; w++) { cmax = fmax(cmax, red[w]); ga += redg[0][w]; gb += redg[1][w]; }
  km_finalize_publish(f, cmax, changed, ga, gb);
}

// Many blocks (large k*d): block maxima meet in st->cmax_bits, the last block to finish publishes.
__device__ __forceinline__ void km_finalize_grid(const KmFinalizeArgs &f) {
  const KmAccLayout L = km_acc_layout(f.k, f.d);
  const bool reads = f.do_stats || f.do_update;
  const long long changed = reads ? __ldcg(f.acc_in + L.scal + KM_ACC_CHANGED) : 0;
  long long ga, gb;
  double lmax = km_finalize_elements(f, (int64_t)blockIdx.x * blockDim.x + threadIdx.x, (int64_t)gridDim.x * blockDim.x, changed, ga, gb);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  ga = km_warp_sum_ll(ga);
  gb = km_warp_sum_ll(gb);
  if ((threadIdx.x & 31) == 0) {
    if (lmax > 0.0) atomicMax(&f.st->cmax_bits, (unsigned long long)__double_as_longlong(lmax));
    if (ga) atomicAdd(&f.st->g_a, (unsigned long long)ga);   // integer adds: any order, same bits
    if (gb) atomicAdd(&f.st->g_b, (unsigned long long)gb);
  }
  __shared__ int is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(&f.st->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last || threadIdx.x != 0) return;
  __threadfence();
  const unsigned long long bits = atomicExch(&f.st->cmax_bits, 0ULL);
  const long long tga = (long long)atomicExch(&f.st->g_a, 0ULL), tgb = (long long)atomicExch(&f.st->g_b, 0ULL);
  f.st->ticket = 0;
  km_finalize_publish(f, __longlong_as_double((long long)bits), changed, tga, tgb);
}
#endif
