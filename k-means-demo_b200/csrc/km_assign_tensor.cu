// km_assign_tensor.cu — KM_PATH_TENSOR: GEMM-form assignment on the 5th-gen tensor cores
// (tcgen05.mma, accumulators in TMEM) for large k*d (BASELINE.json configs[2], configs[3]).
//
// One assignment = five small/large launches:
//
//  prep     |c_j|^2 and a zero-padded BF16 copy of the centroids in the row layout the TMA
//           descriptor reads ([kpad][kblocks*64]).
//  order    a counting sort of this rank's point indices by their PREVIOUS label.  Points of one
//           cluster then sit in the same warps of the coarse stage, so a warp's 32 points find
//           their few candidate centroids in the same columns (points never move; only an index
//           list is rebuilt, and every per-cluster sum is order-free).
//  xbf      once per upload: a BF16 copy of x (rows zero-padded to 64-column blocks, plus the 1.0 columns
//           of the folded h), so that the coarse stage can take its point rows by TMA.
//  coarse   km_tensor_coarse_kernel, warp-specialised, one CTA per SM, 256 points per pass:
//             4 gather warps    cp.async.bulk.tensor ... tile::gather4: the pass's rows of the BF16 copy, picked
//                               by the index list, land K-major / SWIZZLE_128B in two 128-row tiles (2-4
//                               stages, mbarrier complete_tx) — no thread loads, converts or stores a point;
//             1 TMA warp        an elected lane streams 128-centroid tiles of the BF16 copy with
//                               cp.async.bulk.tensor (2-4 stages, mbarrier complete_tx), starting
//                               with the tile that holds the pass's previous label;
//             1 MMA warp        an elected lane issues tcgen05.mma kind::f16, 2 point tiles x 128
//                               centroids per step into 4 x 128 TMEM columns (double-buffered),
//                               tcgen05.commit -> mbarriers;
//             8 epilogue warps  tcgen05.ld 32 columns at a time; score = x~.c~ - |c|^2/2 (packed
//                               f32x2 adds of h from shared memory, or nothing at all when h rides
//                               in the contraction), the chunk maximum by 3-input FMNMX; only a
//                               chunk whose maximum comes within the BF16 error window of the
//                               running maximum is looked at column by column.
//           The N x k matrix never exists.  A point whose only survivor is its previous label is
//           finished right there (nothing is written for it: the inertia comes from the sums);
//           every other point goes to a compact worklist with up to 8 survivors.
//  exact    km_tensor_exact_kernel over the worklist only: binary32 differences + the §5 test, the
//           oracle's binary64 arithmetic where the gap is inside the bound, label write, move
//           between the order-free sums.
//
// Labels are the oracle's bit for bit: a centroid outside a point's survivor set is provably not
// its binary64 argmin (bound in DESIGN.md §6.7; attacked by tests/test_gpu_at_size.py).
#include <cuda.h>   // CUtensorMap and its enums only; the encoder is fetched from the driver at run time
#include <cuda_bf16.h>

#include <algorithm>

#include "km_commit.cuh"
#include "km_internal.h"

#define KT_M 128            // points per MMA tile (UMMA M, cta_group::1)
#define KT_N 128            // centroids per MMA tile (UMMA N)
#define KT_ROWS 256         // points per CTA pass: two MMA tiles
#define KT_LIST 8           // per-point survivor list kept in shared memory while the tiles stream by
#define KT_EPI_WARPS 8      // warp w: point tile w >> 2, TMEM lane quadrant w & 3
#define KT_WARP_GATHER KT_EPI_WARPS          // TMA: rows of the BF16 copy of x, gathered by the index list ...
#define KT_GATHER_WARPS 4                    // ... by four warps, 16 lanes of each (a TMA instruction is issued lane by lane: more warps, more issue)
#define KT_WARP_MMA (KT_EPI_WARPS + KT_GATHER_WARPS)
#define KT_WARP_TMA (KT_WARP_MMA + 1)        // TMA: centroid tiles
#define KT_THREADS ((KT_WARP_TMA + 1) * 32)   // 448
#define KT_KB_BYTES (KT_M * 128)   // one 64-element (128-byte) k-block of a 128-row tile: 16 KiB
#define KT_WORK_INTS 12     // worklist entry: {row, n (or -2: scan all k), 0, 0, c0..c7}

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
// make generic-proxy writes to shared memory visible to the tensor core / TMA (async proxy)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// the same for waiters that are usually early (producers, the TMA thread): sleep between polls so the
// spin does not take issue slots from the epilogue warps of the same scheduler
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "WAIT_%=:\n\t"
      "nanosleep.u32 64;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// TMA: one box of the 2-D tensor map -> shared memory, completion counted in bytes on `bar`
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// TMA gather: four rows of the 2-D tensor map, picked by their row numbers, one box-width of columns each
// -> four consecutive 128-byte rows in shared memory (swizzled by the hardware like any other tile)
__device__ __forceinline__ void tma_gather4(uint32_t smem_dst, const CUtensorMap *map, int c0, int r0, int r1, int r2, int r3, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_dst),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t *slot_in_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, BF16 inputs, FP32 accumulate, K = 16; issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// 32 consecutive accumulator columns of this thread's TMEM lane; valid after the next tmem_ld_wait()
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
// 16-byte read-only load that does not allocate in L1 (data read exactly once)
__device__ __forceinline__ float4 ldg_stream4(const float *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// one lane of a converged warp (the compiler keeps single-thread tcgen05 / TMA issue cheap behind this predicate)
__device__ __forceinline__ uint32_t elect_one_sync() {
  uint32_t pred = 0, laneid = 0;
  asm volatile(
      "{\n\t.reg .b32 %%rx;\n\t.reg .pred %%px;\n\t"
      "elect.sync %%rx|%%px, %2;\n\t"
      "@%%px mov.s32 %1, 1;\n\t"
      "mov.s32 %0, %%rx;\n\t}"
      : "+r"(laneid), "+r"(pred)
      : "r"(0xFFFFFFFFu));
  return pred;
}
// max of three (FMNMX3)
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
// (x + hx, y + hy) in one FADD2
__device__ __forceinline__ void add2(float &x, float &y, float hx, float hy) {
  unsigned long long a, b, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(x), "f"(y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(hx), "f"(hy));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(r));
}

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B: rows of 128 bytes, 8-row (1024 B) swizzle
// atoms stacked along M/N (stride byte offset 1024), version 1 (Blackwell).  Fields as in
// cute::UMMA::SmemDescriptor: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version [46,48),
// layout type [61,64) with SWIZZLE_128B = 2.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// The same for a K-major operand WITHOUT swizzle (layout type 0): 8-row x 16-byte core matrices, 8-row groups
// `sbo` bytes apart, the two 16-byte chunks of a K = 16 step `lbo` bytes apart (canonical layout
// ((8,m),(T,2)):((1T,SBO),(1,LBO)) of cute's make_umma_desc<Major::K>).
__device__ __forceinline__ uint64_t umma_desc_nosw(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 = 1 at [4,6), a/b format
// BF16 = 1 at [7,10)/[10,13), both K-major, N>>3 at [17,23), M>>4 at [24,29).
__device__ __forceinline__ uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

static __host__ __device__ inline int kt_kblocks(int d) { return (d + 63) / 64; }
// FOLD: when the padded row has 16 spare columns, -|c|^2/2 rides in the contraction itself: three extra
// BF16 columns of the centroid copy hold h_j split into three 8-bit pieces (24 bits, as exact as the fp32 h),
// the matching columns of every point row hold 1.0 — the accumulator IS the score, the epilogue adds nothing.
static __host__ __device__ inline bool kt_fold(int d) { return ((d + 15) & ~15) + 16 <= kt_kblocks(d) * 64; }
static __host__ __device__ inline int kt_bstages(int d) { return kt_kblocks(d) == 1 ? 4 : 2; }
static __host__ __device__ inline int kt_astages(int d) { return kt_kblocks(d) == 1 ? 4 : 2; }   // short rows: more passes in flight
// plus, when it fits, a copy of h (one float per padded centroid) — see kt_h_in_smem
static __host__ __device__ inline size_t kt_smem_bytes(int d) {
  return (size_t)(kt_astages(d) * 2 + kt_bstages(d)) * kt_kblocks(d) * KT_KB_BYTES   // A stages (two tiles each) + B stages
         + (size_t)KT_LIST * KT_ROWS * 8                                               // survivor lists
         + 256;                                                                        // mbarriers + TMEM slot
}

// ------------------------------------------------------------------ prep: norms + BF16 copy
// h_j = -|chi_j|^2 / 2 (-inf beyond k, so a padding column never scores), max_j |chi_j|^2, and
// the BF16 copy of the row, zero beyond d and beyond k.
__global__ void __launch_bounds__(256) km_tensor_prep_kernel(const float *__restrict__ chi, int k, int d, int kpad,
                                                             float *__restrict__ h, KmState *st,
                                                             __nv_bfloat16 *__restrict__ cbf, int drow, int fold_col,
                                                             __nv_bfloat16 *__restrict__ hfold) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= kpad) return;
  float s = 0.f;
  if (warp < k)
    for (int t = lane; t < d; t += 32) { const float v = chi[(int64_t)warp * d + t]; s = fmaf(v, v, s); }
  for (int t = lane; t < drow; t += 32)
    cbf[(int64_t)warp * drow + t] = __float2bfloat16_rn((warp < k && t < d) ? chi[(int64_t)warp * d + t] : 0.f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __syncwarp();   // the row's zero padding (written above by other lanes) precedes the three columns written below
  if (lane == 0) {
    const float hv = -0.5f * s;
    h[warp] = warp < k ? hv : __int_as_float(0xff800000);
    if (warp < k) atomicMax(reinterpret_cast<unsigned int *>(&st->cnmax2), __float_as_uint(s));   // s >= 0: bits order = value order
    if (fold_col >= 0) {   // kt_fold: h as hi + mid + lo in three BF16 columns (a padding row scores -3e38, never +-inf)
      const float hf = warp < k ? hv : -3.0e38f;
      const __nv_bfloat16 b0 = __float2bfloat16_rn(hf);
      const float r1 = warp < k ? hf - __bfloat162float(b0) : 0.f;
      const __nv_bfloat16 b1 = __float2bfloat16_rn(r1);
      const __nv_bfloat16 b2 = __float2bfloat16_rn(r1 - __bfloat162float(b1));
      cbf[(int64_t)warp * drow + fold_col] = b0;
      cbf[(int64_t)warp * drow + fold_col + 1] = b1;
      cbf[(int64_t)warp * drow + fold_col + 2] = b2;
    }
    if (hfold) {           // foldx: the same three pieces as a 16-byte row of their own (the extra K-step's centroid operand)
      const float hf = warp < k ? hv : -3.0e38f;
      const __nv_bfloat16 b0 = __float2bfloat16_rn(hf);
      const float r1 = warp < k ? hf - __bfloat162float(b0) : 0.f;
      const __nv_bfloat16 b1 = __float2bfloat16_rn(r1);
      const __nv_bfloat16 b2 = __float2bfloat16_rn(r1 - __bfloat162float(b1));
      const __nv_bfloat16 z = __float2bfloat16_rn(0.f);
      __nv_bfloat16 *o = hfold + (int64_t)warp * 8;
      o[0] = b0; o[1] = b1; o[2] = b2; o[3] = z; o[4] = z; o[5] = z; o[6] = z; o[7] = z;
    }
  }
}

// ------------------------------------------------------------------ |x_i| per point, once per upload
// Eight lanes per row; the BF16 window of a point scales with its own norm instead of the job's largest.
__global__ void __launch_bounds__(256) km_row_norm_kernel(const float *__restrict__ x, int64_t n, int d, float *__restrict__ xn) {
  const int lane = threadIdx.x & 31, sl = lane & 7;
  const int64_t grp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3, ngrp = ((int64_t)gridDim.x * blockDim.x) >> 3;
  for (int64_t base = grp0 - (lane >> 3); base < n; base += ngrp) {   // base: first row of this warp's four
    const int64_t i = base + (lane >> 3);
    float s = 0.f;
    if (i < n)
      for (int t = sl; t < d; t += 8) { const float v = x[i * (int64_t)d + t]; s = fmaf(v, v, s); }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (i < n && sl == 0) xn[i] = sqrtf(s);
  }
}

cudaError_t km_launch_row_norms(const float *x, int64_t n, int32_t d, float *xn, int sm_count, cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  const int64_t want = (n * 8 + 255) / 256, cap = (int64_t)sm_count * 16;
  km_row_norm_kernel<<<(int)(want < cap ? want : cap), 256, 0, s>>>(x, n, d, xn);
  (*launches)++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ BF16 copy of x, once per upload
// Row i: the d coordinates rounded to BF16 (round to nearest even — the rounding the error window of the
// coarse stage assumes), zero up to the padded width drow = 64 * kblocks, and — when h rides in the
// contraction (kt_fold) — 1.0 in the three columns that multiply the split h.  The coarse stage gathers
// these rows by TMA straight into its operand tiles: half the bytes of the binary32 rows, no conversion
// in the kernel.  Eight elements (16 bytes) per thread.
__global__ void __launch_bounds__(256) km_tensor_xbf_kernel(const float *__restrict__ x, int64_t n, int d, int drow, int fold_col,
                                                            __nv_bfloat16 *__restrict__ xbf) {
  const int per_row = drow >> 3;
  const int64_t total = n * per_row;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = w / per_row;
    const int c0 = (int)(w - i * per_row) * 8;
    __align__(16) __nv_bfloat16 o[8];
#pragma unroll
    for (int t = 0; t < 8; t++) {
      const int c = c0 + t;
      float v = c < d ? x[i * d + c] : 0.f;
      if (fold_col >= 0 && c >= fold_col && c < fold_col + 3) v = 1.f;
      o[t] = __float2bfloat16_rn(v);
    }
    *reinterpret_cast<uint4 *>(xbf + i * drow + c0) = *reinterpret_cast<const uint4 *>(o);
  }
}

size_t km_tensor_xbf_bytes(int64_t n, int32_t d) { return (size_t)(n > 0 ? n : 1) * kt_kblocks(d) * 64 * sizeof(__nv_bfloat16); }
cudaError_t km_launch_xbf(const float *x, int64_t n, int32_t d, void *xbf, int sm_count, cudaStream_t s, int64_t *launches) {
  if (n == 0) return cudaSuccess;
  const int drow = kt_kblocks(d) * 64;
  const int64_t want = (n * (drow / 8) + 255) / 256, cap = (int64_t)sm_count * 16;
  km_tensor_xbf_kernel<<<(int)(want < cap ? want : cap), 256, 0, s>>>(x, n, d, drow, kt_fold(d) ? ((d + 15) & ~15) : -1,
                                                                       (__nv_bfloat16 *)xbf);
  (*launches)++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ order: counting sort by previous label
// Bucket k holds the never-assigned points (label < 0).  Any order inside a bucket will do.
#define KO_THREADS 256
#define KO_SMEM_BUCKETS 8192

__global__ void __launch_bounds__(KO_THREADS) km_order_hist_kernel(const int32_t *__restrict__ labels, int64_t n, int k,
                                                                   int64_t per_block, unsigned *__restrict__ hist) {
  extern __shared__ unsigned ko_sm[];
  const bool sm = k + 1 <= KO_SMEM_BUCKETS;
  if (sm) {
    for (int j = threadIdx.x; j <= k; j += KO_THREADS) ko_sm[j] = 0;
    __syncthreads();
  }
  const int64_t b0 = (int64_t)blockIdx.x * per_block, b1 = min(n, b0 + per_block);
  for (int64_t i = b0 + threadIdx.x; i < b1; i += KO_THREADS) {
    const int l = labels[i];
    const int j = (l < 0 || l >= k) ? k : l;
    atomicAdd(sm ? &ko_sm[j] : &hist[j], 1u);
  }
  if (sm) {
    __syncthreads();
    for (int j = threadIdx.x; j <= k; j += KO_THREADS)
      if (ko_sm[j]) atomicAdd(&hist[j], ko_sm[j]);
  }
}

// exclusive scan of hist[0..k] in place (one block)
__global__ void __launch_bounds__(1024) km_order_scan_kernel(unsigned *__restrict__ hist, int k) {
  __shared__ unsigned warp_tot[32];
  __shared__ unsigned carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base <= k; base += 1024) {
    const int j = base + threadIdx.x;
    const unsigned v = j <= k ? hist[j] : 0;
    unsigned incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      unsigned w = warp_tot[threadIdx.x], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, wi, o);
        if (threadIdx.x >= o) wi += t;
      }
      warp_tot[threadIdx.x] = wi - w;   // exclusive prefix of the warp totals
    }
    __syncthreads();
    const unsigned excl = carry + warp_tot[threadIdx.x >> 5] + incl - v;
    if (j <= k) hist[j] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
}

// every block places its own contiguous share of the points: one global atomic per (block, label present)
__global__ void __launch_bounds__(KO_THREADS) km_order_scatter_kernel(const int32_t *__restrict__ labels, int64_t n, int k,
                                                                      int64_t per_block, unsigned *__restrict__ cursor,
                                                                      unsigned *__restrict__ order) {
  extern __shared__ unsigned ko_sm[];
  const bool sm = k + 1 <= KO_SMEM_BUCKETS;
  const int64_t b0 = (int64_t)blockIdx.x * per_block, b1 = min(n, b0 + per_block);
  if (!sm) {
    for (int64_t i = b0 + threadIdx.x; i < b1; i += KO_THREADS) {
      const int l = labels[i];
      order[atomicAdd(&cursor[(l < 0 || l >= k) ? k : l], 1u)] = (unsigned)i;
    }
    return;
  }
  for (int j = threadIdx.x; j <= k; j += KO_THREADS) ko_sm[j] = 0;
  __syncthreads();
  for (int64_t i = b0 + threadIdx.x; i < b1; i += KO_THREADS) {
    const int l = labels[i];
    atomicAdd(&ko_sm[(l < 0 || l >= k) ? k : l], 1u);
  }
  __syncthreads();
  for (int j = threadIdx.x; j <= k; j += KO_THREADS) {
    const unsigned c = ko_sm[j];
    ko_sm[j] = c ? atomicAdd(&cursor[j], c) : 0u;   // this block's range of bucket j starts here
  }
  __syncthreads();
  for (int64_t i = b0 + threadIdx.x; i < b1; i += KO_THREADS) {
    const int l = labels[i];
    order[atomicAdd(&ko_sm[(l < 0 || l >= k) ? k : l], 1u)] = (unsigned)i;
  }
}

// ------------------------------------------------------------------ coarse stage
struct KtParams {
  const float *x;          // [n][d]
  int64_t n;
  int d, k;
  const unsigned *order;   // [n] row processed at position r, or nullptr = r itself
  const int32_t *labels;   // previous labels
  const float *xn;         // [n] |x_i| (binary32, km_row_norm_kernel)
  const float *h;          // [kpad] -|chi_j|^2 / 2, -inf beyond k
  int h_in_smem;           // the launcher found room for a shared-memory copy of h
  int foldx;               // h rides in one extra K-step whose operands are small no-swizzle tiles (rows without spare columns)
  const KmState *st;       // st->cnmax2 = max_j |chi_j|^2
  float xnmax;             // upper bound of |x_i| over the job
  int *work;               // worklist entries, KT_WORK_INTS ints each
  unsigned *work_count;
  int all_to_work;         // full-update mode: every point goes to the exact stage
};

// rare: the survivor list of a point is full.  Drop what the tighter threshold has since excluded;
// if it is still full keep the KT_LIST best scores and remember the best one that had to go
// (`lostmax`): the list is complete exactly if that one ends up below the final threshold.
// Returns {entries now in the list, lostmax} — by value, so that neither lives in local memory in the caller.
__device__ __noinline__ int2 kt_insert_full(int2 *lst, int row, float thr, int col, float val, float lostmax) {
  int w = 0;
  for (int s = 0; s < KT_LIST; s++) {
    const int2 e = lst[s * KT_ROWS + row];
    if (__int_as_float(e.y) >= thr) lst[(w++) * KT_ROWS + row] = e;
  }
  if (w < KT_LIST) {
    lst[(w++) * KT_ROWS + row] = make_int2(col, __float_as_int(val));
    return make_int2(w, __float_as_int(lostmax));
  }
  int smin = 0;
  float vmin = __int_as_float(lst[row].y);
  for (int s = 1; s < KT_LIST; s++) {
    const float v = __int_as_float(lst[s * KT_ROWS + row].y);
    if (v < vmin) { vmin = v; smin = s; }
  }
  if (val > vmin) {
    lst[smin * KT_ROWS + row] = make_int2(col, __float_as_int(val));
    lostmax = fmaxf(lostmax, vmin);
  } else {
    lostmax = fmaxf(lostmax, val);
  }
  return make_int2(KT_LIST, __float_as_int(lostmax));
}

// rare (a point with more than one recorded column): keep the entries within the window of the FINAL
// maximum, sort them by centroid index (tie rule: lowest j first) and write the indices back into the
// list's first slots.  Returns how many (1..8), or -2 when the exact stage has to scan all k: more than
// eight, a dropped score that the final threshold would have kept, or none at all (a non-finite score).
__device__ __noinline__ int kt_collect(int2 *lst, int row, int cnt, float thr, float lostmax) {
  int out[8];
  int w = 0;
  for (int s2 = 0; s2 < cnt; s2++) {
    const int2 e = lst[s2 * KT_ROWS + row];
    if (__int_as_float(e.y) >= thr) {
      if (w < 8) out[w] = e.x;
      w++;
    }
  }
  if (lostmax >= thr || w > 8 || w == 0) return -2;
  for (int a = 1; a < w; a++)
    for (int q = a; q > 0 && out[q] < out[q - 1]; q--) { const int tmp = out[q]; out[q] = out[q - 1]; out[q - 1] = tmp; }
  for (int s2 = 0; s2 < w; s2++) lst[s2 * KT_ROWS + row].x = out[s2];
  return w;
}

// HSMEM: h has a copy in shared memory, read with ld.shared (its own queue — a generic or global load of h
// would wait in line behind the producers' stream of point loads)
template <bool HSMEM, bool FOLD>
__global__ void __launch_bounds__(KT_THREADS, 1) km_tensor_coarse_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                                         const __grid_constant__ CUtensorMap tmH, const KtParams p) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];   // SWIZZLE_128B operand tiles need 1024-byte alignment
  const int d = p.d, kblocks = kt_kblocks(d), bstages = kt_bstages(d);
  const bool foldx = FOLD && p.foldx != 0;
  const int nks = ((d + 15) >> 4) + ((FOLD && !foldx) ? 1 : 0);  // K = 16 MMA steps over the (zero-padded) dimension [+ the h columns]
  const uint32_t tile_bytes = (uint32_t)kblocks * KT_KB_BYTES;   // one 128-row operand tile
  const int astages = kt_astages(d);
  const uint32_t sA = smem_u32(smem_raw);                        // [astages][2 tiles]
  const uint32_t sB = sA + astages * 2 * tile_bytes;             // [bstages]
  unsigned char *after = smem_raw + (size_t)(astages * 2 + bstages) * tile_bytes;
  int2 *lst = reinterpret_cast<int2 *>(after);                   // [KT_LIST][KT_ROWS] (centroid, score bits)
  uint64_t *bars = reinterpret_cast<uint64_t *>(lst + KT_LIST * KT_ROWS);
  uint64_t *a_full = bars, *a_empty = bars + 4, *b_full = bars + 8, *b_empty = bars + 12, *acc_full = bars + 16,
           *acc_empty = bars + 18;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 20);
  float *hs = reinterpret_cast<float *>(bars + 32);             // [kpad] copy of h when p.h_in_smem
  const int kpad_h = ((p.k + KT_N - 1) / KT_N) * KT_N;
  if (HSMEM && !FOLD)
    for (int j = threadIdx.x; j < kpad_h; j += KT_THREADS) hs[j] = p.h[j];
  const uint32_t hs_addr = smem_u32(hs);
  // foldx: the extra K-step's operands, no swizzle, rows 16 bytes apart.  Chunk 0 of the point side is the same for
  // every point ({1, 1, 1, 0...}: it multiplies the three BF16 pieces of h), chunk 0 of the centroid side arrives by
  // TMA with each centroid tile, chunk 1 of both is one block of zeros.
  const uint32_t sAf = hs_addr, sBf = sAf + 2048, sZ = sBf + 2048 * (uint32_t)bstages;
  if (foldx) {
    for (int r = threadIdx.x; r < KT_M; r += KT_THREADS) {
      reinterpret_cast<uint4 *>(hs)[r] = make_uint4(0x3F803F80u, 0x00003F80u, 0u, 0u);
      reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(hs) + 2048 * (1 + bstages))[r] = make_uint4(0u, 0u, 0u, 0u);
    }
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // (the point tiles arrive whole by TMA from the BF16 copy of x: zero padding and — FOLD — the 1.0 columns are part of its rows)
  fence_proxy_async();
  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; s++) { mbar_init(&a_full[s], KT_GATHER_WARPS); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < 4; s++) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < 2; s++) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], KT_EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) tmem_alloc(tmem_slot, 512);              // 2 point tiles x 2 buffers x 128 columns
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int ntiles = (p.k + KT_N - 1) / KT_N;
  const int64_t passes = (p.n + KT_ROWS - 1) / KT_ROWS;

  // The centroid tiles of a pass stream by starting with the tile that holds the previous label of the
  // pass's first point (the index list is label-ordered: nearly all 256 points share it), so the running
  // maximum is tight from the first chunk on.  Every role derives the same rotation from global memory.
  // (Both look-ups are two dependent loads — index list, then label.  They are software-pipelined:
  //  the index of pass p+2 and the label behind the index of pass p+1 are requested during pass p, so
  //  neither latency is ever waited for.)
  auto idx_at = [&](int64_t ps, int rowoff) -> int {          // level 1: which row of x sits at this position
    const int64_t r = ps * KT_ROWS + rowoff;
    if (ps >= passes || r >= p.n) return -1;
    return p.order ? (int)__ldg(p.order + r) : (int)r;
  };
  // level 2: previous label of a pass's first point — the RAW load; it is range-checked (start_ok) only where
  // it is used, a pass later, so that nothing waits for the load in the pass that issues it
  auto start_of = [&](int gi0) -> int { return gi0 < 0 ? 0 : __ldg(p.labels + gi0); };
  auto start_ok = [&](int lab) -> int { return (lab >= 0 && lab < p.k) ? lab : 0; };

  if (warp >= KT_WARP_GATHER && warp < KT_WARP_MMA) {
    // ===================== TMA: the pass's 256 rows of the BF16 copy of x, gathered through the index list =====================
    // 64 lanes (16 of each gather warp) own four consecutive rows of the pass each: a lane reads their row
    // numbers (16 contiguous bytes of the list, one pass ahead) and issues one tile::gather4 per 64-column block.
    // The rows land in the K-major SWIZZLE_128B layout the MMA reads; nothing is converted or stored by any thread.
    constexpr int GL = 64 / KT_GATHER_WARPS;                        // issuing lanes per gather warp
    const bool on = lane < GL;
    const int gl = (warp - KT_WARP_GATHER) * GL + (on ? lane : 0);  // 0 .. 63: rows 4 gl .. 4 gl + 3
    auto rows_of = [&](int64_t ps, int (&idx)[4]) {
      const int64_t r0 = ps * KT_ROWS + gl * 4;
#pragma unroll
      for (int i = 0; i < 4; i++) idx[i] = (int)min(p.n, r0 + i);            // past the end: row n, out of bounds -> zero fill
      if (ps < passes && p.order != nullptr) {
        if (r0 + 4 <= p.n) {
          const uint4 a4 = __ldg(reinterpret_cast<const uint4 *>(p.order + r0));
          idx[0] = (int)a4.x; idx[1] = (int)a4.y; idx[2] = (int)a4.z; idx[3] = (int)a4.w;
        } else {
#pragma unroll
          for (int i = 0; i < 4; i++)
            if (r0 + i < p.n) idx[i] = (int)__ldg(p.order + r0 + i);
        }
      }
    };
    int it = 0;
    int idx[4];
    rows_of(blockIdx.x, idx);
    const uint32_t lane_dst = (uint32_t)((gl >> 5) * tile_bytes + (gl & 31) * 4 * 128);   // rows 4 gl .. : tile gl >> 5, row (gl & 31) * 4 in it
    for (int64_t ps = blockIdx.x; ps < passes; ps += gridDim.x, it++) {
      const int s = it % astages;
      mbar_wait_relaxed(&a_empty[s], ((it / astages) & 1) ^ 1);
      if (lane == 0) mbar_expect_tx(&a_full[s], 2 * tile_bytes / KT_GATHER_WARPS);   // this warp's share of the two point tiles, every k-block
      __syncwarp();
      const uint32_t dst = sA + (uint32_t)s * 2 * tile_bytes + lane_dst;
      if (on)
        for (int kb = 0; kb < kblocks; kb++)
          tma_gather4(dst + kb * KT_KB_BYTES, &tmA, kb * 64, idx[0], idx[1], idx[2], idx[3], &a_full[s]);
      rows_of(ps + gridDim.x, idx);             // the next pass's row numbers travel while this one is consumed
    }
  } else if (warp == KT_WARP_TMA) {
    // ===================== TMA: centroid tiles (the whole warp walks the loop, one elected lane issues) =====================
    uint32_t g = 0;
    int sg1 = idx_at((int64_t)blockIdx.x + gridDim.x, 0);
    int start_next = start_of(idx_at(blockIdx.x, 0));
    for (int64_t ps = blockIdx.x; ps < passes; ps += gridDim.x) {
      const int t0 = start_ok(start_next) / KT_N;
      start_next = start_of(sg1);                               // its index arrived a pass ago
      sg1 = idx_at(ps + 2 * (int64_t)gridDim.x, 0);
      for (int t = 0; t < ntiles; t++, g++) {
        const uint32_t sb = g % bstages;
        int tt = t0 + t;
        if (tt >= ntiles) tt -= ntiles;
        mbar_wait_relaxed(&b_empty[sb], ((g / bstages) & 1) ^ 1);
        if (elect_one_sync()) {
          mbar_expect_tx(&b_full[sb], tile_bytes + (foldx ? 2048u : 0u));
          for (int kbi = 0; kbi < kblocks; kbi++)
            tma_load_2d(sB + sb * tile_bytes + kbi * KT_KB_BYTES, &tmB, kbi * 64, tt * KT_N, &b_full[sb]);
          if (foldx) tma_load_2d(sBf + sb * 2048, &tmH, 0, tt * KT_N, &b_full[sb]);   // the tile's 128 x {h_hi, h_mid, h_lo, 0 x 5}
        }
        __syncwarp();
      }
    }
  } else if (warp == KT_WARP_MMA) {
    // ===================== MMA issue (whole warp in the loop, one elected lane issues) =====================
    const uint32_t idesc = umma_idesc_bf16(KT_M, KT_N);
    uint32_t g = 0;
    int it = 0;
    for (int64_t ps = blockIdx.x; ps < passes; ps += gridDim.x, it++) {
      const int sa = it % astages;
      mbar_wait_relaxed(&a_full[sa], (it / astages) & 1);
      for (int t = 0; t < ntiles; t++, g++) {
        const uint32_t sb = g % bstages, b = g & 1;
        mbar_wait_relaxed(&b_full[sb], (g / bstages) & 1);
        mbar_wait_relaxed(&acc_empty[b], ((g >> 1) & 1) ^ 1);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint64_t bdesc = umma_desc_sw128(sB + sb * tile_bytes);
#pragma unroll
          for (int m = 0; m < 2; m++) {
            const uint64_t adesc = umma_desc_sw128(sA + (sa * 2 + m) * tile_bytes);
            const uint32_t tacc = tmem_base + (uint32_t)((m * 2 + b) * KT_N);
#pragma unroll
            for (int ks = 0; ks < 8; ks++) {
              if (ks < nks) {
                // K step ks: k-block ks >> 2 (16 KiB apart), 32 bytes per step inside it; descriptor addresses are in 16-byte units
                const uint64_t koff = (uint64_t)((ks >> 2) * (KT_KB_BYTES >> 4) + (ks & 3) * 2);
                umma_bf16(tacc, adesc + koff, bdesc + koff, idesc, ks != 0);
              }
            }
            // + h: one more K = 16 step, {1, 1, 1, 0...} against the tile's {h_hi, h_mid, h_lo, 0...}
            if (foldx)
              umma_bf16(tacc, umma_desc_nosw(sAf, sZ - sAf, 128), umma_desc_nosw(sBf + sb * 2048, sZ - (sBf + sb * 2048), 128), idesc, 1);
          }
          umma_commit(&b_empty[sb]);          // the centroid tile may be overwritten once these MMAs are done
          umma_commit(&acc_full[b]);          // ... and the epilogue may read the accumulators
          if (t == ntiles - 1) umma_commit(&a_empty[sa]);
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue: 8 warps, one TMEM lane (= one point) per thread =====================
    const int m = warp >> 2, quad = warp & 3;
    const int row = m * KT_M + quad * 32 + lane;          // row within the pass
    // BF16 keeps 8 significant bits: unit roundoff u = 2^-8 per operand, a product errs by (2u + u^2)|x_t c_t|,
    // the score x~.c~ - |c|^2/2 by E/2 with E = (2^-6 + 2^-15) |x||c|.  Survivors: within
    // 2 * E/2 of the best score (plus slack for the fp32 accumulation and the rounding of h).
    // Per point: |x| is its own norm (computed once per upload, rounded up), |c| <= max_j |chi_j|.
    const float cnmax2 = p.st->cnmax2, cnmax = sqrtf(cnmax2);
    const float NEG_INF = __int_as_float(0xff800000);
    uint32_t g = 0;
    int it = 0;
    // this point's row of x (two passes ahead), its previous label and norm (one pass ahead)
    auto point_of = [&](int gi_, int &lab_, float &xn_) {
      lab_ = -1; xn_ = 0.f;
      if (gi_ >= 0) { lab_ = __ldg(p.labels + gi_); xn_ = __ldg(p.xn + gi_); }
    };
    int gi_next = idx_at(blockIdx.x, row), gi_n2 = idx_at((int64_t)blockIdx.x + gridDim.x, row);
    int sg1 = idx_at((int64_t)blockIdx.x + gridDim.x, 0);
    int lab_next;
    float xn_next;
    point_of(gi_next, lab_next, xn_next);
    int start_next = start_of(idx_at(blockIdx.x, 0));
    for (int64_t ps = blockIdx.x; ps < passes; ps += gridDim.x, it++) {
      const int gi = gi_next, lab = lab_next, start = start_ok(start_next);
      const bool live = gi >= 0;
      const float xn = fminf(xn_next * 1.0001f, p.xnmax);      // |x_i| (km_row_norm_kernel), rounded up
      gi_next = gi_n2;
      point_of(gi_next, lab_next, xn_next);                     // their index arrived a pass ago
      start_next = start_of(sg1);
      gi_n2 = idx_at(ps + 2 * (int64_t)gridDim.x, row);
      sg1 = idx_at(ps + 2 * (int64_t)gridDim.x, 0);
      const int t0 = start / KT_N, c0 = (start % KT_N) >> 5;   // first tile, and the first chunk within it
      const float win = 0.0156556f * xn * cnmax + 1e-6f * (xn * xn + cnmax2);
      float runmax = NEG_INF, thr = NEG_INF;
      int cnt = 0;
      float lostmax = NEG_INF;                // best score the full list had to let go

      // one chunk of 32 columns: scores, chunk maximum, and — only if some lane's maximum comes within
      // the window of its running maximum — a look at the columns one by one
      auto chunk = [&](uint32_t (&v)[32], int col0) {
        float s[32];
        if (FOLD) {
#pragma unroll
          for (int i = 0; i < 32; i++) s[i] = __uint_as_float(v[i]);   // the accumulator is the score
        }
#pragma unroll
        for (int i4 = 0; i4 < (FOLD ? 0 : 8); i4++) {
          float4 h4;                                  // warp-uniform address: one broadcast
          if (HSMEM)
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(h4.x), "=f"(h4.y), "=f"(h4.z), "=f"(h4.w)
                         : "r"(hs_addr + (uint32_t)(col0 + i4 * 4) * 4u));
          else
            h4 = __ldg(reinterpret_cast<const float4 *>(p.h + col0) + i4);
          s[i4 * 4 + 0] = __uint_as_float(v[i4 * 4 + 0]); s[i4 * 4 + 1] = __uint_as_float(v[i4 * 4 + 1]);
          s[i4 * 4 + 2] = __uint_as_float(v[i4 * 4 + 2]); s[i4 * 4 + 3] = __uint_as_float(v[i4 * 4 + 3]);
          add2(s[i4 * 4 + 0], s[i4 * 4 + 1], h4.x, h4.y);
          add2(s[i4 * 4 + 2], s[i4 * 4 + 3], h4.z, h4.w);
        }
        float mx[11];
#pragma unroll
        for (int i = 0; i < 10; i++) mx[i] = max3(s[3 * i], s[3 * i + 1], s[3 * i + 2]);
        mx[10] = fmaxf(s[30], s[31]);
        const float m1 = max3(mx[0], mx[1], mx[2]), m2 = max3(mx[3], mx[4], mx[5]), m3 = max3(mx[6], mx[7], mx[8]);
        const float cm = max3(max3(m1, m2, m3), mx[9], mx[10]);
        if (cm > runmax) { runmax = cm; thr = cm - win; }
        if (!__any_sync(0xffffffffu, live && cm >= thr)) return;
        unsigned mk[4] = {0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < 32; i++)
          if (s[i] >= thr) mk[i & 3] |= 1u << i;
        const unsigned mask = live ? (mk[0] | mk[1] | mk[2] | mk[3]) : 0u;
        unsigned U = __reduce_or_sync(0xffffffffu, mask);
        while (U) {                                 // warp-uniform walk over the columns any lane hit
          const int i = __ffs(U) - 1;
          U &= U - 1;
          float val;
          switch (i) {
#define KT_CASE(I) case I: val = s[I]; break;
            KT_CASE(0) KT_CASE(1) KT_CASE(2) KT_CASE(3) KT_CASE(4) KT_CASE(5) KT_CASE(6) KT_CASE(7)
            KT_CASE(8) KT_CASE(9) KT_CASE(10) KT_CASE(11) KT_CASE(12) KT_CASE(13) KT_CASE(14) KT_CASE(15)
            KT_CASE(16) KT_CASE(17) KT_CASE(18) KT_CASE(19) KT_CASE(20) KT_CASE(21) KT_CASE(22) KT_CASE(23)
            KT_CASE(24) KT_CASE(25) KT_CASE(26) KT_CASE(27) KT_CASE(28) KT_CASE(29) KT_CASE(30)
#undef KT_CASE
            default: val = s[31]; break;
          }
          if ((mask >> i) & 1u) {
            if (cnt < KT_LIST) lst[(cnt++) * KT_ROWS + row] = make_int2(col0 + i, __float_as_int(val));
            else {
              const int2 r2 = kt_insert_full(lst, row, thr, col0 + i, val, lostmax);
              cnt = r2.x;
              lostmax = __int_as_float(r2.y);
            }
          }
        }
      };

      for (int t = 0; t < ntiles; t++, g++) {
        const uint32_t b = g & 1;
        int tt = t0 + t;                          // the tile that arrives t-th (same rotation as the TMA thread's)
        if (tt >= ntiles) tt -= ntiles;
        // chunk order inside a tile: start where the previous label sits (it matters in the first tile only)
        const int q0 = c0 * 32, q1 = ((c0 + 1) & 3) * 32, q2 = ((c0 + 2) & 3) * 32, q3 = ((c0 + 3) & 3) * 32;
        mbar_wait(&acc_full[b], (g >> 1) & 1);
        tc_fence_after();
        const uint32_t tcol = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)((m * 2 + b) * KT_N);
        uint32_t va[32], vb[32];
        __syncwarp();                             // tcgen05.ld / wait::ld are warp-collective: issue them converged
        tmem_ld32_issue(tcol + q0, va);
        tmem_ld_wait();
        tmem_ld32_issue(tcol + q1, vb);           // the TMEM read of chunk c+1 is in flight while chunk c is processed
        chunk(va, tt * KT_N + q0);
        __syncwarp();
        tmem_ld_wait();
        tmem_ld32_issue(tcol + q2, va);
        chunk(vb, tt * KT_N + q1);
        __syncwarp();
        tmem_ld_wait();
        tmem_ld32_issue(tcol + q3, vb);
        chunk(va, tt * KT_N + q2);
        __syncwarp();
        tmem_ld_wait();
        tc_fence_before();
        if (lane == 0) mbar_arrive(&acc_empty[b]);   // this warp holds its last columns in registers: the buffer is free
        chunk(vb, tt * KT_N + q3);
      }

      // ---- survivors of this point: within the window of the final maximum, ascending index.
      // Nearly always exactly one column was ever recorded — the final maximum's (a column is recorded
      // whenever it raises the maximum) — and nothing was dropped: nothing to filter or sort.
      int nout = 1;
      const bool simple = cnt == 1 && !(lostmax >= thr);
      if (!__all_sync(0xffffffffu, simple || !live)) nout = simple ? 1 : kt_collect(lst, row, cnt, thr, lostmax);
      // a lone survivor equal to the previous label: the point keeps its label, its sums stay where they are
      const bool need = live && (p.all_to_work || nout != 1 || lst[row].x != lab);
      const unsigned bal = __ballot_sync(0xffffffffu, need);
      if (bal) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(p.work_count, (unsigned)__popc(bal));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (need) {
          int4 *e = reinterpret_cast<int4 *>(p.work + (int64_t)(base + __popc(bal & ((1u << lane) - 1u))) * KT_WORK_INTS);
          int o[8];
#pragma unroll
          for (int s2 = 0; s2 < 8; s2++) o[s2] = s2 < nout ? lst[s2 * KT_ROWS + row].x : -1;
          e[0] = make_int4(gi, nout, 0, 0);
          e[1] = make_int4(o[0], o[1], o[2], o[3]);
          e[2] = make_int4(o[4], o[5], o[6], o[7]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------ exact stage (worklist only)
// Eight lanes per worklist entry (four entries per warp); each lane owns every eighth float4 of the
// row.  Survivors are re-evaluated with direct binary32 differences; the binary64 re-score (oracle
// operation order, one lane, sequential) settles what the error bound cannot.
__device__ __forceinline__ float kt_sum8(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}

// rare, out of line (keeps the main loop's register count low): the oracle's own arithmetic over the
// candidates — binary64, dims ascending, separate multiply and add (A1/A2); strict <, lowest index first
__device__ __noinline__ int kt_exact_rescore(const float *xr, const double *__restrict__ c64, int d, int ncand, bool scan_all,
                                             int4 cw, int4 cw2) {
  double best = __longlong_as_double(0x7FF0000000000000LL);
  int lab = 0;
  for (int s = 0; s < ncand; s++) {
    const int j = scan_all ? s : s == 0 ? cw.x : s == 1 ? cw.y : s == 2 ? cw.z : s == 3 ? cw.w : s == 4 ? cw2.x : s == 5 ? cw2.y : s == 6 ? cw2.z : cw2.w;
    const double dj = km_dist2_f64(xr, c64 + (int64_t)j * d, d);
    if (dj < best) { best = dj; lab = j; }
  }
  return lab;
}

// a point that changed its label (or every point, in full-update mode) moves between the order-free sums;
// called by the 8 lanes of its group: consecutive lanes take consecutive dimensions
__device__ __noinline__ void kt_exact_move(const float *xr, int d, int sl, int lab, int from, long long *__restrict__ acc,
                                           int64_t sa, int64_t sb, int64_t cnt, double scale_a, double scale_b) {
  for (int t = sl; t < d; t += 8) {
    long long a, b;
    km_fixed_split((double)xr[t], scale_a, scale_b, a, b);
    km_atomic_add_ll(acc + sa + (int64_t)lab * d + t, a);
    km_atomic_add_ll(acc + sb + (int64_t)lab * d + t, b);
    if (from >= 0) {
      km_atomic_add_ll(acc + sa + (int64_t)from * d + t, -a);
      km_atomic_add_ll(acc + sb + (int64_t)from * d + t, -b);
    }
  }
  if (sl == 0) {
    km_atomic_add_ll(acc + cnt + lab, 1);
    if (from >= 0) km_atomic_add_ll(acc + cnt + from, -1);
  }
}

// QN: float4 pieces of a row per lane (d <= 32 * QN).  The loop is a chain of dependent loads (worklist
// entry -> row of x and previous label -> candidate centroid rows): the next entry is requested one
// iteration ahead, and the register budget is kept low enough for 3-4 blocks per SM.
template <int QN>
__global__ void __launch_bounds__(256, 3)
km_tensor_exact_kernel(const float *__restrict__ x, int d, int k, const int *__restrict__ work,
                       const unsigned *__restrict__ work_count, const float *__restrict__ chi,
                       const double *__restrict__ c64, int32_t *__restrict__ labels, long long *__restrict__ acc,
                       const KmState *__restrict__ st, int incremental) {
  const KmAccLayout L = km_acc_layout(k, d);
  const float tau2 = st->tau_hi;
  const float rho2 = 2.5f * (float)(d + 6) * 0x1.0p-24f;
  const int lane = threadIdx.x & 31, g = lane >> 3, sl = lane & 7;
  const bool vec = (d & 3) == 0;
  const int64_t nwork = *work_count;
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  KmThreadScalars S;
  // this lane's pieces of a row: float4 index q * 8 + sl, q = 0..QN-1, zero beyond d
  auto load_row = [&](const float *row, float4 (&v)[QN]) {
#pragma unroll
    for (int q = 0; q < QN; q++) {
      const int e = (q * 8 + sl) * 4;
      v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (e < d) {
        if (vec) {
          v[q] = __ldg(reinterpret_cast<const float4 *>(row + e));
        } else {
          v[q].x = row[e];
          if (e + 1 < d) v[q].y = row[e + 1];
          if (e + 2 < d) v[q].z = row[e + 2];
          if (e + 3 < d) v[q].w = row[e + 3];
        }
      }
    }
  };
  auto load_entry = [&](int64_t w, int4 &hd, int4 &cw, int4 &cw2) {
    hd = make_int4(0, 0, 0, 0); cw = make_int4(-1, -1, -1, -1); cw2 = cw;
    if (w < nwork) {
      const int4 *e = reinterpret_cast<const int4 *>(work + w * KT_WORK_INTS);
      hd = __ldg(e); cw = __ldg(e + 1); cw2 = __ldg(e + 2);
    }
  };
  int4 hd_n, cw_n, cw2_n;
  load_entry(warp0 * 4 + g, hd_n, cw_n, cw2_n);
  for (int64_t base = warp0 * 4; base < nwork; base += nwarps * 4) {
    const int64_t w = base + g;
    const bool live = w < nwork;
    const int4 hd = hd_n, cw = cw_n, cw2 = cw2_n;
    const int64_t i = hd.x;
    const float *xr = x + (live ? i : 0) * (int64_t)d;
    float4 xq[QN];
    int old = 0;
    if (live) { load_row(xr, xq); old = labels[i]; }
    load_entry(w + nwarps * 4, hd_n, cw_n, cw2_n);      // the next iteration's entry travels while this one is worked on
    const bool scan_all = hd.y == -2;
    const int ncand = !live ? 0 : scan_all ? k : hd.y;
    auto pick = [&](int s) { return s == 0 ? cw.x : s == 1 ? cw.y : s == 2 ? cw.z : s == 3 ? cw.w : s == 4 ? cw2.x : s == 5 ? cw2.y : s == 6 ? cw2.z : cw2.w; };
    // a lone survivor is the argmin by the coarse bound: no binary32 pass needed
    const int nloop = __reduce_max_sync(0xffffffffu, ncand > 1 ? ncand : 0);
    float b1 = __int_as_float(0x7f800000), b2 = b1;
    int i1 = scan_all ? 0 : cw.x;
    for (int s = 0; s < nloop; s++) {
      const bool act = ncand > 1 && s < ncand;
      const int j = scan_all ? s : pick(s);
      float part = 0.f;
      if (act) {
        float4 cv[QN];
        load_row(chi + (int64_t)j * d, cv);
#pragma unroll
        for (int q = 0; q < QN; q++) {
          float df = xq[q].x - cv[q].x; part = fmaf(df, df, part);
          df = xq[q].y - cv[q].y; part = fmaf(df, df, part);
          df = xq[q].z - cv[q].z; part = fmaf(df, df, part);
          df = xq[q].w - cv[q].w; part = fmaf(df, df, part);
        }
      }
      part = kt_sum8(part);
      if (act) { if (part < b1) { b2 = b1; b1 = part; i1 = j; } else if (part < b2) b2 = part; }
    }
    int lab = i1;
    const bool clear = ncand <= 1 || (b2 - b1) > fmaf(rho2, b2, tau2);
    if (live && !clear && sl == 0) {
      lab = kt_exact_rescore(xr, c64, d, ncand, scan_all, cw, cw2);
      S.rescored++;
    }
    lab = __shfl_sync(0xffffffffu, lab, g * 8);     // the group's lane 0 holds the decision
    if (live && sl == 0 && lab != old) { S.changed++; labels[i] = lab; }
    if (live && (incremental ? (lab != old) : true))
      kt_exact_move(xr, d, sl, lab, incremental ? old : -1, acc, L.sa, L.sb, L.cnt, st->scale_a, st->scale_b);
  }
  km_flush_scalars(S, acc, L);
}

// ------------------------------------------------------------------ launcher
bool km_tensor_path_supported(int32_t d, int32_t k, int dtype, int max_smem_optin) {
  if (dtype != KM_DTYPE_F32 || d < 9 || d > 128 || k < 2) return false;
  return kt_smem_bytes(d) + 1024 <= (size_t)max_smem_optin;
}

static int kt_kpad(int32_t k) { return ((k + KT_N - 1) / KT_N) * KT_N; }
// the BF16 centroid rows, then one 16-byte row per centroid with the three BF16 pieces of h (foldx)
size_t km_tensor_cbf_bytes(int32_t d, int32_t k) { return (size_t)kt_kpad(k) * (kt_kblocks(d) * 64 + 8) * sizeof(__nv_bfloat16); }
size_t km_tensor_h_bytes(int32_t k) { return (size_t)kt_kpad(k) * sizeof(float); }
size_t km_tensor_work_bytes(int64_t n) { return (size_t)(n > 0 ? n : 1) * KT_WORK_INTS * sizeof(int); }

// cuTensorMapEncodeTiled, fetched from the driver through the runtime: nothing links against libcuda
typedef CUresult (*kt_encode_tiled_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                      const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                      CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static cudaError_t kt_encoder(kt_encode_tiled_t *out) {
  static kt_encode_tiled_t fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess) return e;
    if (q != cudaDriverEntryPointSuccess || !p) return cudaErrorNotSupported;
    fn = (kt_encode_tiled_t)p;
  }
  *out = fn;
  return cudaSuccess;
}

cudaError_t km_launch_assign_tensor(const KmAssignArgs &a, int64_t *launches) {
  const int kpad = kt_kpad(a.k);
  const int drow = kt_kblocks(a.d) * 64;
  cudaError_t e = cudaMemsetAsync(&a.state->cnmax2, 0, sizeof(float), a.stream);
  if (e != cudaSuccess) return e;
  const bool fold = kt_fold(a.d);
  __nv_bfloat16 *hfold = (__nv_bfloat16 *)a.t_cbf + (size_t)kpad * drow;
  const bool foldx = !fold && kt_smem_bytes(a.d) + 2048 * (2 + kt_bstages(a.d)) + 1024 <= (size_t)a.max_smem_optin;
  km_tensor_prep_kernel<<<(kpad * 32 + 255) / 256, 256, 0, a.stream>>>(a.chi, a.k, a.d, kpad, a.t_cn, a.state,
                                                                       (__nv_bfloat16 *)a.t_cbf, drow, fold ? ((a.d + 15) & ~15) : -1,
                                                                       foldx ? hfold : nullptr);
  (*launches)++;
  if (a.n == 0) return cudaGetLastError();
  if (a.n >= ((int64_t)1 << 31)) return cudaErrorInvalidValue;   // worklist rows are 32-bit

  // ---- index list ordered by previous label (skipped while every label is still -1)
  const unsigned *order = nullptr;
  if (a.t_order && !a.t_labels_fresh && a.t_order_reuse) {
    order = a.t_order;   // a list from a few passes ago is still nearly label-ordered (and any order is correct)
  } else if (a.t_order && !a.t_labels_fresh) {
    e = cudaMemsetAsync(a.t_hist, 0, ((size_t)a.k + 1) * sizeof(unsigned), a.stream);
    if (e != cudaSuccess) return e;
    const int blocks = (int)std::min<int64_t>((a.n + 16383) / 16384, (int64_t)a.sm_count * 8);
    const int64_t per_block = (a.n + blocks - 1) / blocks;
    const size_t osm = (a.k + 1 <= KO_SMEM_BUCKETS) ? ((size_t)a.k + 1) * sizeof(unsigned) : 0;
    km_order_hist_kernel<<<blocks, KO_THREADS, osm, a.stream>>>(a.labels, a.n, a.k, per_block, a.t_hist);
    km_order_scan_kernel<<<1, 1024, 0, a.stream>>>(a.t_hist, a.k);
    km_order_scatter_kernel<<<blocks, KO_THREADS, osm, a.stream>>>(a.labels, a.n, a.k, per_block, a.t_hist, a.t_order);
    (*launches) += 3;
    order = a.t_order;
  }

  // ---- coarse stage
  kt_encode_tiled_t encode = nullptr;
  if ((e = kt_encoder(&encode)) != cudaSuccess) return e;
  CUtensorMap tmB;
  const cuuint64_t gdim[2] = {(cuuint64_t)drow, (cuuint64_t)kpad};
  const cuuint64_t gstride[1] = {(cuuint64_t)drow * sizeof(__nv_bfloat16)};
  const cuuint32_t box[2] = {64, KT_N};
  const cuuint32_t estr[2] = {1, 1};
  if (encode(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.t_cbf, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return cudaErrorInvalidValue;
  // the BF16 copy of x, row by row (box = 64 columns x 1 row; tile::gather4 takes four such rows per instruction)
  CUtensorMap tmA;
  const cuuint64_t adim[2] = {(cuuint64_t)drow, (cuuint64_t)a.n};
  const cuuint32_t abox[2] = {64, 1};
  if (encode(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(a.t_xbf), adim, gstride, abox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return cudaErrorInvalidValue;
  // the three BF16 pieces of h, a 16-byte row per centroid, no swizzle (foldx)
  CUtensorMap tmH;
  const cuuint64_t hdim[2] = {8, (cuuint64_t)kpad};
  const cuuint64_t hstride[1] = {16};
  const cuuint32_t hbox[2] = {8, KT_N};
  if (encode(&tmH, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, hfold, hdim, hstride, hbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return cudaErrorInvalidValue;
  e = cudaMemsetAsync(a.t_work_count, 0, sizeof(unsigned), a.stream);
  if (e != cudaSuccess) return e;
  KtParams p;
  p.x = (const float *)a.x; p.n = a.n; p.d = a.d; p.k = a.k; p.order = order; p.labels = a.labels; p.h = a.t_cn;
  p.xn = a.t_xn;
  p.st = a.state; p.xnmax = a.t_xnmax; p.work = a.t_work; p.work_count = a.t_work_count;
  p.all_to_work = a.incremental ? 0 : 1;
  size_t smem = kt_smem_bytes(a.d);
  p.foldx = foldx ? 1 : 0;
  if (foldx) smem += 2048 * (2 + kt_bstages(a.d));
  p.h_in_smem = !fold && !foldx && smem + (size_t)kpad * sizeof(float) + 1024 <= (size_t)a.max_smem_optin;
  if (p.h_in_smem) smem += (size_t)kpad * sizeof(float);
  auto kern = (fold || foldx) ? km_tensor_coarse_kernel<false, true>
                   : (p.h_in_smem ? km_tensor_coarse_kernel<true, false> : km_tensor_coarse_kernel<false, false>);
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int64_t passes = (a.n + KT_ROWS - 1) / KT_ROWS;
  kern<<<(int)(passes < a.sm_count ? passes : a.sm_count), KT_THREADS, smem, a.stream>>>(tmA, tmB, tmH, p);
  (*launches)++;
  if ((e = cudaGetLastError()) != cudaSuccess) return e;

  // ---- exact stage over the worklist (its length is read on the device)
  const int64_t want = (a.n * 8 + 255) / 256;
  const int64_t cap2 = (int64_t)a.sm_count * 8;
  const int eg = (int)(want < cap2 ? want : cap2);
#define KT_EXACT(QN)                                                                                                  \
  km_tensor_exact_kernel<QN><<<eg, 256, 0, a.stream>>>((const float *)a.x, a.d, a.k, a.t_work, a.t_work_count, a.chi, \
                                                       a.c64, a.labels, a.acc, a.state, a.incremental)
  if (a.d <= 32) KT_EXACT(1);
  else if (a.d <= 64) KT_EXACT(2);
  else KT_EXACT(4);
#undef KT_EXACT
  (*launches)++;
  return cudaGetLastError();
}
