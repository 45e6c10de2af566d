// km_assign_tensor.cu — KM_PATH_TENSOR: GEMM-form assignment on the 5th-gen tensor cores
// (tcgen05.mma, accumulators in TMEM) for large k*d (BASELINE.json configs[2], configs[3]).
//
// Two stages:
//
//  1. COARSE, tensor cores.  For a tile of 128 points the kernel forms
//         dist~(i,j) = |c_j|^2 - 2 x_i.c_j            (|x_i|^2 is the same for every j)
//     with x.c from `tcgen05.mma.kind::tf32` (operands straight from the float32 rows, staged
//     in shared memory in the canonical K-major 128-byte-swizzled layout; 128x128 fp32
//     accumulator tiles double-buffered in TMEM so the MMA of centroid tile t+1 runs under
//     the epilogue of tile t).  The N x k distance matrix never leaves the SM: the epilogue
//     (tcgen05.ld, one TMEM lane = one point per thread) keeps, per point, only the
//     centroids whose coarse distance is within 2E of the running minimum, where E is a
//     rigorous bound on the TF32 error.  Any centroid outside that set is provably not the
//     binary64 argmin.  At most KT_CAND survivors per point are written out (4 ints).
//
//  2. EXACT, CUDA cores (km_tensor_stage2_kernel, one warp per point).  The survivors —
//     almost always 1 to 3 — are re-evaluated with direct binary32 differences and the same
//     error-bound test as the small-d path, falling back to binary64 in the oracle's exact
//     operation order when the gap is inside the bound.  Label compare, order-free sums
//     (global int64 atomics), change count and inertia follow.
//
// So the labels are the oracle's bit for bit, and the tensor cores do all O(N k d) work.
#include <cuda_bf16.h>
#include <stdlib.h>

#include "km_commit.cuh"
#include "km_internal.h"

#define KT_M 128         // points per tile (UMMA M, cta_group::1)
#define KT_N 128         // centroids per MMA tile (UMMA N)
#define KT_THREADS 128   // 4 warps: warp w owns TMEM lanes [32w, 32w+32)
#define KT_LIST 8        // per-point survivor list kept in shared memory while the tiles stream by
#define KT_KBLK 32       // floats per 128-byte swizzled row

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// make generic-proxy writes to shared memory visible to the tensor core (async proxy)
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

__device__ __forceinline__ void tmem_alloc(uint32_t *slot_in_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, TF32 inputs, FP32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// issue only: the registers are valid after the next tmem_ld_wait()
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
// hits |= bit  iff  dist <= thr   (one compare + one predicated OR)
__device__ __forceinline__ void kt_hit(unsigned &hits, float dist, float thr, unsigned bit) {
  asm("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p or.b32 %0, %0, %3;\n\t}" : "+r"(hits) : "f"(dist), "f"(thr), "r"(bit));
}
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
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
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B: rows of 128 bytes, 8-row (1024 B) swizzle
// atoms stacked along M/N (stride byte offset 1024), version 1 (Blackwell).  Fields as in
// cute::UMMA::SmemDescriptor: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version [46,48),
// layout type [61,64) with SWIZZLE_128B = 2.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 = 1 at [4,6), a/b format
// TF32 = 2 at [7,10)/[10,13), both K-major, N>>3 at [17,23), M>>4 at [24,29).
__device__ __forceinline__ uint32_t umma_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// same, BF16 inputs (kind::f16), K = 16 per instruction
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// rare: the survivor list of a point is full -> drop what the tighter threshold has since excluded
__device__ __noinline__ int kt_compact(int2 *lst, int row, float thr, int stride = KT_M) {
  int w = 0;
  for (int s = 0; s < KT_LIST; s++) {
    const int2 e = lst[s * stride + row];
    if (__int_as_float(e.y) <= thr) lst[(w++) * stride + row] = e;
  }
  return w;
}

// ------------------------------------------------------------------ stage 1: coarse kernel
struct KmTensorParams {
  const float *x;      // [n][d]
  int64_t n;
  int d, k;
  const float *chi;    // [k][d] binary32 centroids
  const float *cn;     // [kpad] |chi_j|^2, +inf beyond k
  const KmState *st;   // st->cnmax2 = max_j |chi_j|^2
  float xnmax;         // upper bound of |x_i| over the job
  int *cand;           // [n][KT_CAND]: survivors (ascending index), -1 = unused; cand[i][0] = -2: scan all k
  const int32_t *labels;  // previous labels: the old centroid's distance seeds the running minimum
};

// Copy `rows` x d floats (row stride d) into the swizzled K-major layout the UMMA descriptor
// describes: k-block kb (32 floats) of row r lives at kb*(R*128) + r*128 + ((chunk ^ (r & 7)) * 16).
// Rows at or beyond `valid` are zero-filled (cp.async with src-size 0).
__device__ __forceinline__ void load_tile_sw128(uint32_t smem_base, const float *g, int64_t row0, int64_t valid_rows,
                                                int d) {
  // lanes sweep the 16-byte chunks of a row (d/4 of them: 8, 16, 24 or 32), several rows per
  // warp pass when a row is shorter than a warp; nothing in the loop divides
  const int cpr = d >> 2;
  const int rpp = 32 / cpr;                       // rows per warp pass (1 for d = 96, 128)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r_off = lane / cpr, q = lane - r_off * cpr;
  if (r_off >= rpp) return;                       // d = 96: lanes 24..31 idle
  const int kb = q >> 3, ch = q & 7;
  for (int r = warp * rpp + r_off; r < KT_M; r += (KT_THREADS / 32) * rpp) {
    const uint32_t dst = smem_base + kb * (KT_M * 128) + r * 128 + ((ch ^ (r & 7)) << 4);
    const bool ok = row0 + r < valid_rows;
    const float *src = g + (ok ? (row0 + r) * (int64_t)d + q * 4 : 0);
    cp_async16(dst, src, ok ? 16 : 0);
  }
}

// shared-memory bytes of the coarse kernel for dimension d
static __host__ __device__ inline size_t kt_smem_bytes(int d) {
  return 3 * (size_t)KT_M * d * 4          // A tile + two B tiles
         + 2 * KT_N * sizeof(float)        // centroid norms of the two B tiles
         + 36 * KT_M * sizeof(float)       // one chunk of coarse distances per point (row stride 36)
         + (size_t)KT_LIST * KT_M * 8      // survivor lists
         + 64;                             // mbarriers + TMEM slot
}

__global__ void __launch_bounds__(KT_THREADS) km_tensor_coarse_kernel(const KmTensorParams p) {
  // SWIZZLE_128B operand tiles need 1024-byte alignment
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int d = p.d, kblocks = d / KT_KBLK;
  const uint32_t tile_bytes = (uint32_t)KT_M * d * 4;            // one operand tile
  const uint32_t sA = smem_u32(smem_raw), sB0 = sA + tile_bytes, sB1 = sB0 + tile_bytes;
  float *cnb = reinterpret_cast<float *>(smem_raw + 3 * (size_t)tile_bytes);     // [2][KT_N] centroid norms of the tile
  float *dsm = cnb + 2 * KT_N;                                                   // [KT_M][36] distances of one chunk
  int2 *lst = reinterpret_cast<int2 *>(dsm + 36 * KT_M);                         // [KT_LIST][KT_M] (index, dist bits)
  uint64_t *bars = reinterpret_cast<uint64_t *>(lst + KT_LIST * KT_M);           // [2] MMA-done barriers
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2);

  const int warp = threadIdx.x >> 5, row = threadIdx.x;
  if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); }
  if (warp == 0) tmem_alloc(tmem_slot, 2 * KT_N);     // two 128-column fp32 accumulator tiles
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc = umma_idesc_tf32(KT_M, KT_N);
  const int ntiles = (p.k + KT_N - 1) / KT_N;
  // rigorous TF32 bound: |dist~ - dist| <= E = 2^-7.5 |x_i| max_j|c_j| (+ fp32 slop)
  const float cnmax = sqrtf(p.st->cnmax2);
  const float two_e = 2.0f * (0.0055243f * p.xnmax * cnmax) + 1e-6f * (p.xnmax * p.xnmax + p.st->cnmax2);
  uint32_t phase[2] = {0, 0};

  const int64_t mtiles = (p.n + KT_M - 1) / KT_M;
  // operands of the first point tile
  if ((int64_t)blockIdx.x < mtiles) {
    load_tile_sw128(sA, p.x, (int64_t)blockIdx.x * KT_M, p.n, d);
    load_tile_sw128(sB0, p.chi, 0, p.k, d);
    cp_async_commit();
  }
  for (int64_t mt = blockIdx.x; mt < mtiles; mt += gridDim.x) {
    const int64_t row0 = mt * KT_M;
    if (threadIdx.x < KT_N) cnb[threadIdx.x] = p.cn[threadIdx.x];
    cp_async_wait<0>();
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      for (int kb = 0; kb < kblocks; kb++)
        for (int ks = 0; ks < 4; ks++)
          umma_tf32(tmem_base, umma_desc_sw128(sA + kb * (KT_M * 128) + ks * 32),
                    umma_desc_sw128(sB0 + kb * (KT_N * 128) + ks * 32), idesc, (kb | ks) != 0);
      umma_commit(&bars[0]);
    }
    if (ntiles > 1) {
      load_tile_sw128(sB1, p.chi, KT_N, p.k, d);
      if (threadIdx.x < KT_N) cnb[KT_N + threadIdx.x] = p.cn[KT_N + threadIdx.x];
      cp_async_commit();
    }
    // Seed the running minimum with the coarse distance to the point's PREVIOUS centroid
    // (binary32 dot product of this thread's row of the A tile with chi[old], + E for the
    // TF32 stage's own error): the true minimum can only be lower, so nothing is lost, and the
    // threshold is tight from the first centroid tile on instead of after the nearest one
    // happens to stream by.
    float runmin = __int_as_float(0x7f800000), thr = runmin;
    {
      const int old = (row0 + row < p.n) ? p.labels[row0 + row] : -1;
      if (old >= 0 && old < p.k) {
        const float *cr = p.chi + (int64_t)old * d;
        float dot = 0.f;
        for (int kb = 0; kb < kblocks; kb++) {
          const unsigned char *arow = smem_raw + kb * (KT_M * 128) + row * 128;
#pragma unroll
          for (int ch = 0; ch < 8; ch++) {
            const float4 xv = *reinterpret_cast<const float4 *>(arow + ((ch ^ (row & 7)) << 4));
            const float4 cv = __ldg(reinterpret_cast<const float4 *>(cr + kb * 32 + ch * 4));
            dot = fmaf(xv.x, cv.x, dot); dot = fmaf(xv.y, cv.y, dot);
            dot = fmaf(xv.z, cv.z, dot); dot = fmaf(xv.w, cv.w, dot);
          }
        }
        runmin = fmaf(-2.0f, dot, p.cn[old]) + 0.5f * two_e;
        thr = runmin + two_e;
      }
    }
    int cnt = 0, overflow = 0;

    for (int t = 0; t < ntiles; t++) {
      const int b = t & 1;
      if (t + 1 < ntiles) {
        // B(t+1) has landed -> issue MMA(t+1) into the other TMEM buffer; it runs under epilogue(t)
        cp_async_wait<0>();
        fence_proxy_async();
        __syncthreads();
        if (threadIdx.x == 0) {
          tc_fence_after();
          const uint32_t sB = b ? sB0 : sB1;
          for (int kb = 0; kb < kblocks; kb++)
            for (int ks = 0; ks < 4; ks++)
              umma_tf32(tmem_base + (uint32_t)((b ^ 1) * KT_N), umma_desc_sw128(sA + kb * (KT_M * 128) + ks * 32),
                        umma_desc_sw128(sB + kb * (KT_N * 128) + ks * 32), idesc, (kb | ks) != 0);
          umma_commit(&bars[b ^ 1]);
        }
      }
      mbar_wait(&bars[b], phase[b]);     // MMA(t) done: TMEM buffer b is readable, B smem buffer b is free
      phase[b] ^= 1;
      tc_fence_after();
      if (t == ntiles - 1 && mt + gridDim.x < mtiles) {
        // every MMA of this point tile has completed: the operand buffers are free, so the
        // next point tile's A and first B tile stream in under this last epilogue
        load_tile_sw128(sA, p.x, (mt + gridDim.x) * KT_M, p.n, d);
        load_tile_sw128(sB0, p.chi, 0, p.k, d);
        cp_async_commit();
      }
      // epilogue(t): this thread's point against the 128 centroids of tile t.  The hot loop is
      // branch-free (distance, store, compare, set a bit); hits are handled after the chunk.
      const float *cnt_tile = cnb + b * KT_N;
#pragma unroll 1
      for (int c = 0; c < KT_N; c += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(b * KT_N + c), v);
        unsigned hits = 0;
        float *drow = dsm + row * 36;
#pragma unroll
        for (int i4 = 0; i4 < 8; i4++) {
          const float4 cn4 = *reinterpret_cast<const float4 *>(cnt_tile + c + i4 * 4);   // warp-uniform: one broadcast LDS.128
          float4 d4;
          d4.x = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 0]), cn4.x);
          d4.y = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 1]), cn4.y);
          d4.z = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 2]), cn4.z);
          d4.w = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 3]), cn4.w);
          *reinterpret_cast<float4 *>(drow + i4 * 4) = d4;
          if (d4.x <= thr) hits |= 1u << (i4 * 4 + 0);
          if (d4.y <= thr) hits |= 1u << (i4 * 4 + 1);
          if (d4.z <= thr) hits |= 1u << (i4 * 4 + 2);
          if (d4.w <= thr) hits |= 1u << (i4 * 4 + 3);
        }
        while (hits) {
          const int i = __ffs(hits) - 1;
          hits &= hits - 1;
          const float dist = drow[i];
          if (dist <= thr) {             // the threshold may have tightened inside this chunk
            if (cnt == KT_LIST) cnt = kt_compact(lst, row, thr);
            if (cnt < KT_LIST) lst[(cnt++) * KT_M + row] = make_int2(t * KT_N + c + i, __float_as_int(dist));
            else overflow = 1;
            if (dist < runmin) { runmin = dist; thr = runmin + two_e; }
          }
        }
      }
      tc_fence_before();
      __syncthreads();                   // every warp is done with TMEM buffer b and with cnb[b]
      if (t + 2 < ntiles) {
        load_tile_sw128(b ? sB1 : sB0, p.chi, (int64_t)(t + 2) * KT_N, p.k, d);
        if (threadIdx.x < KT_N) cnb[b * KT_N + threadIdx.x] = p.cn[(t + 2) * KT_N + threadIdx.x];
        cp_async_commit();
      }
    }
    // ---- survivors of this point: within 2E of the final minimum, ascending index
    if (row0 + row < p.n) {
      int out[KT_CAND];
      int w = 0;
      for (int s2 = 0; s2 < cnt; s2++) {
        const int2 e = lst[s2 * KT_M + row];
        if (__int_as_float(e.y) <= thr) {
          if (w < KT_CAND) out[w] = e.x;
          w++;
        }
      }
      if (overflow || w > KT_CAND || w == 0) {
        out[0] = -2;                       // too many to list: stage 2 scans all k
        for (int s2 = 1; s2 < KT_CAND; s2++) out[s2] = -1;
      } else {
        for (int s2 = w; s2 < KT_CAND; s2++) out[s2] = -1;
        for (int a = 1; a < w; a++)        // insertion sort, ascending index (tie rule: lowest j first)
          for (int q = a; q > 0 && out[q] < out[q - 1]; q--) { const int tmp = out[q]; out[q] = out[q - 1]; out[q - 1] = tmp; }
      }
      int4 *cdst = reinterpret_cast<int4 *>(p.cand + (row0 + row) * KT_CAND);
      cdst[0] = make_int4(out[0], out[1], out[2], out[3]);
      cdst[1] = make_int4(out[4], out[5], out[6], out[7]);
    }
    __syncthreads();                       // lst is reused by the next point tile
  }
  cp_async_wait<0>();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 2 * KT_N);
}

// ------------------------------------------------------------------ stage 1, BF16, two point tiles per CTA
// The production coarse kernel.  Differences from the TF32 kernel above:
//   * operands are BF16 (x converted on the way into shared memory, centroids pre-converted
//     once per pass): half the shared memory per tile and twice the MMA rate;
//   * a CTA holds TWO point tiles (256 points, 8 warps, all 512 TMEM columns as 2 x 2
//     accumulator buffers), so every centroid tile streamed from L2 feeds two MMAs — half the
//     L2 traffic per point — and the SM has 8 epilogue warps instead of 4.
// E grows to 2^-6.5 |x||c| (BF16 keeps 8 significant bits, round to nearest); everything
// downstream (survivor lists, stage 2) is unchanged.
#define KB_THREADS 256
#define KB_MT 2

static __host__ __device__ inline int kb_kblocks(int d) { return (d + 63) / 64; }
static __host__ __device__ inline size_t kb_smem_bytes(int d) {
  return (size_t)(KB_MT + 2) * kb_kblocks(d) * (KT_M * 128)   // two A tiles + two B buffers
         + 2 * KT_N * sizeof(float) + 36 * (KB_MT * KT_M) * sizeof(float) + (size_t)KT_LIST * (KB_MT * KT_M) * 8 + 64;
}

struct KmTensorBf16Params {
  KmTensorParams base;
  const __nv_bfloat16 *cbf;   // [kpad][dpad] BF16 centroids, zero padded
  int dpad;                   // kblocks * 64
};

__global__ void __launch_bounds__(KB_THREADS, 1) km_tensor_coarse_bf16_kernel(const KmTensorBf16Params pp) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const KmTensorParams &p = pp.base;
  const int d = p.d, kblocks = kb_kblocks(d), dpad = pp.dpad;
  constexpr int ROWS = KB_MT * KT_M;                       // 256 points per CTA pass
  const uint32_t kb_bytes = KT_M * 128;                    // one 64-element k-block of one 128-row tile
  const uint32_t tile_bytes = kblocks * kb_bytes;
  const uint32_t sA = smem_u32(smem_raw), sB0 = sA + KB_MT * tile_bytes, sB1 = sB0 + tile_bytes;
  float *cnb = reinterpret_cast<float *>(smem_raw + (KB_MT + 2) * (size_t)tile_bytes);
  float *dsm = cnb + 2 * KT_N;                                                   // [ROWS][36]
  int2 *lst = reinterpret_cast<int2 *>(dsm + 36 * ROWS);                         // [KT_LIST][ROWS]
  uint64_t *bars = reinterpret_cast<uint64_t *>(lst + KT_LIST * ROWS);
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, row = threadIdx.x;
  const int mi = warp >> 2, quad = warp & 3;               // point tile and TMEM lane quadrant of this warp
  // zero the A region once: the padding columns (d..dpad) are never written again
  for (uint32_t o = threadIdx.x * 16; o < KB_MT * tile_bytes; o += KB_THREADS * 16)
    *reinterpret_cast<uint4 *>(smem_raw + o) = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); }
  if (warp == 0) tmem_alloc(tmem_slot, 512);              // 2 point tiles x 2 buffers x 128 columns
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc = umma_idesc_bf16(KT_M, KT_N);
  const int ntiles = (p.k + KT_N - 1) / KT_N;
  const float cnmax = sqrtf(p.st->cnmax2);
  // BF16 keeps 8 significant bits: unit roundoff u = 2^-8 per operand, so a product errs by
  // (2u + u^2)|x_t c_t| and dist~ = |c|^2 - 2 x.c by E = (2^-6 + 2^-15) sum|x_t c_t| <= (2^-6 + 2^-15)|x||c|
  // (DESIGN.md §6.7); survivors are kept within 2E of the running minimum.
  const float two_e = 2.0f * (0.0156556f * p.xnmax * cnmax) + 1e-6f * (p.xnmax * p.xnmax + p.st->cnmax2);
  uint32_t phase[2] = {0, 0};

  // x rows -> BF16, swizzled K-major (two tiles); lanes sweep float4 chunks, warps sweep rows
  auto load_a = [&](int64_t row0) {
    const int cpr = d >> 2, rpp = 32 / cpr, r_off = lane / cpr, q = lane - r_off * cpr;
    if (r_off >= rpp) return;
    const int e = q * 4, kb = e >> 6, eb = e & 63, ch = eb >> 3, within = (eb & 7) * 2;
    const int rstep = (KB_THREADS / 32) * rpp;
    // batches of 8 rows: all eight global loads are in flight before the first conversion
    for (int r0 = warp * rpp + r_off; r0 < ROWS; r0 += 8 * rstep) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int r = r0 + u * rstep;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < ROWS && row0 + r < p.n) v[u] = __ldg(reinterpret_cast<const float4 *>(p.x + (row0 + r) * (int64_t)d + e));
      }
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int r = r0 + u * rstep;
        if (r >= ROWS) break;
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v[u].x, v[u].y), hi = __floats2bfloat162_rn(v[u].z, v[u].w);
        const int rr = r & (KT_M - 1);
        unsigned char *dst = smem_raw + (r >> 7) * tile_bytes + kb * kb_bytes + rr * 128 + ((ch ^ (rr & 7)) << 4) + within;
        *reinterpret_cast<uint2 *>(dst) = make_uint2(*reinterpret_cast<const unsigned *>(&lo), *reinterpret_cast<const unsigned *>(&hi));
      }
    }
  };
  auto load_b = [&](uint32_t sB, int64_t j0) {
    const int cpr = dpad >> 3, rpp = 32 / cpr, r_off = lane / cpr, q = lane - r_off * cpr;   // 16-byte chunks of 8 BF16
    const int kb = q >> 3, ch = q & 7;
    for (int r = warp * rpp + r_off; r < KT_N; r += (KB_THREADS / 32) * rpp)
      cp_async16(sB + kb * kb_bytes + r * 128 + ((ch ^ (r & 7)) << 4), pp.cbf + (j0 + r) * (int64_t)dpad + q * 8, 16);
  };
  auto issue_mma = [&](uint32_t sB, int b) {
    for (int m = 0; m < KB_MT; m++)
      for (int kb = 0; kb < kblocks; kb++)
        for (int ks = 0; ks < 4; ks++)
          umma_bf16(tmem_base + (uint32_t)((m * 2 + b) * KT_N), umma_desc_sw128(sA + m * tile_bytes + kb * kb_bytes + ks * 32),
                    umma_desc_sw128(sB + kb * kb_bytes + ks * 32), idesc, (kb | ks) != 0);
    umma_commit(&bars[b]);
  };

  const int64_t passes = (p.n + ROWS - 1) / ROWS;
  if ((int64_t)blockIdx.x < passes) {
    load_a((int64_t)blockIdx.x * ROWS);
    load_b(sB0, 0);
    cp_async_commit();
  }
  for (int64_t ps = blockIdx.x; ps < passes; ps += gridDim.x) {
    const int64_t row0 = ps * ROWS;
    if (threadIdx.x < KT_N) cnb[threadIdx.x] = p.cn[threadIdx.x];
    cp_async_wait<0>();
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x == 0) { tc_fence_after(); issue_mma(sB0, 0); }
    if (ntiles > 1) {
      load_b(sB1, KT_N);
      if (threadIdx.x < KT_N) cnb[KT_N + threadIdx.x] = p.cn[KT_N + threadIdx.x];
      cp_async_commit();
    }
    // seed the running minimum with the coarse distance to the previous centroid (see above)
    float runmin = __int_as_float(0x7f800000), thr = runmin;
    {
      const int old = (row0 + row < p.n) ? p.labels[row0 + row] : -1;
      if (old >= 0 && old < p.k) {
        const __nv_bfloat16 *cr = pp.cbf + (int64_t)old * dpad;
        const int rr = row & (KT_M - 1);
        float dot = 0.f;
        for (int kb = 0; kb < kblocks; kb++) {
          const unsigned char *arow = smem_raw + mi * tile_bytes + kb * kb_bytes + rr * 128;
#pragma unroll
          for (int ch = 0; ch < 8; ch++) {
            const uint4 xa = *reinterpret_cast<const uint4 *>(arow + ((ch ^ (rr & 7)) << 4));
            const uint4 cb = __ldg(reinterpret_cast<const uint4 *>(cr + kb * 64 + ch * 8));
            const unsigned xw[4] = {xa.x, xa.y, xa.z, xa.w}, cw[4] = {cb.x, cb.y, cb.z, cb.w};
#pragma unroll
            for (int u = 0; u < 4; u++) {
              const float2 xf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&xw[u]));
              const float2 cf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&cw[u]));
              dot = fmaf(xf.x, cf.x, dot);
              dot = fmaf(xf.y, cf.y, dot);
            }
          }
        }
        runmin = fmaf(-2.0f, dot, p.cn[old]) + 0.5f * two_e;
        thr = runmin + two_e;
      }
    }
    int cnt = 0, overflow = 0;

    for (int t = 0; t < ntiles; t++) {
      const int b = t & 1;
      if (t + 1 < ntiles) {
        cp_async_wait<0>();
        fence_proxy_async();
        __syncthreads();
        if (threadIdx.x == 0) { tc_fence_after(); issue_mma(b ? sB0 : sB1, b ^ 1); }
      }
      mbar_wait(&bars[b], phase[b]);
      phase[b] ^= 1;
      tc_fence_after();
      if (t == ntiles - 1 && ps + gridDim.x < passes) {
        // all MMAs of this pass are done: the next pass's operands stream in under this epilogue
        load_a((ps + gridDim.x) * ROWS);
        load_b(sB0, 0);
        cp_async_commit();
      } else if (t + 2 < ntiles) {
        // B buffer b is free (MMA(t) has completed): tile t+2 streams in under epilogue(t)
        load_b(b ? sB1 : sB0, (int64_t)(t + 2) * KT_N);
        cp_async_commit();
      }
      const float *cnt_tile = cnb + b * KT_N;
      float *drow = dsm + row * 36;
      const uint32_t tcol = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)((mi * 2 + b) * KT_N);
      // one chunk of 32 columns: distances to shared memory, hit mask, then the rare hits
      auto chunk = [&](const uint32_t (&v)[32], int c) {
        unsigned hits = 0;
#pragma unroll
        for (int i4 = 0; i4 < 8; i4++) {
          const float4 cn4 = *reinterpret_cast<const float4 *>(cnt_tile + c + i4 * 4);
          float4 d4;
          d4.x = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 0]), cn4.x);
          d4.y = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 1]), cn4.y);
          d4.z = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 2]), cn4.z);
          d4.w = fmaf(-2.0f, __uint_as_float(v[i4 * 4 + 3]), cn4.w);
          *reinterpret_cast<float4 *>(drow + i4 * 4) = d4;
          kt_hit(hits, d4.x, thr, 1u << (i4 * 4 + 0));
          kt_hit(hits, d4.y, thr, 1u << (i4 * 4 + 1));
          kt_hit(hits, d4.z, thr, 1u << (i4 * 4 + 2));
          kt_hit(hits, d4.w, thr, 1u << (i4 * 4 + 3));
        }
        while (hits) {
          const int i = __ffs(hits) - 1;
          hits &= hits - 1;
          const float dist = drow[i];
          if (dist <= thr) {
            if (cnt == KT_LIST) cnt = kt_compact(lst, row, thr, ROWS);
            if (cnt < KT_LIST) lst[(cnt++) * ROWS + row] = make_int2(t * KT_N + c + i, __float_as_int(dist));
            else overflow = 1;
            if (dist < runmin) { runmin = dist; thr = runmin + two_e; }
          }
        }
      };
      {
        // the TMEM read of chunk c+1 is in flight while chunk c is processed
        uint32_t va[32], vb[32];
        tmem_ld32_issue(tcol + 0, va);
        tmem_ld_wait();
        tmem_ld32_issue(tcol + 32, vb);
        chunk(va, 0);
        tmem_ld_wait();
        tmem_ld32_issue(tcol + 64, va);
        chunk(vb, 32);
        tmem_ld_wait();
        tmem_ld32_issue(tcol + 96, vb);
        chunk(va, 64);
        tmem_ld_wait();
        chunk(vb, 96);
      }
      tc_fence_before();
      __syncthreads();
      if (t + 2 < ntiles && threadIdx.x < KT_N) cnb[b * KT_N + threadIdx.x] = p.cn[(t + 2) * KT_N + threadIdx.x];
    }
    if (row0 + row < p.n) {
      int out[KT_CAND];
      int w = 0;
      for (int s2 = 0; s2 < cnt; s2++) {
        const int2 e = lst[s2 * ROWS + row];
        if (__int_as_float(e.y) <= thr) {
          if (w < KT_CAND) out[w] = e.x;
          w++;
        }
      }
      if (overflow || w > KT_CAND || w == 0) {
        out[0] = -2;
        for (int s2 = 1; s2 < KT_CAND; s2++) out[s2] = -1;
      } else {
        for (int s2 = w; s2 < KT_CAND; s2++) out[s2] = -1;
        for (int a = 1; a < w; a++)
          for (int q = a; q > 0 && out[q] < out[q - 1]; q--) { const int tmp = out[q]; out[q] = out[q - 1]; out[q - 1] = tmp; }
      }
      int4 *cdst = reinterpret_cast<int4 *>(p.cand + (row0 + row) * KT_CAND);
      cdst[0] = make_int4(out[0], out[1], out[2], out[3]);
      cdst[1] = make_int4(out[4], out[5], out[6], out[7]);
    }
    __syncthreads();
  }
  cp_async_wait<0>();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

// |chi_j|^2 per centroid (padded with +inf to a whole number of MMA tiles) and its maximum
__global__ void __launch_bounds__(256) km_tensor_cnorm_kernel(const float *__restrict__ chi, int k, int d, int kpad,
                                                              float *__restrict__ cn, KmState *st,
                                                              __nv_bfloat16 *__restrict__ cbf, int dpad) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= kpad) return;
  float s = 0.f;
  if (warp < k)
    for (int t = lane; t < d; t += 32) { const float v = chi[(int64_t)warp * d + t]; s = fmaf(v, v, s); }
  if (cbf)   // BF16 copy of the row, zero beyond d and beyond k
    for (int t = lane; t < dpad; t += 32)
      cbf[(int64_t)warp * dpad + t] = __float2bfloat16_rn((warp < k && t < d) ? chi[(int64_t)warp * d + t] : 0.f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    cn[warp] = warp < k ? s : __int_as_float(0x7f800000);
    if (warp < k) atomicMax(reinterpret_cast<unsigned int *>(&st->cnmax2), __float_as_uint(s));   // s >= 0: bits order = value order
  }
}

// ------------------------------------------------------------------ stage 2: exact kernel
// Eight lanes per point (four points per warp); each lane owns d/32 float4 slices of the row.
// Survivors are re-evaluated with direct binary32 differences; the binary64 re-score (oracle
// operation order, one lane, sequential) settles what the error bound cannot.
__device__ __forceinline__ float kt_sum8(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}
__device__ __forceinline__ double kt_sum8(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}

__global__ void __launch_bounds__(256)
km_tensor_stage2_kernel(const float *__restrict__ x, int64_t n, int d, int k, const int *__restrict__ cand,
                        const float *__restrict__ chi, const double *__restrict__ c64, int32_t *__restrict__ labels,
                        long long *__restrict__ acc, const KmState *__restrict__ st, int incremental, int exact_inertia) {
  const KmAccLayout L = km_acc_layout(k, d);
  const double scale_a = st->scale_a, scale_b = st->scale_b;
  const float tau2 = st->tau_hi;
  const float rho2 = 2.5f * (float)(d + 6) * 0x1.0p-24f;
  const int lane = threadIdx.x & 31, g = lane >> 3, sl = lane & 7;
  const int nq = d >> 5;                                   // float4 slices per lane (1..4)
  const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  KmThreadScalars S;
  for (int64_t base = warp0 * 4; base < n; base += nwarps * 4) {
    const int64_t i = base + g;
    const bool live = i < n;
    const float *xr = x + (live ? i : 0) * (int64_t)d;
    int4 cw = make_int4(-1, -1, -1, -1), cw2 = make_int4(-1, -1, -1, -1);
    int old = 0;
    float4 xq[4];
    if (live) {
      cw = *reinterpret_cast<const int4 *>(cand + i * KT_CAND);
      cw2 = *reinterpret_cast<const int4 *>(cand + i * KT_CAND + 4);
      old = labels[i];
#pragma unroll
      for (int q = 0; q < 4; q++)
        if (q < nq) xq[q] = *reinterpret_cast<const float4 *>(xr + (q * 8 + sl) * 4);
    }
    const bool scan_all = cw.x == -2;
    const int ncand = !live ? 0 : scan_all ? k : ((cw.y < 0) ? 1 : (cw.z < 0) ? 2 : (cw.w < 0) ? 3 : (cw2.x < 0) ? 4 :
                                                  (cw2.y < 0) ? 5 : (cw2.z < 0) ? 6 : (cw2.w < 0) ? 7 : 8);
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
        const float *cr = chi + (int64_t)j * d;
#pragma unroll
        for (int q = 0; q < 4; q++)
          if (q < nq) {
            const float4 cv = *reinterpret_cast<const float4 *>(cr + (q * 8 + sl) * 4);
            float df = xq[q].x - cv.x; part = fmaf(df, df, part);
            df = xq[q].y - cv.y; part = fmaf(df, df, part);
            df = xq[q].z - cv.z; part = fmaf(df, df, part);
            df = xq[q].w - cv.w; part = fmaf(df, df, part);
          }
      }
      part = kt_sum8(part);
      if (act) { if (part < b1) { b2 = b1; b1 = part; i1 = j; } else if (part < b2) b2 = part; }
    }
    int lab = i1;
    const bool clear = ncand <= 1 || (b2 - b1) > fmaf(rho2, b2, tau2);
    double best = 0.0;
    bool have_best = false;
    if (live && !clear) {
      // binary64, dims ascending, separate multiply and add: the oracle's own arithmetic (A1/A2)
      if (sl == 0) {
        best = __longlong_as_double(0x7FF0000000000000LL);
        lab = 0;
        for (int s = 0; s < ncand; s++) {
          const int j = scan_all ? s : pick(s);
          const double dj = km_dist2_f64(xr, c64 + (int64_t)j * d, d);
          if (dj < best) { best = dj; lab = j; }
        }
        S.rescored++;
      }
      have_best = true;
    }
    lab = __shfl_sync(0xffffffffu, lab, g * 8);     // the group's lane 0 holds the decision
    if (live && sl == 0 && lab != old) { S.changed++; labels[i] = lab; }
    const bool move = live && (incremental ? (lab != old) : true);
    if (move) {
      // consecutive lanes take consecutive dimensions: 8 lanes = 64 contiguous bytes per limb row
      const int from = incremental ? old : -1;
      for (int t = sl; t < d; t += 8) {
        long long a, b;
        km_fixed_split((double)xr[t], scale_a, scale_b, a, b);
        km_atomic_add_ll(acc + L.sa + (int64_t)lab * d + t, a);
        km_atomic_add_ll(acc + L.sb + (int64_t)lab * d + t, b);
        if (from >= 0) {
          km_atomic_add_ll(acc + L.sa + (int64_t)from * d + t, -a);
          km_atomic_add_ll(acc + L.sb + (int64_t)from * d + t, -b);
        }
      }
      if (sl == 0) {
        km_atomic_add_ll(acc + L.cnt + lab, 1);
        if (from >= 0) km_atomic_add_ll(acc + L.cnt + from, -1);
      }
    }
  }
  km_flush_scalars(S, acc, L);
}

// ------------------------------------------------------------------ launcher

bool km_tensor_path_supported(int32_t d, int32_t k, int dtype, int max_smem_optin) {
  if (dtype != KM_DTYPE_F32 || d % KT_KBLK != 0 || d < KT_KBLK || d > 128 || k < 2) return false;
  return kt_smem_bytes(d) <= (size_t)max_smem_optin && kb_smem_bytes(d) <= (size_t)max_smem_optin;
}

size_t km_tensor_cbf_bytes(int32_t d, int32_t k) {
  const int kpad = ((k + KT_N - 1) / KT_N + 2) * KT_N;
  return (size_t)kpad * kb_kblocks(d) * 64 * sizeof(__nv_bfloat16);
}

cudaError_t km_launch_assign_tensor(const KmAssignArgs &a, int64_t *launches) {
  const int kpad = ((a.k + KT_N - 1) / KT_N + 2) * KT_N;   // whole tiles, plus slack for the prefetch of cn
  static const bool use_tf32 = getenv("KM_TENSOR_TF32") != nullptr;   // A/B switch for measurements
  const int dpad = kb_kblocks(a.d) * 64;
  cudaError_t e = cudaMemsetAsync(&a.state->cnmax2, 0, sizeof(float), a.stream);
  if (e != cudaSuccess) return e;
  km_tensor_cnorm_kernel<<<(kpad * 32 + 255) / 256, 256, 0, a.stream>>>(a.chi, a.k, a.d, kpad, a.t_cn, a.state,
                                                                        use_tf32 ? nullptr : (__nv_bfloat16 *)a.t_cbf, dpad);
  (*launches)++;
  if (a.n > 0) {
    KmTensorParams p;
    p.x = (const float *)a.x; p.n = a.n; p.d = a.d; p.k = a.k; p.chi = a.chi; p.cn = a.t_cn; p.st = a.state;
    p.xnmax = a.t_xnmax; p.cand = a.t_cand; p.labels = a.labels;
    if (use_tf32) {
      const size_t smem = kt_smem_bytes(a.d);
      e = cudaFuncSetAttribute(km_tensor_coarse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      // two CTAs per SM when shared memory allows (each holds 256 of the SM's 512 TMEM columns)
      const int per_sm = (2 * (smem + 1024) <= (size_t)a.max_smem_optin) ? 2 : 1;
      const int64_t mtiles = (a.n + KT_M - 1) / KT_M;
      const int64_t cap = (int64_t)a.sm_count * per_sm;
      km_tensor_coarse_kernel<<<(int)(mtiles < cap ? mtiles : cap), KT_THREADS, smem, a.stream>>>(p);
    } else {
      KmTensorBf16Params pp;
      pp.base = p; pp.cbf = (const __nv_bfloat16 *)a.t_cbf; pp.dpad = dpad;
      const size_t smem = kb_smem_bytes(a.d);
      e = cudaFuncSetAttribute(km_tensor_coarse_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      const int64_t passes = (a.n + KB_MT * KT_M - 1) / (KB_MT * KT_M);
      km_tensor_coarse_bf16_kernel<<<(int)(passes < a.sm_count ? passes : a.sm_count), KB_THREADS, smem, a.stream>>>(pp);
    }
    (*launches)++;
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const int64_t want = (a.n * 8 + 255) / 256;
    const int64_t cap2 = (int64_t)a.sm_count * 8;
    km_tensor_stage2_kernel<<<(int)(want < cap2 ? want : cap2), 256, 0, a.stream>>>(
        (const float *)a.x, a.n, a.d, a.k, a.t_cand, a.chi, a.c64, a.labels, a.acc, a.state, a.incremental, a.exact_inertia);
    (*launches)++;
  }
  return cudaGetLastError();
}
