// km_noalloc.cu — km_lloyd_step_noalloc(): one Lloyd iteration that allocates nothing.
//
// Who this is for.  Some hosts hand a library a fixed set of device buffers and then forbid
// it to touch the allocator at all: a serving process whose whole device memory belongs to
// its own pool, a control loop that cannot absorb the device-wide stall a cudaMalloc or
// cudaFree may cause, a plugin sandbox that is billed for — and capped at — the footprint it
// declared at start-up.  The context API (km_create / km_upload_points / ...) is not usable
// there, because it sizes and allocates its tables on demand.  This entry point is: the
// caller passes every input and output as a device pointer, and the only memory the library
// itself needs — the order-free sum table and a few scalars — lives in ONE fixed array
// compiled into the library image.
//
// What it costs.  That array (KM_NOALLOC_SCRATCH_BYTES, a little over 256 KiB) is part of the
// module: every GPU that loads libkmeans.so holds it from then until the process exits,
// whether or not this function is ever called.  Problems whose table does not fit
// (k*(2d+1)+4 > KM_NOALLOC_MAX_ENTRIES) are refused, not spilled.
//
// Sharing.  Every call on a device uses the same array, and the array still holds whatever
// the previous call left there.  The function therefore clears the part that must start at
// zero as its first enqueued operation.  Two calls whose GPU work can overlap (different
// streams on one device, or two host threads) would scribble over each other's sums: run
// them one after the other.
//
// Set-up.  Host code cannot use the array's name as an address — on the host that name is
// only a handle for the runtime.  The first call on each device (with that device current)
// asks the runtime for the real device address with cudaGetSymbolAddress and keeps it; later
// calls reuse it.  Make that first call before starting any stream capture on the device.
// That first call is also the one place where the CUDA runtime itself may still allocate on
// the caller's behalf: it may create the device's primary context, and — module loading being
// lazy in CUDA 12 — it is when this file's kernels and the array are actually placed in device
// memory.  A host with a no-allocation phase therefore makes one warm-up call per device at
// start-up (any tiny problem will do) and enters that phase afterwards; from the second call
// on nothing is allocated by anyone.  The caller's current device is left as it was found.
#include <math.h>
#include <mutex>
#include <string>

#include "km_commit.cuh"
#include "km_finalize.cuh"
#include "km_internal.h"

#define KM_NOALLOC_MAX_ENTRIES 32768                 // int64 entries of the sum table
#define KM_NOALLOC_HEADER_BYTES 512
#define KM_NOALLOC_SCRATCH_BYTES (KM_NOALLOC_HEADER_BYTES + KM_NOALLOC_MAX_ENTRIES * 8)
#define KM_NOALLOC_MAX_DEVICES 64
#define KM_NA_THREADS 128

// The one piece of library-owned device memory this entry point uses.
static __device__ __align__(256) unsigned char km_noalloc_scratch[KM_NOALLOC_SCRATCH_BYTES];

struct KmNaHeader {
  unsigned long long maxabs_bits;   // max |x| of this call's points (binary64 bits)
  int nonfinite;                    // set if a point is NaN / Inf
  int pad;
  KmState st;                       // fixed-point scales of this call
};
static_assert(sizeof(KmNaHeader) <= KM_NOALLOC_HEADER_BYTES, "header must fit its slot");

void km_internal_set_thread_error(const char *msg);   // km_api.cu: text behind km_last_error(NULL)

// ---- set-up: the array's device address, fetched once per device ------------------------
static std::mutex g_na_mutex;
static void *g_na_ptr[KM_NOALLOC_MAX_DEVICES];

static cudaError_t na_scratch_address(int device, void **out) {
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[device]) {
    void *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&ptr, km_noalloc_scratch);   // `device` is current here
    if (e != cudaSuccess) return e;
    g_na_ptr[device] = ptr;
  }
  *out = g_na_ptr[device];
  return cudaSuccess;
}

// ---- kernels (variants of the scan, direct-assignment and finalize kernels that take the
//      scratch address instead of context-owned tables) ------------------------------------
__global__ void __launch_bounds__(256) km_na_scan_kernel(const float *__restrict__ x, int64_t count, KmNaHeader *hdr) {
  double m = 0.0;
  int bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    double v = fabs((double)x[i]);
    if (!(v <= 1.7976931348623157e308)) bad = 1; else m = fmax(m, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    bad |= __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (m > 0.0) atomicMax(&hdr->maxabs_bits, (unsigned long long)__double_as_longlong(m));
    if (bad) atomicOr(&hdr->nonfinite, 1);
  }
}

// one block: fixed-point parameters from (max|x|, n), filter/inertia scales from the centroids
__global__ void __launch_bounds__(256) km_na_setup_kernel(KmNaHeader *hdr, const double *__restrict__ c, int32_t k,
                                                          int32_t d, int64_t n) {
  __shared__ double red[8];
  double lmax = 0.0;
  for (int64_t e = threadIdx.x; e < (int64_t)k * d; e += blockDim.x) {
    const double v = fabs(c[e]);
    if (!(v <= 1.7976931348623157e308)) atomicOr(&hdr->nonfinite, 1);   // NaN / Inf among the centroids (fmax would drop a NaN)
    else lmax = fmax(lmax, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lmax = fmax(lmax, __shfl_xor_sync(0xffffffffu, lmax, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = lmax;
  __syncthreads();
  if (threadIdx.x != 0) return;
  double cmax = 0.0;
  for (int w = 0; w < 8; w++) cmax = fmax(cmax, red[w]);
  const double maxabs = __longlong_as_double((long long)hdr->maxabs_bits);
  int E = maxabs > 0.0 ? ilogb(maxabs) + 1 : 0;
  int NB = 1;
  while (NB < 62 && ((long long)1 << NB) <= n) NB++;
  KmState &st = hdr->st;
  st.E = E; st.QB = 62 - NB; st.qa = st.QB - E;
  st.scale_a = scalbn(1.0, st.qa);
  st.scale_b = scalbn(1.0, st.QB);
  km_refresh_scales(&st, cmax, maxabs, d);
}

// assignment in binary64 with the oracle's operation order (A1/A2), every point added to the
// scratch table from zero (the "full" update), change count and inertia fused
__global__ void __launch_bounds__(KM_NA_THREADS)
km_na_assign_kernel(const float *__restrict__ x, int64_t n, int32_t d, int32_t k, int32_t *__restrict__ labels,
                    const double *__restrict__ c64, long long *__restrict__ table, const KmNaHeader *__restrict__ hdr,
                    int chunk_rows, int tab_in_smem) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *cs = reinterpret_cast<double *>(smem_raw);
  long long *stab = reinterpret_cast<long long *>(cs + (size_t)chunk_rows * d);
  const KmAccLayout L = km_acc_layout(k, d);
  KmBlockTab T;
  T.L = L;
  T.tab = tab_in_smem ? stab : table;
  if (tab_in_smem) km_tab_zero(stab, L.scal);
  const double scale_a = hdr->st.scale_a, scale_b = hdr->st.scale_b, inertia_scale = hdr->st.inertia_scale;
  KmThreadScalars S;
  const int64_t ntiles = (n + KM_NA_THREADS - 1) / KM_NA_THREADS;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t i = tile * KM_NA_THREADS + threadIdx.x;
    const bool live = i < n;
    const float *xrow = x + (live ? i : 0) * (int64_t)d;
    double best = __longlong_as_double(0x7FF0000000000000LL);
    int lab = 0;
    for (int j0 = 0; j0 < k; j0 += chunk_rows) {
      const int rows = min(chunk_rows, k - j0);
      __syncthreads();
      for (int e = threadIdx.x; e < rows * d; e += KM_NA_THREADS) cs[e] = c64[(int64_t)j0 * d + e];
      __syncthreads();
      if (live)
        for (int j = 0; j < rows; j++) {
          const double dj = km_dist2_f64(xrow, cs + (size_t)j * d, d);
          if (dj < best) { best = dj; lab = j0 + j; }
        }
    }
    if (live) {
      const int old = labels[i];
      if (lab != old) labels[i] = lab;
      km_commit_point(T, S, xrow, d, old, lab, best, /*incremental=*/0, scale_a, scale_b, inertia_scale);
    }
  }
  __syncthreads();
  if (tab_in_smem) km_tab_flush(stab, table, L.scal);
  km_flush_scalars(S, table, L);
}

// one block: c = sum / count (empty cluster keeps its centroid), counts, and the call's result
__global__ void __launch_bounds__(256)
km_na_finalize_kernel(const long long *__restrict__ table, const KmNaHeader *__restrict__ hdr, const double *c_in,
                      double *c_out, int64_t *counts_out, km_noalloc_result *result, int32_t k, int32_t d, int64_t n) {
  const KmAccLayout L = km_acc_layout(k, d);
  const int QB = hdr->st.QB, shift = hdr->st.qa + hdr->st.QB;
  for (int64_t idx = threadIdx.x; idx < (int64_t)k * d; idx += blockDim.x) {
    const long long cnt = table[L.cnt + idx / d];
    double c = c_in[idx];
    if (cnt > 0) c = __ddiv_rn(km_fixed_to_double(table[L.sa + idx], table[L.sb + idx], QB, shift), (double)cnt);
    c_out[idx] = c;
  }
  for (int j = threadIdx.x; j < k; j += blockDim.x) counts_out[j] = table[L.cnt + j];
  if (threadIdx.x == 0) {
    result->changed = table[L.scal + KM_ACC_CHANGED];
    result->inertia = km_fixed_to_double(table[L.scal + KM_ACC_INERTIA_A], table[L.scal + KM_ACC_INERTIA_B], QB,
                                         hdr->st.qi + QB);
    result->n = n;
    result->status = hdr->nonfinite ? KM_ERR_NONFINITE : KM_OK;
    result->pad = 0;
  }
}

static km_status na_fail(km_status code, const std::string &msg) {
  km_internal_set_thread_error(msg.c_str());
  return code;
}
#define KM_NA_CUDA(expr)                                                                       \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) return na_fail(KM_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// One Lloyd iteration (assign every point; centroids = cluster means, an empty cluster keeps
// its centroid) on caller-owned device buffers, enqueued on `stream` without waiting for it.
// Calls no cudaMalloc / cudaFree / cudaHostAlloc or any routine that allocates.
extern "C" km_status km_lloyd_step_noalloc(int32_t device, void *stream, const float *x, int64_t n, int32_t d,
                                           const double *centroids_in, int32_t k, int32_t *labels,
                                           double *centroids_out, int64_t *counts_out, km_noalloc_result *result) {
  if (device < 0 || device >= KM_NOALLOC_MAX_DEVICES) return na_fail(KM_ERR_INVALID_ARG, "km_lloyd_step_noalloc: device out of range");
  if (n < 0 || d < 1 || k < 1 || !centroids_in || !centroids_out || !counts_out || !result || (n > 0 && (!x || !labels)))
    return na_fail(KM_ERR_INVALID_ARG, "km_lloyd_step_noalloc: need n >= 0, d >= 1, k >= 1 and non-NULL device pointers");
  const KmAccLayout L = km_acc_layout(k, d);
  if ((int64_t)k * (2 * (int64_t)d + 1) + KM_ACC_SCALARS > KM_NOALLOC_MAX_ENTRIES || L.len > KM_NOALLOC_MAX_ENTRIES)
    return na_fail(KM_ERR_UNSUPPORTED,
                   "km_lloyd_step_noalloc: k*(2d+1)+4 = " + std::to_string((int64_t)k * (2 * (int64_t)d + 1) + KM_ACC_SCALARS) +
                       " exceeds the built-in scratch of " + std::to_string(KM_NOALLOC_MAX_ENTRIES) +
                       " entries (256 KiB); use the context API for this shape");
  // centroid staging must fit the default 48 KB of shared memory (no attribute calls here)
  if ((size_t)d * sizeof(double) > 40 * 1024) return na_fail(KM_ERR_UNSUPPORTED, "km_lloyd_step_noalloc: d too large for one centroid row in shared memory");
  // run on `device`, and hand the caller's current device back on every way out
  struct DeviceGuard {
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  } guard;
  int cur = -1;
  KM_NA_CUDA(cudaGetDevice(&cur));
  if (cur != device) {
    KM_NA_CUDA(cudaSetDevice(device));
    guard.prev = cur;
  }
  void *base = nullptr;
  KM_NA_CUDA(na_scratch_address(device, &base));
  cudaStream_t s = (cudaStream_t)stream;
  KmNaHeader *hdr = reinterpret_cast<KmNaHeader *>(base);
  long long *table = reinterpret_cast<long long *>(reinterpret_cast<unsigned char *>(base) + KM_NOALLOC_HEADER_BYTES);
  // the array still holds the previous call's sums: clear the header and this shape's table
  KM_NA_CUDA(cudaMemsetAsync(base, 0, KM_NOALLOC_HEADER_BYTES + (size_t)L.len * sizeof(long long), s));
  int sm_count = 148;
  cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device);
  const int64_t count = n * (int64_t)d;
  if (count > 0) {
    int64_t want = (count + 2047) / 2048;
    int grid = (int)(want < (int64_t)sm_count * 8 ? want : (int64_t)sm_count * 8);
    km_na_scan_kernel<<<grid, 256, 0, s>>>(x, count, hdr);
  }
  km_na_setup_kernel<<<1, 256, 0, s>>>(hdr, centroids_in, k, d, n);
  if (n > 0) {
    int chunk_rows = (int)(32768 / ((size_t)d * sizeof(double)));
    if (chunk_rows < 1) chunk_rows = 1;
    if (chunk_rows > k) chunk_rows = k;
    size_t smem = (size_t)chunk_rows * d * sizeof(double);
    int tab_in_smem = 0;
    if (smem + (size_t)L.scal * sizeof(long long) <= 46 * 1024) {
      tab_in_smem = 1;
      smem += (size_t)L.scal * sizeof(long long);
    }
    const int64_t ntiles = (n + KM_NA_THREADS - 1) / KM_NA_THREADS;
    const int64_t cap = (int64_t)sm_count * 8;
    km_na_assign_kernel<<<(int)(ntiles < cap ? ntiles : cap), KM_NA_THREADS, smem, s>>>(x, n, d, k, labels, centroids_in,
                                                                                        table, hdr, chunk_rows, tab_in_smem);
  }
  km_na_finalize_kernel<<<1, 256, 0, s>>>(table, hdr, centroids_in, centroids_out, counts_out, result, k, d, n);
  KM_NA_CUDA(cudaGetLastError());
  return KM_OK;
}
