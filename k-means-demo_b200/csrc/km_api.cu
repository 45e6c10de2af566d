// km_api.cu — the C ABI of libkmeans.so (include/kmeans.h): context, point/centroid
// store, Lloyd driver loop, NCCL plumbing.  No CPU fallback anywhere: every compute
// entry point enqueues CUDA kernels or fails.
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "km_internal.h"

// ------------------------------------------------------------------ errors
static thread_local std::string g_create_err;

static km_status fail(km_ctx *ctx, km_status code, const std::string &msg) {
  if (ctx) ctx->err = msg; else g_create_err = msg;
  return code;
}
// for the translation units that have no ctx to hang an error text on
void km_internal_set_thread_error(const char *msg) { g_create_err = msg ? msg : ""; }

#define KM_CUDA(ctx, expr)                                                                      \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      return fail(ctx, _e == cudaErrorMemoryAllocation ? KM_ERR_OOM : KM_ERR_CUDA,              \
                  std::string(#expr) + ": " + cudaGetErrorString(_e));                          \
    }                                                                                           \
  } while (0)

// ------------------------------------------------------------------ NCCL, loaded lazily
// Nothing links against NCCL: single-GPU users never need it, and a process that already
// carries an NCCL (e.g. a Python host with torch) shares that copy through the SONAME.
struct KmNcclUniqueId { char internal[128]; };
typedef int (*nccl_get_unique_id_t)(KmNcclUniqueId *);
typedef int (*nccl_comm_init_rank_t)(void **, int, KmNcclUniqueId, int);
typedef int (*nccl_comm_destroy_t)(void *);
typedef int (*nccl_all_reduce_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*nccl_all_gather_t)(const void *, void *, size_t, int, void *, cudaStream_t);
typedef const char *(*nccl_get_error_string_t)(int);
enum { KM_NCCL_INT64 = 4, KM_NCCL_FLOAT64 = 8, KM_NCCL_SUM = 0, KM_NCCL_MAX = 2 };

static struct {
  void *handle = nullptr;
  nccl_get_unique_id_t get_unique_id = nullptr;
  nccl_comm_init_rank_t comm_init_rank = nullptr;
  nccl_comm_destroy_t comm_destroy = nullptr;
  nccl_all_reduce_t all_reduce = nullptr;
  nccl_all_gather_t all_gather = nullptr;
  nccl_get_error_string_t get_error_string = nullptr;
} g_nccl;

static bool nccl_load(std::string &why) {
  if (g_nccl.handle) return true;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { why = std::string("dlopen libnccl.so.2: ") + dlerror(); return false; }
  g_nccl.get_unique_id = (nccl_get_unique_id_t)dlsym(h, "ncclGetUniqueId");
  g_nccl.comm_init_rank = (nccl_comm_init_rank_t)dlsym(h, "ncclCommInitRank");
  g_nccl.comm_destroy = (nccl_comm_destroy_t)dlsym(h, "ncclCommDestroy");
  g_nccl.all_reduce = (nccl_all_reduce_t)dlsym(h, "ncclAllReduce");
  g_nccl.all_gather = (nccl_all_gather_t)dlsym(h, "ncclAllGather");
  g_nccl.get_error_string = (nccl_get_error_string_t)dlsym(h, "ncclGetErrorString");
  if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.comm_destroy || !g_nccl.all_reduce ||
      !g_nccl.all_gather || !g_nccl.get_error_string) {
    why = "libnccl.so.2 lacks an expected symbol";
    return false;
  }
  g_nccl.handle = h;
  return true;
}
#define KM_NCCL(ctx, expr)                                                                       \
  do {                                                                                           \
    int _r = (expr);                                                                             \
    if (_r != 0) return fail(ctx, KM_ERR_NCCL, std::string(#expr) + ": " + g_nccl.get_error_string(_r)); \
  } while (0)

// ------------------------------------------------------------------ helpers
static size_t elt_size(int dtype) { return dtype == KM_DTYPE_F32 ? sizeof(float) : sizeof(double); }

static km_status bind(km_ctx *ctx) {
  if (!ctx) return KM_ERR_INVALID_ARG;
  KM_CUDA(ctx, cudaSetDevice(ctx->device));
  return KM_OK;
}
#define KM_BIND(ctx)                         \
  do {                                       \
    km_status _s = bind(ctx);                \
    if (_s != KM_OK) return _s;              \
  } while (0)

static void free_points(km_ctx *ctx) {
  if (ctx->x && ctx->own_x) cudaFree(ctx->x);
  if (ctx->labels) cudaFree(ctx->labels);
  if (ctx->perm) cudaFree(ctx->perm);
  if (ctx->unperm_buf) cudaFree(ctx->unperm_buf);
  ctx->perm = nullptr; ctx->unperm_buf = nullptr; ctx->passes_since_upload = 0;
  if (ctx->t_work) cudaFree(ctx->t_work);
  if (ctx->t_order) cudaFree(ctx->t_order);
  if (ctx->t_xn) cudaFree(ctx->t_xn);
  if (ctx->t_xbf) cudaFree(ctx->t_xbf);
  ctx->t_xbf = nullptr; ctx->t_xbf_bytes = 0;
  ctx->t_work = nullptr; ctx->t_order = nullptr; ctx->t_xn = nullptr; ctx->t_work_n = 0; ctx->t_xn_valid = false;
  ctx->x = nullptr; ctx->labels = nullptr; ctx->own_x = false; ctx->have_points = false;
  ctx->n = 0; ctx->n_total = 0; ctx->i0 = 0;
}
static void free_centroids(km_ctx *ctx) {
  if (ctx->c64) cudaFree(ctx->c64);
  if (ctx->chi) cudaFree(ctx->chi);
  if (ctx->clo) cudaFree(ctx->clo);
  if (ctx->acc_global && ctx->acc_global != ctx->acc) cudaFree(ctx->acc_global);
  if (ctx->acc) cudaFree(ctx->acc);
  ctx->c64 = nullptr; ctx->chi = ctx->clo = nullptr; ctx->acc = ctx->acc_global = nullptr;
  for (int q = 0; q < 2; q++) { if (ctx->p1S[q]) cudaFree(ctx->p1S[q]); if (ctx->p1C[q]) cudaFree(ctx->p1C[q]); ctx->p1S[q] = nullptr; ctx->p1C[q] = nullptr; }
  for (int q = 0; q < 3; q++) {
    if (ctx->p1D[q] && !ctx->p1D_in_peer) cudaFree(ctx->p1D[q]);
    if (ctx->p1R[q]) cudaFree(ctx->p1R[q]);
    ctx->p1D[q] = ctx->p1R[q] = nullptr;
  }
  ctx->p1D_in_peer = false;
  if (ctx->t_cn) cudaFree(ctx->t_cn);
  if (ctx->t_cbf) cudaFree(ctx->t_cbf);
  if (ctx->t_hist) cudaFree(ctx->t_hist);
  ctx->t_cn = nullptr; ctx->t_cbf = nullptr; ctx->t_hist = nullptr; ctx->t_work_count = nullptr;
  ctx->p1 = false; ctx->p1_pending = false;
  ctx->have_centroids = false; ctx->k = 0; ctx->acc_len = 0;
}

static km_status ensure_pinned(km_ctx *ctx, size_t bytes) {
  if (ctx->pinned_bytes >= bytes) return KM_OK;
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  ctx->pinned = nullptr; ctx->pinned_bytes = 0;
  KM_CUDA(ctx, cudaHostAlloc(&ctx->pinned, bytes, cudaHostAllocDefault));
  ctx->pinned_bytes = bytes;
  return KM_OK;
}

static bool is_pinned_host(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

// host -> device through pinned staging (two halves in flight), or directly when the
// caller's buffer is already page-locked
static km_status copy_h2d(km_ctx *ctx, void *dst, const void *src, size_t bytes) {
  if (bytes == 0) return KM_OK;
  if (is_pinned_host(src)) {
    KM_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return KM_OK;
  }
  const size_t half = 32u << 20;
  km_status s = ensure_pinned(ctx, 2 * half);
  if (s != KM_OK) return s;
  cudaEvent_t done[2];
  KM_CUDA(ctx, cudaEventCreateWithFlags(&done[0], cudaEventDisableTiming));
  KM_CUDA(ctx, cudaEventCreateWithFlags(&done[1], cudaEventDisableTiming));
  size_t off = 0;
  int slot = 0;
  bool used[2] = {false, false};
  cudaError_t e = cudaSuccess;
  while (off < bytes && e == cudaSuccess) {
    size_t len = std::min(half, bytes - off);
    char *stage = (char *)ctx->pinned + (size_t)slot * half;
    if (used[slot]) e = cudaEventSynchronize(done[slot]);
    if (e != cudaSuccess) break;
    memcpy(stage, (const char *)src + off, len);
    e = cudaMemcpyAsync((char *)dst + off, stage, len, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaEventRecord(done[slot], ctx->stream);
    used[slot] = true;
    off += len;
    slot ^= 1;
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaEventDestroy(done[0]);
  cudaEventDestroy(done[1]);
  KM_CUDA(ctx, e);
  return KM_OK;
}

static km_status copy_d2h(km_ctx *ctx, void *dst, const void *src, size_t bytes) {
  if (bytes == 0) return KM_OK;
  if (is_pinned_host(dst)) {
    KM_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return KM_OK;
  }
  // pageable destination: two pinned halves, the copy of chunk i+1 flies while chunk i is
  // memcpy'd out of the staging buffer
  const size_t chunk = 32u << 20;
  km_status s = ensure_pinned(ctx, 2 * chunk);
  if (s != KM_OK) return s;
  cudaEvent_t done[2];
  KM_CUDA(ctx, cudaEventCreateWithFlags(&done[0], cudaEventDisableTiming));
  KM_CUDA(ctx, cudaEventCreateWithFlags(&done[1], cudaEventDisableTiming));
  cudaError_t e = cudaSuccess;
  size_t issued = 0, drained = 0;
  int slot_issue = 0, slot_drain = 0;
  size_t len_of[2] = {0, 0};
  while (drained < bytes && e == cudaSuccess) {
    while (issued < bytes && issued - drained < 2 * chunk && e == cudaSuccess) {
      const size_t len = std::min(chunk, bytes - issued);
      e = cudaMemcpyAsync((char *)ctx->pinned + (size_t)slot_issue * chunk, (const char *)src + issued, len,
                          cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaEventRecord(done[slot_issue], ctx->stream);
      len_of[slot_issue] = len;
      issued += len;
      slot_issue ^= 1;
    }
    if (e != cudaSuccess) break;
    e = cudaEventSynchronize(done[slot_drain]);
    if (e != cudaSuccess) break;
    memcpy((char *)dst + drained, (const char *)ctx->pinned + (size_t)slot_drain * chunk, len_of[slot_drain]);
    drained += len_of[slot_drain];
    slot_drain ^= 1;
  }
  if (e != cudaSuccess) cudaStreamSynchronize(ctx->stream);
  cudaEventDestroy(done[0]);
  cudaEventDestroy(done[1]);
  KM_CUDA(ctx, e);
  return KM_OK;
}

static int bitlen64(int64_t v) { int b = 1; while (b < 62 && ((int64_t)1 << b) <= v) b++; return b; }

// fixed-point parameters of the order-free sums from the job's max|x| and N
static int bitlen32(int32_t v) { int b = 0; while ((1 << b) <= v) b++; return b; }

// grids of the order-free sums and of T = sum |x|^2 (§4.1) from the job's max|x| and N
static void compute_grids(km_ctx *ctx) {
  int E = 0;
  if (ctx->maxabs > 0.0) E = ilogb(ctx->maxabs) + 1;
  int NB = bitlen64(ctx->n_total);
  ctx->E = E; ctx->NB = NB; ctx->QB = 62 - NB; ctx->qa = ctx->QB - E;
  ctx->QBt = 61 - NB - bitlen32(ctx->d);
  ctx->qt = std::min(ctx->QBt - 2 * E, 900);
}
// grid of the inertia terms c (2S - n c): |c| is bounded by max(max|x|, max|c0|)
static void inertia_grid(const km_ctx *ctx, int *qg, int *QBg) {
  const double M = std::max(ctx->cbound, ctx->maxabs);
  const int EM = M > 0.0 ? ilogb(M) + 1 : 0;
  *QBg = 60 - ctx->NB - bitlen32(ctx->d);
  *qg = std::min(*QBg - 2 * EM, 900);
}

static km_status push_state(km_ctx *ctx) {
  const int E = ctx->E;
  KmState st;
  memset(&st, 0, sizeof(st));
  st.QB = ctx->QB; st.qa = ctx->qa; st.E = E;
  st.Ta = ctx->Ta; st.Tb = ctx->Tb; st.qt = ctx->qt; st.QBt = ctx->QBt;
  st.scale_a = ldexp(1.0, ctx->qa);
  st.scale_b = ldexp(1.0, ctx->QB);
  st.qi = ctx->QB;  // refreshed with the centroids
  st.inertia_scale = ldexp(1.0, st.qi);
  KM_CUDA(ctx, cudaMemcpyAsync(ctx->state, &st, sizeof(st), cudaMemcpyHostToDevice, ctx->stream));
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return KM_OK;
}

static km_status peer_barrier(km_ctx *ctx);
static km_status reset_labels(km_ctx *ctx) {
  if (ctx->labels) KM_CUDA(ctx, km_launch_fill_i32(ctx->labels, ctx->n, -1, ctx->sm_count, ctx->stream, &ctx->launches));
  if (ctx->acc) KM_CUDA(ctx, cudaMemsetAsync(ctx->acc, 0, (size_t)ctx->acc_len * sizeof(long long), ctx->stream));
  if (ctx->acc_global && ctx->acc_global != ctx->acc)
    KM_CUDA(ctx, cudaMemsetAsync(ctx->acc_global, 0, (size_t)ctx->acc_len * sizeof(long long), ctx->stream));
  if (ctx->p1) {
    if (ctx->p1D_in_peer) {  // a peer may still be summing the last pass's table
      km_status bs = peer_barrier(ctx);
      if (bs != KM_OK) return bs;
    }
    for (int q = 0; q < 2; q++) KM_CUDA(ctx, cudaMemsetAsync(ctx->p1S[q], 0, (size_t)ctx->acc_len * sizeof(long long), ctx->stream));
    for (int q = 0; q < 3; q++) KM_CUDA(ctx, cudaMemsetAsync(ctx->p1D[q], 0, (size_t)ctx->acc_len * sizeof(long long), ctx->stream));
    ctx->p1_pending = false;
    ctx->p1_dw = 0;
  }
  if (ctx->state) KM_CUDA(ctx, cudaMemsetAsync(&ctx->state->it, 0, sizeof(int32_t), ctx->stream));
  ctx->host_it = 0;
  ctx->labels_fresh = true;
  ctx->t_order_valid = false;
  return KM_OK;
}

static KmFinalizeArgs make_fin(km_ctx *ctx, KmStatSlot *slot, int do_stats, int do_update, int needs_change) {
  KmFinalizeArgs f;
  f.fused = 0;
  f.acc_in = ctx->acc_global; f.acc_local = ctx->acc;
  f.c64 = ctx->c64; f.chi = ctx->chi; f.clo = ctx->clo;
  f.k = ctx->k; f.d = ctx->d; f.st = ctx->state; f.slot = slot; f.maxabs_x = ctx->maxabs;
  f.do_stats = do_stats; f.do_update = do_update; f.update_needs_change = needs_change;
  inertia_grid(ctx, &f.qg, &f.QBg);
  return f;
}

// after the shard is on the device: scan it, agree on the job-wide scale, reset labels
static km_status adopt_points(km_ctx *ctx) {
  double *d_max = ctx->scan_max;
  int *d_bad = ctx->scan_bad;
  cudaError_t e = km_launch_scan(ctx->x, ctx->dtype, ctx->n * (int64_t)ctx->d, d_max, d_bad, ctx->sm_count, ctx->stream,
                                 &ctx->launches);
  double hmax = 0.0;
  int hbad = 0;
  unsigned hbox[4] = {0xFFFFFFFFu, 0u, 0xFFFFFFFFu, 0u};
  const bool want_bbox = ctx->d == 2 && ctx->dtype == KM_DTYPE_F32 && (reinterpret_cast<uintptr_t>(ctx->x) & 7) == 0;
  if (e == cudaSuccess && want_bbox) {
    e = km_launch_bbox2((const float *)ctx->x, ctx->n, ctx->scan_bbox, ctx->sm_count, ctx->stream, &ctx->launches);
    if (e == cudaSuccess) e = cudaMemcpyAsync(hbox, ctx->scan_bbox, sizeof(hbox), cudaMemcpyDeviceToHost, ctx->stream);
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(&hmax, d_max, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(&hbad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  int64_t n_total = ctx->n, i0 = 0;
  auto ord2f = [](unsigned u) { unsigned b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u; float f; memcpy(&f, &b, 4); return (double)f; };
  // this rank's box; an empty shard contributes an inverted (ignored) one
  double box[4] = {INFINITY, -INFINITY, INFINITY, -INFINITY};
  if (want_bbox && ctx->n > 0 && e == cudaSuccess) for (int q = 0; q < 4; q++) box[q] = ord2f(hbox[q]);
  const int NG = 7;
  if (e == cudaSuccess && ctx->world > 1) {
    // job-wide: max of max|x|, any non-finite, shard offsets, bounding box — one all-gather
    double *d_in = nullptr, *d_out = nullptr;
    e = cudaMalloc(&d_in, NG * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d_out, NG * sizeof(double) * ctx->world);
    double h_in[NG] = {hmax, (double)hbad, (double)ctx->n, box[0], box[1], box[2], box[3]};
    std::vector<double> h_out(NG * (size_t)ctx->world);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, h_in, sizeof(h_in), cudaMemcpyHostToDevice, ctx->stream);
    int r = 0;
    if (e == cudaSuccess) r = g_nccl.all_gather(d_in, d_out, NG, KM_NCCL_FLOAT64, ctx->nccl_comm, ctx->stream);
    if (e == cudaSuccess && r == 0)
      e = cudaMemcpyAsync(h_out.data(), d_out, h_out.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && r == 0) e = cudaStreamSynchronize(ctx->stream);
    if (d_in) cudaFree(d_in);
    if (d_out) cudaFree(d_out);
    if (r != 0) { return fail(ctx, KM_ERR_NCCL, std::string("ncclAllGather: ") + g_nccl.get_error_string(r)); }
    if (e == cudaSuccess) {
      n_total = 0;
      for (int q = 0; q < ctx->world; q++) {
        hmax = std::max(hmax, h_out[NG * q]);
        hbad |= (h_out[NG * q + 1] != 0.0);
        if (q < ctx->rank) i0 += (int64_t)h_out[NG * q + 2];
        n_total += (int64_t)h_out[NG * q + 2];
        box[0] = std::min(box[0], h_out[NG * q + 3]); box[1] = std::max(box[1], h_out[NG * q + 4]);
        box[2] = std::min(box[2], h_out[NG * q + 5]); box[3] = std::max(box[3], h_out[NG * q + 6]);
      }
    }
  }
  KM_CUDA(ctx, e);
  if (hbad) { free_points(ctx); return fail(ctx, KM_ERR_NONFINITE, "points contain NaN or Inf"); }
  ctx->maxabs = hmax;
  ctx->have_bbox = want_bbox && box[0] <= box[1] && box[2] <= box[3];
  for (int q = 0; q < 4; q++) ctx->bbox[q] = box[q];
  ctx->n_total = n_total;
  ctx->i0 = i0;
  ctx->have_points = true;
  compute_grids(ctx);
  {
    // T = sum |x|^2 over the whole job, once per upload: exact limbs on the grid just fixed
    KM_CUDA(ctx, km_launch_sumsq(ctx->x, ctx->dtype, ctx->n * (int64_t)ctx->d, ldexp(1.0, ctx->qt), ldexp(1.0, ctx->QBt),
                                 ctx->scan_t, ctx->sm_count, ctx->stream, &ctx->launches));
    if (ctx->world > 1)
      KM_NCCL(ctx, g_nccl.all_reduce(ctx->scan_t, ctx->scan_t, 2, KM_NCCL_INT64, KM_NCCL_SUM, ctx->nccl_comm, ctx->stream));
    long long ht[2] = {0, 0};
    KM_CUDA(ctx, cudaMemcpyAsync(ht, ctx->scan_t, sizeof(ht), cudaMemcpyDeviceToHost, ctx->stream));
    KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->Ta = ht[0]; ctx->Tb = ht[1];
  }
  km_status s = push_state(ctx);
  if (s != KM_OK) return s;
  s = reset_labels(ctx);
  if (s != KM_OK) return s;
  if (ctx->have_centroids) {
    if (ctx->d * (int64_t)ctx->k * 2 + ctx->k + KM_ACC_SCALARS != ctx->acc_len) {
      free_centroids(ctx);  // shape changed under the old centroids
    } else if (!ctx->p1) {
      KM_CUDA(ctx, km_launch_finalize(make_fin(ctx, nullptr, 0, 0, 0), ctx->stream, &ctx->launches));
    }
  }
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return KM_OK;
}

static km_status check_shape(km_ctx *ctx, int64_t n, int32_t d, int32_t dtype) {
  if (n < 0 || d < 1 || d > KM_MAX_D) return fail(ctx, KM_ERR_INVALID_ARG, "need n >= 0 and 1 <= d <= 65536");
  if (dtype != KM_DTYPE_F32 && dtype != KM_DTYPE_F64) return fail(ctx, KM_ERR_INVALID_ARG, "dtype must be KM_DTYPE_F32 or KM_DTYPE_F64");
  return KM_OK;
}

static km_status alloc_points(km_ctx *ctx, int64_t n, int32_t d, int32_t dtype, bool alloc_x) {
  // same shape as last time and we own the buffer: keep the allocations (an upload per
  // step must not pay cudaMalloc/cudaFree)
  if (alloc_x && ctx->own_x && ctx->x && ctx->labels && ctx->n == n && ctx->d == d && ctx->dtype == dtype) {
    ctx->have_points = false;
    ctx->t_xn_valid = false;              // new rows: their norms are not known yet
    if (ctx->perm) cudaFree(ctx->perm);   // new rows arrive in the caller's order
    ctx->perm = nullptr; ctx->passes_since_upload = 0;
    return KM_OK;
  }
  free_points(ctx);
  if (ctx->have_centroids && ctx->d != d) free_centroids(ctx);
  ctx->n = n; ctx->d = d; ctx->dtype = dtype;
  size_t bytes = (size_t)n * d * elt_size(dtype);
  if (alloc_x) {
    KM_CUDA(ctx, cudaMalloc(&ctx->x, std::max<size_t>(bytes, 16)));
    ctx->own_x = true;
  }
  KM_CUDA(ctx, cudaMalloc(&ctx->labels, std::max<size_t>((size_t)n * sizeof(int32_t), 16)));
  return KM_OK;
}

// ------------------------------------------------------------------ life cycle
extern "C" int32_t km_abi_version(void) { return KM_ABI_VERSION; }

extern "C" const char *km_last_error(const km_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" km_status km_create(const km_config *cfg, km_ctx **out) {
  if (!out) return fail(nullptr, KM_ERR_INVALID_ARG, "km_create: out is NULL");
  *out = nullptr;
  km_config c;
  memset(&c, 0, sizeof(c));
  c.struct_size = sizeof(km_config);
  c.exact_inertia = 1;
  if (cfg) {
    if (cfg->struct_size != (int32_t)sizeof(km_config)) return fail(nullptr, KM_ERR_INVALID_ARG, "km_config.struct_size mismatch");
    c = *cfg;
  }
  if (c.path < KM_PATH_AUTO || c.path > KM_PATH_TENSOR) return fail(nullptr, KM_ERR_INVALID_ARG, "km_config.path out of range");
  if (c.update_mode != KM_UPDATE_INCREMENTAL && c.update_mode != KM_UPDATE_FULL)
    return fail(nullptr, KM_ERR_INVALID_ARG, "km_config.update_mode out of range");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, KM_ERR_CUDA, std::string("no usable CUDA device (libkmeans has no CPU fallback): ") +
                                          (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
  if (c.device < 0 || c.device >= ndev) return fail(nullptr, KM_ERR_INVALID_ARG, "km_config.device out of range");
  km_ctx *ctx = new km_ctx();
  ctx->cfg = c;
  ctx->device = c.device;
  for (int q = 0; q < KM_RING; q++) ctx->ring_ev[q] = nullptr;
  for (int q = 0; q < KM_PROF_RING; q++) ctx->prof_ev[q][0] = ctx->prof_ev[q][1] = nullptr;
#define KM_CREATE_CUDA(expr)                                                                  \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      g_create_err = std::string(#expr) + ": " + cudaGetErrorString(_e);                      \
      km_destroy(ctx);                                                                        \
      return _e == cudaErrorMemoryAllocation ? KM_ERR_OOM : KM_ERR_CUDA;                      \
    }                                                                                         \
  } while (0)
  KM_CREATE_CUDA(cudaSetDevice(ctx->device));
  KM_CREATE_CUDA(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, ctx->device));
  KM_CREATE_CUDA(cudaDeviceGetAttribute(&ctx->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, ctx->device));
  KM_CREATE_CUDA(cudaDeviceGetAttribute(&ctx->coop_ok, cudaDevAttrCooperativeLaunch, ctx->device));
  if (c.stream) {
    ctx->stream = (cudaStream_t)c.stream;
  } else {
    KM_CREATE_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    ctx->own_stream = true;
  }
  KM_CREATE_CUDA(cudaMalloc(&ctx->state, sizeof(KmState)));
  KM_CREATE_CUDA(cudaMemset(ctx->state, 0, sizeof(KmState)));
  KM_CREATE_CUDA(cudaMalloc(&ctx->scan_max, sizeof(double)));
  KM_CREATE_CUDA(cudaMalloc(&ctx->scan_bad, sizeof(int)));
  KM_CREATE_CUDA(cudaMalloc(&ctx->scan_bbox, 4 * sizeof(unsigned)));
  KM_CREATE_CUDA(cudaMalloc(&ctx->scan_t, 2 * sizeof(long long)));
  for (int q = 0; q < KM_PROF_RING; q++) {
    KM_CREATE_CUDA(cudaEventCreate(&ctx->prof_ev[q][0]));
    KM_CREATE_CUDA(cudaEventCreate(&ctx->prof_ev[q][1]));
  }
  KM_CREATE_CUDA(cudaHostAlloc((void **)&ctx->ring, sizeof(KmStatSlot) * KM_RING, cudaHostAllocMapped));
  memset(ctx->ring, 0, sizeof(KmStatSlot) * KM_RING);
  KM_CREATE_CUDA(cudaHostGetDevicePointer((void **)&ctx->ring_dev, ctx->ring, 0));
  for (int q = 0; q < KM_RING; q++) KM_CREATE_CUDA(cudaEventCreateWithFlags(&ctx->ring_ev[q], cudaEventDisableTiming));
  KM_CREATE_CUDA(cudaEventCreate(&ctx->ev0));
  KM_CREATE_CUDA(cudaEventCreate(&ctx->ev1));
#undef KM_CREATE_CUDA
  *out = ctx;
  return KM_OK;
}

// ------------------------------------------------------------------ peer regions (km_peer.cu)
// Each rank allocates one region, the ranks exchange CUDA IPC handles through the communicator and
// map each other's.  Any failure on any rank (ranks on different nodes, no peer access, IPC
// unavailable) leaves every rank on the NCCL all-reduce: the decision is itself all-reduced.
static void peer_teardown(km_ctx *ctx) {
  for (int r = 0; r < KM_PEER_MAX_WORLD; r++) {
    if (!ctx->peer_base[r]) continue;
    if (r == ctx->rank) cudaFree(ctx->peer_base[r]);
    else cudaIpcCloseMemHandle(ctx->peer_base[r]);
    ctx->peer_base[r] = nullptr;
  }
  if (ctx->peer_err) cudaFreeHost(ctx->peer_err);
  ctx->peer_err = nullptr;
  ctx->peer_ok = false;
  cudaGetLastError();
}

static km_status peer_setup(km_ctx *ctx) {
  ctx->peer_ok = false;
  if (ctx->world < 2 || ctx->world > KM_PEER_MAX_WORLD) return KM_OK;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle travels as 8 int64");
  const int W = ctx->world, NH = 9;  // 8 words of handle + "this rank is fine so far"
  const size_t alloc_bytes = KM_PEER_REGION_BYTES;  // 2 MB: a whole allocation block of its own, so the handle maps exactly this pointer
  static_assert(KM_PEER_SNAP_OFF + 2 * KM_PEER_SNAP_BYTES <= KM_PEER_REGION_BYTES, "the snapshots fit the region");
  bool ok = true;
  void *base = nullptr;
  ok = cudaMalloc(&base, alloc_bytes) == cudaSuccess && cudaMemset(base, 0, alloc_bytes) == cudaSuccess &&
       cudaDeviceSynchronize() == cudaSuccess;
  ctx->peer_base[ctx->rank] = ok ? base : nullptr;
  long long h_in[NH];
  memset(h_in, 0, sizeof(h_in));
  cudaIpcMemHandle_t mine;
  if (ok) ok = cudaIpcGetMemHandle(&mine, base) == cudaSuccess;
  if (ok) memcpy(h_in, &mine, sizeof(mine));
  if (ok) ok = cudaHostAlloc(&ctx->peer_err, sizeof(int), cudaHostAllocMapped) == cudaSuccess;
  if (ok) *ctx->peer_err = 0;
  h_in[8] = ok ? 1 : 0;
  cudaGetLastError();
  // the two collectives below are joined by every rank whatever happened above
  long long *d_buf = nullptr;
  std::vector<long long> h_out((size_t)NH * W);
  KM_CUDA(ctx, cudaMalloc(&d_buf, sizeof(long long) * NH * (W + 1)));
  cudaError_t e = cudaMemcpyAsync(d_buf, h_in, sizeof(h_in), cudaMemcpyHostToDevice, ctx->stream);
  int rc = g_nccl.all_gather(d_buf, d_buf + NH, NH, KM_NCCL_INT64, ctx->nccl_comm, ctx->stream);
  if (e == cudaSuccess && rc == 0)
    e = cudaMemcpyAsync(h_out.data(), d_buf + NH, h_out.size() * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && rc == 0) e = cudaStreamSynchronize(ctx->stream);
  if (rc != 0 || e != cudaSuccess) {
    cudaFree(d_buf);
    peer_teardown(ctx);
    if (rc != 0) return fail(ctx, KM_ERR_NCCL, std::string("ncclAllGather: ") + g_nccl.get_error_string(rc));
    return fail(ctx, KM_ERR_CUDA, cudaGetErrorString(e));
  }
  for (int r = 0; r < W; r++) ok = ok && h_out[(size_t)NH * r + 8] == 1;
  for (int r = 0; r < W && ok; r++) {
    if (r == ctx->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, &h_out[(size_t)NH * r], sizeof(h));
    void *p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = false; cudaGetLastError(); break; }
    ctx->peer_base[r] = p;
  }
  long long bad = ok ? 0 : 1;
  e = cudaMemcpyAsync(d_buf, &bad, sizeof(bad), cudaMemcpyHostToDevice, ctx->stream);
  rc = g_nccl.all_reduce(d_buf, d_buf, 1, KM_NCCL_INT64, KM_NCCL_SUM, ctx->nccl_comm, ctx->stream);
  if (e == cudaSuccess && rc == 0) e = cudaMemcpyAsync(&bad, d_buf, sizeof(bad), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && rc == 0) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d_buf);
  if (rc != 0 || e != cudaSuccess || bad != 0) {
    peer_teardown(ctx);
    if (rc != 0) return fail(ctx, KM_ERR_NCCL, std::string("ncclAllReduce: ") + g_nccl.get_error_string(rc));
    if (e != cudaSuccess) return fail(ctx, KM_ERR_CUDA, cudaGetErrorString(e));
    return KM_OK;  // some rank could not map a peer: everybody stays on NCCL
  }
  ctx->peer_ok = true;
  ctx->peer_epoch = 0;
  return KM_OK;
}

// every rank has finished reading every table published so far (needed before a table is
// cleared outside the three-pass rotation)
static km_status peer_barrier(km_ctx *ctx) {
  if (!ctx->peer_ok) return KM_OK;
  ctx->peer_epoch++;
  KM_CUDA(ctx, km_launch_peer_allreduce(ctx->peer_base, ctx->world, ctx->rank, 0, KM_PEER_FLAGS_OFF, 0, ctx->peer_epoch, nullptr,
                                        ctx->peer_err, ctx->sm_count, ctx->stream, &ctx->launches));
  return KM_OK;
}

extern "C" km_status km_destroy(km_ctx *ctx) {
  if (!ctx) return KM_OK;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->nccl_comm && g_nccl.comm_destroy) g_nccl.comm_destroy(ctx->nccl_comm);
  free_points(ctx);
  free_centroids(ctx);
  peer_teardown(ctx);
  if (ctx->state) cudaFree(ctx->state);
  if (ctx->scan_max) cudaFree(ctx->scan_max);
  if (ctx->scan_bad) cudaFree(ctx->scan_bad);
  if (ctx->scan_bbox) cudaFree(ctx->scan_bbox);
  if (ctx->scan_t) cudaFree(ctx->scan_t);
  if (ctx->cells_g) cudaFree(ctx->cells_g);
  for (int q = 0; q < KM_PROF_RING; q++)
    for (int w = 0; w < 2; w++) if (ctx->prof_ev[q][w]) cudaEventDestroy(ctx->prof_ev[q][w]);
  if (ctx->ring) cudaFreeHost(ctx->ring);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (int q = 0; q < KM_RING; q++) if (ctx->ring_ev[q]) cudaEventDestroy(ctx->ring_ev[q]);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return KM_OK;
}

extern "C" km_status km_host_alloc(void **out, uint64_t bytes) {
  if (!out) return KM_ERR_INVALID_ARG;
  *out = nullptr;
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
  if (e != cudaSuccess) return fail(nullptr, e == cudaErrorMemoryAllocation ? KM_ERR_OOM : KM_ERR_CUDA,
                                    std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
  return KM_OK;
}
extern "C" km_status km_host_free(void *p) {
  if (p) cudaFreeHost(p);
  return KM_OK;
}

// ------------------------------------------------------------------ multi-GPU
extern "C" km_status km_comm_unique_id(void *out128) {
  if (!out128) return KM_ERR_INVALID_ARG;
  std::string why;
  if (!nccl_load(why)) return fail(nullptr, KM_ERR_NCCL, why);
  KmNcclUniqueId id;
  int r = g_nccl.get_unique_id(&id);
  if (r != 0) return fail(nullptr, KM_ERR_NCCL, std::string("ncclGetUniqueId: ") + g_nccl.get_error_string(r));
  memcpy(out128, &id, sizeof(id));
  return KM_OK;
}

extern "C" km_status km_comm_init(km_ctx *ctx, const void *id128, int32_t rank, int32_t world) {
  KM_BIND(ctx);
  if (!id128 || world < 1 || rank < 0 || rank >= world) return fail(ctx, KM_ERR_INVALID_ARG, "km_comm_init: bad rank/world/id");
  if (ctx->have_points) return fail(ctx, KM_ERR_STATE, "km_comm_init must precede km_upload_points / km_generate_blobs");
  if (ctx->nccl_comm) return fail(ctx, KM_ERR_STATE, "communicator already initialised");
  std::string why;
  if (!nccl_load(why)) return fail(ctx, KM_ERR_NCCL, why);
  KmNcclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  KM_NCCL(ctx, g_nccl.comm_init_rank(&ctx->nccl_comm, world, id, rank));
  ctx->rank = rank;
  ctx->world = world;
  return peer_setup(ctx);
}

extern "C" km_status km_comm_info(const km_ctx *ctx, int32_t *rank, int32_t *world, int32_t *peer_allreduce) {
  if (!ctx) return KM_ERR_INVALID_ARG;
  if (rank) *rank = ctx->rank;
  if (world) *world = ctx->world;
  const bool acc_peer = ctx->have_centroids && !ctx->p1 && ctx->world > 1 && (size_t)ctx->acc_len * sizeof(long long) <= KM_PEER_SNAP_BYTES;
  if (peer_allreduce) *peer_allreduce = !ctx->peer_ok ? 0 : ((ctx->p1D_in_peer || acc_peer) ? 2 : 1);
  return KM_OK;
}

// ------------------------------------------------------------------ a1: store
extern "C" km_status km_upload_points(km_ctx *ctx, const void *host, int64_t n, int32_t d, int32_t dtype) {
  KM_BIND(ctx);
  km_status s = check_shape(ctx, n, d, dtype);
  if (s != KM_OK) return s;
  if (n > 0 && !host) return fail(ctx, KM_ERR_INVALID_ARG, "km_upload_points: host is NULL");
  s = alloc_points(ctx, n, d, dtype, true);
  if (s != KM_OK) return s;
  s = copy_h2d(ctx, ctx->x, host, (size_t)n * d * elt_size(dtype));
  if (s != KM_OK) return s;
  return adopt_points(ctx);
}

extern "C" km_status km_attach_points(km_ctx *ctx, const void *device, int64_t n, int32_t d, int32_t dtype) {
  KM_BIND(ctx);
  km_status s = check_shape(ctx, n, d, dtype);
  if (s != KM_OK) return s;
  if (n > 0 && !device) return fail(ctx, KM_ERR_INVALID_ARG, "km_attach_points: device pointer is NULL");
  if (reinterpret_cast<uintptr_t>(device) & 15) return fail(ctx, KM_ERR_INVALID_ARG, "km_attach_points: pointer must be 16-byte aligned");
  s = alloc_points(ctx, n, d, dtype, false);
  if (s != KM_OK) return s;
  ctx->x = const_cast<void *>(device);
  ctx->own_x = false;
  return adopt_points(ctx);
}

extern "C" km_status km_generate_blobs(km_ctx *ctx, int64_t n, int64_t i0, int32_t d, int32_t k_true, uint64_t seed,
                                       double sigma) {
  KM_BIND(ctx);
  km_status s = check_shape(ctx, n, d, KM_DTYPE_F32);
  if (s != KM_OK) return s;
  if (k_true < 1 || i0 < 0) return fail(ctx, KM_ERR_INVALID_ARG, "km_generate_blobs: need k_true >= 1 and i0 >= 0");
  s = alloc_points(ctx, n, d, KM_DTYPE_F32, true);
  if (s != KM_OK) return s;
  KM_CUDA(ctx, km_launch_gen_blobs((float *)ctx->x, n, i0, d, k_true, seed, sigma, ctx->sm_count, ctx->stream, &ctx->launches));
  s = adopt_points(ctx);
  if (s != KM_OK) return s;
  if (ctx->world == 1) ctx->i0 = 0;  // a single rank's rows are the job's rows 0..n-1 whatever slice it generated
  return KM_OK;
}

extern "C" km_status km_get_points(km_ctx *ctx, void *host, int64_t i0, int64_t n) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "no points");
  if (i0 < 0 || n < 0 || i0 + n > ctx->n || (n > 0 && !host)) return fail(ctx, KM_ERR_INVALID_ARG, "km_get_points: range outside the shard");
  size_t row = (size_t)ctx->d * elt_size(ctx->dtype);
  if (ctx->perm && n > 0) {  // stored in cell order: put the requested rows back in the caller's
    // small requests (a Forgy pick, a prefix) go through the buffer the store keeps for un-permuting labels
    const bool own = (size_t)n * row > (size_t)ctx->n * sizeof(int32_t) || !ctx->unperm_buf;
    float *tmp = reinterpret_cast<float *>(ctx->unperm_buf);
    if (own) KM_CUDA(ctx, cudaMalloc(&tmp, (size_t)n * row));
    cudaError_t e = km_unpermute_rows2((const float *)ctx->x, ctx->perm, ctx->n, i0, n, tmp, ctx->sm_count, ctx->stream, &ctx->launches);
    km_status s = KM_OK;
    if (e == cudaSuccess) s = copy_d2h(ctx, host, tmp, (size_t)n * row);
    if (own) cudaFree(tmp);
    KM_CUDA(ctx, e);
    return s;
  }
  return copy_d2h(ctx, host, (const char *)ctx->x + (size_t)i0 * row, (size_t)n * row);
}

// ------------------------------------------------------------------ a2: initialisation
static int pick_path_for(km_ctx *ctx, int32_t k) {
  int want = ctx->cfg.path;
  bool small_ok = km_small_path_supported(ctx->d, k, ctx->dtype, ctx->max_smem_optin);
  bool tensor_ok = km_tensor_path_supported(ctx->d, k, ctx->dtype, ctx->max_smem_optin);
  if (want == KM_PATH_AUTO) return small_ok ? KM_PATH_FILTER_F32 : (tensor_ok && k >= 16) ? KM_PATH_TENSOR : KM_PATH_DIRECT_F64;
  return want;
}

static km_status p1_flush(km_ctx *ctx, int override_update, int override_nc);

static km_status install_centroids(km_ctx *ctx, const double *c_host, int32_t k) {
  const int32_t d = ctx->d;
  const KmAccLayout L = km_acc_layout(k, d);
  const int path = pick_path_for(ctx, k);
  if (path == KM_PATH_FILTER_F32 && !km_small_path_supported(d, k, ctx->dtype, ctx->max_smem_optin))
    return fail(ctx, KM_ERR_UNSUPPORTED, "KM_PATH_FILTER_F32 needs float32 points, d <= 8, k <= 4096, k*d <= 8192");
  if (path == KM_PATH_TENSOR && !km_tensor_path_supported(d, k, ctx->dtype, ctx->max_smem_optin))
    return fail(ctx, KM_ERR_UNSUPPORTED, "KM_PATH_TENSOR needs float32 points, 9 <= d <= 128, k >= 2");
  // binary64 brute force stages one centroid row (d doubles) in shared memory: say so here, not as an opaque
  // launch failure at the first assign
  if (path == KM_PATH_DIRECT_F64 && (size_t)d * sizeof(double) + 4096 > (size_t)ctx->max_smem_optin)
    return fail(ctx, KM_ERR_UNSUPPORTED, "KM_PATH_DIRECT_F64 needs d * 8 bytes of shared memory per centroid row: d <= " +
                                             std::to_string((ctx->max_smem_optin - 4096) / 8) + " on this device");
  const bool p1 = path == KM_PATH_FILTER_F32;
  if (!ctx->have_centroids || ctx->k != k || ctx->acc_len != L.len || ctx->p1 != p1 || ctx->path_now != path) {
    free_centroids(ctx);
    ctx->path_now = path;
    size_t kd = (size_t)k * d;
    if (p1) {
      for (int q = 0; q < 2; q++) {
        KM_CUDA(ctx, cudaMalloc(&ctx->p1S[q], (size_t)L.len * sizeof(long long)));
        KM_CUDA(ctx, cudaMalloc(&ctx->p1C[q], kd * sizeof(double)));
      }
      // tables small enough for the one-shot peer all-reduce live in the region the peers have mapped
      ctx->p1D_in_peer = ctx->peer_ok && (size_t)L.len * sizeof(long long) <= KM_PEER_TABLE_BYTES;
      for (int q = 0; q < 3; q++) {
        if (ctx->p1D_in_peer) {
          ctx->p1D[q] = reinterpret_cast<long long *>(static_cast<unsigned char *>(ctx->peer_base[ctx->rank]) + (size_t)q * KM_PEER_TABLE_BYTES);
          KM_CUDA(ctx, cudaMalloc(&ctx->p1R[q], (size_t)L.len * sizeof(long long)));
        } else KM_CUDA(ctx, cudaMalloc(&ctx->p1D[q], (size_t)L.len * sizeof(long long)));
      }
      ctx->p1 = true;
      ctx->p1_s = 0;
    } else {
      KM_CUDA(ctx, cudaMalloc(&ctx->c64, kd * sizeof(double)));
      KM_CUDA(ctx, cudaMalloc(&ctx->chi, kd * sizeof(float)));
      KM_CUDA(ctx, cudaMalloc(&ctx->clo, kd * sizeof(float)));
      KM_CUDA(ctx, cudaMalloc(&ctx->acc, (size_t)L.len * sizeof(long long)));
      if (ctx->world > 1) KM_CUDA(ctx, cudaMalloc(&ctx->acc_global, (size_t)L.len * sizeof(long long)));
      else ctx->acc_global = ctx->acc;
      if (path == KM_PATH_TENSOR) {
        KM_CUDA(ctx, cudaMalloc(&ctx->t_cn, km_tensor_h_bytes(k)));
        KM_CUDA(ctx, cudaMalloc(&ctx->t_cbf, km_tensor_cbf_bytes(d, k)));
        KM_CUDA(ctx, cudaMalloc(&ctx->t_hist, ((size_t)k + 2) * sizeof(unsigned)));   // k + 1 buckets, then the worklist length
        ctx->t_work_count = ctx->t_hist + (k + 1);
      }
    }
    ctx->k = k;
    ctx->acc_len = L.len;
    ctx->have_centroids = true;
    km_status s = reset_labels(ctx);  // k changed: old labels mean nothing
    if (s != KM_OK) return s;
  }
  if (ctx->p1) {
    // merge whatever pass is pending under the old centroids — with the scales that pass was
    // issued with, i.e. BEFORE the bound below changes — then overwrite them
    km_status s = p1_flush(ctx, -1, -1);
    if (s != KM_OK) return s;
  }
  ctx->use_hilo = false;
  {
    // every centroid an update produces is a mean of points: max(max|x|, max|c0|) bounds max|c|
    double cb = ctx->maxabs;
    for (int64_t e = 0; e < (int64_t)k * d; e++) cb = std::max(cb, fabs(c_host[e]));
    ctx->cbound = cb;
  }
  if (ctx->p1) {
    KM_CUDA(ctx, cudaMemcpyAsync(ctx->p1C[ctx->p1_s], c_host, (size_t)k * d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // c_host is borrowed for this call only
    return KM_OK;
  }
  KM_CUDA(ctx, cudaMemcpyAsync(ctx->c64, c_host, (size_t)k * d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // c_host is borrowed for this call only
  // finalize with nothing to consume: rewrites chi/clo from c64 and refreshes the scales
  KM_CUDA(ctx, km_launch_finalize(make_fin(ctx, nullptr, 0, 0, 0), ctx->stream, &ctx->launches));
  return KM_OK;
}

extern "C" km_status km_set_centroids(km_ctx *ctx, const double *c, int32_t k) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "km_set_centroids: upload or generate points first (d is taken from them)");
  if (!c || k < 1 || k > KM_MAX_K) return fail(ctx, KM_ERR_INVALID_ARG, "km_set_centroids: need c != NULL and 1 <= k <= 2^20");
  for (int64_t e = 0; e < (int64_t)k * ctx->d; e++)
    if (!std::isfinite(c[e])) return fail(ctx, KM_ERR_NONFINITE, "centroids contain NaN or Inf");
  return install_centroids(ctx, c, k);
}

// A7: first k picks of a seeded partial Fisher-Yates over the job's indices — the loop of
// /root/reference/app.mjs:163 cut to k swaps, Math.random() replaced by splitmix64.
static void forgy_indices(int64_t n, int32_t k, uint64_t seed, std::vector<int64_t> &out) {
  std::vector<std::pair<int64_t, int64_t>> moved;  // displaced slots only
  auto get = [&](int64_t key) { for (auto &m : moved) if (m.first == key) return m.second; return key; };
  auto set = [&](int64_t key, int64_t val) { for (auto &m : moved) if (m.first == key) { m.second = val; return; } moved.emplace_back(key, val); };
  uint64_t state = seed;
  out.resize(k);
  for (int32_t s = 0; s < k; s++) {
    int64_t i = n - 1 - s;
    uint64_t r = km_splitmix64(state);
    state += KM_GAMMA;
    double u = (double)(r >> 11) * 0x1.0p-53;
    int64_t j = (int64_t)floor(u * (double)(i + 1));
    int64_t vi = get(i), vj = get(j);
    set(i, vj);
    set(j, vi);
    out[s] = vj;
  }
}

extern "C" km_status km_init_forgy(km_ctx *ctx, int32_t k, uint64_t seed) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "km_init_forgy: no points");
  if (k < 1 || k > KM_MAX_K || k > ctx->n_total) return fail(ctx, KM_ERR_INVALID_ARG, "km_init_forgy: need 1 <= k <= N");
  std::vector<int64_t> idx;
  forgy_indices(ctx->n_total, k, seed, idx);
  const int32_t d = ctx->d;
  const size_t row = (size_t)d * elt_size(ctx->dtype);
  std::vector<double> c((size_t)k * d, 0.0);
  std::vector<char> tmp(row);
  for (int32_t s = 0; s < k; s++) {
    int64_t g = idx[s];
    if (g < ctx->i0 || g >= ctx->i0 + ctx->n) continue;  // another rank owns this row
    if (ctx->perm) {
      km_status gs = km_get_points(ctx, tmp.data(), g - ctx->i0, 1);
      if (gs != KM_OK) return gs;
    } else {
      KM_CUDA(ctx, cudaMemcpyAsync(tmp.data(), (const char *)ctx->x + (size_t)(g - ctx->i0) * row, row, cudaMemcpyDeviceToHost, ctx->stream));
      KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    for (int32_t t = 0; t < d; t++)
      c[(size_t)s * d + t] = ctx->dtype == KM_DTYPE_F32 ? (double)((const float *)tmp.data())[t] : ((const double *)tmp.data())[t];
  }
  if (ctx->world > 1) {
    // each row is owned by exactly one rank; the others contribute exact zeros
    double *d_buf = nullptr;
    KM_CUDA(ctx, cudaMalloc(&d_buf, c.size() * sizeof(double)));
    cudaError_t e = cudaMemcpyAsync(d_buf, c.data(), c.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
    int r = 0;
    if (e == cudaSuccess) r = g_nccl.all_reduce(d_buf, d_buf, c.size(), KM_NCCL_FLOAT64, KM_NCCL_SUM, ctx->nccl_comm, ctx->stream);
    if (e == cudaSuccess && r == 0) e = cudaMemcpyAsync(c.data(), d_buf, c.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && r == 0) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_buf);
    if (r != 0) return fail(ctx, KM_ERR_NCCL, std::string("ncclAllReduce: ") + g_nccl.get_error_string(r));
    KM_CUDA(ctx, e);
  }
  return install_centroids(ctx, c.data(), k);
}

// put a cell-sorted store back into the caller's row order (k-means++ walks rows in that order)
static km_status restore_caller_order(km_ctx *ctx) {
  if (!ctx->perm) return KM_OK;
  float *x2 = nullptr;
  int32_t *l2 = nullptr;
  KM_CUDA(ctx, cudaMalloc(&x2, std::max<size_t>((size_t)ctx->n * 2 * sizeof(float), 16)));
  cudaError_t e = cudaMalloc(&l2, std::max<size_t>((size_t)ctx->n * sizeof(int32_t), 16));
  if (e == cudaSuccess) e = km_unpermute_rows2((const float *)ctx->x, ctx->perm, ctx->n, 0, ctx->n, x2, ctx->sm_count, ctx->stream, &ctx->launches);
  if (e == cudaSuccess) e = km_unpermute_labels(ctx->labels, ctx->perm, ctx->n, l2, ctx->sm_count, ctx->stream, &ctx->launches);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { cudaFree(x2); if (l2) cudaFree(l2); KM_CUDA(ctx, e); }
  cudaFree(ctx->x);
  cudaFree(ctx->labels);
  cudaFree(ctx->perm);
  ctx->x = x2; ctx->labels = l2; ctx->perm = nullptr;
  ctx->passes_since_upload = 0;
  return KM_OK;
}

// A7b (oracle: ko_kmeanspp_indices).  Init is not the timed path: one pass over the shard per
// centre, a few small copies and (world > 1) two tiny all-reduces per centre.
extern "C" km_status km_init_kmeanspp(km_ctx *ctx, int32_t k, uint64_t seed) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "km_init_kmeanspp: no points");
  if (k < 1 || k > KM_MAX_K || k > ctx->n_total) return fail(ctx, KM_ERR_INVALID_ARG, "km_init_kmeanspp: need 1 <= k <= N");
  km_status st = restore_caller_order(ctx);
  if (st != KM_OK) return st;
  const int32_t d = ctx->d;
  const int64_t n = ctx->n;
  const size_t row = (size_t)d * elt_size(ctx->dtype);
  const int64_t tiles = (n + KM_PP_TILE - 1) / KM_PP_TILE;
  int dbits = 0;
  while ((1 << dbits) <= d) dbits++;
  const double scale = ldexp(1.0, 62 - ctx->NB - (2 * ctx->E + 2 + dbits));
  double *D = nullptr, *c_dev = nullptr;
  long long *partial = nullptr, *small = nullptr;   // small: [0] = found index, [1 .. world] = per-rank totals
  std::vector<long long> h_partial((size_t)std::max<int64_t>(tiles, 1)), h_small((size_t)ctx->world + 1);
  std::vector<double> c((size_t)k * d, 0.0), crow((size_t)d);
  std::vector<char> tmp(row);
  auto cleanup = [&]() { if (D) cudaFree(D); if (c_dev) cudaFree(c_dev); if (partial) cudaFree(partial); if (small) cudaFree(small); };
#define KM_PP_CUDA(expr)                                                          \
  do {                                                                            \
    cudaError_t _pe = (expr);                                                     \
    if (_pe != cudaSuccess) { cleanup(); KM_CUDA(ctx, _pe); }                     \
  } while (0)
  KM_PP_CUDA(cudaMalloc(&D, std::max<size_t>((size_t)n * sizeof(double), 16)));
  KM_PP_CUDA(cudaMalloc(&c_dev, (size_t)d * sizeof(double)));
  KM_PP_CUDA(cudaMalloc(&partial, (size_t)std::max<int64_t>(tiles, 1) * sizeof(long long)));
  KM_PP_CUDA(cudaMalloc(&small, ((size_t)ctx->world + 1) * sizeof(long long)));
  uint64_t state = seed;
  for (int32_t s = 0; s < k; s++) {
    const uint64_t r = km_splitmix64(state);
    state += KM_GAMMA;
    const double u = (double)(r >> 11) * 0x1.0p-53;
    int64_t pick_local = -1;   // row of this shard that becomes centre s, or -1 if another rank owns it
    bool by_global_index = (s == 0);
    if (s > 0) {
      KM_PP_CUDA(cudaMemcpyAsync(c_dev, c.data() + (size_t)(s - 1) * d, (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
      KM_PP_CUDA(km_launch_pp_update(ctx->x, ctx->dtype, n, d, c_dev, D, scale, s == 1, partial, ctx->stream, &ctx->launches));
      if (tiles > 0)
        KM_PP_CUDA(cudaMemcpyAsync(h_partial.data(), partial, (size_t)tiles * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
      KM_PP_CUDA(cudaStreamSynchronize(ctx->stream));
      long long mine = 0;
      for (int64_t b = 0; b < tiles; b++) mine += h_partial[(size_t)b];
      std::vector<long long> totals((size_t)ctx->world, 0);
      totals[(size_t)ctx->rank] = mine;
      if (ctx->world > 1) {   // every rank fills its own slot: the sum is the list of totals
        KM_PP_CUDA(cudaMemcpyAsync(small + 1, totals.data(), (size_t)ctx->world * sizeof(long long), cudaMemcpyHostToDevice, ctx->stream));
        int rc = g_nccl.all_reduce(small + 1, small + 1, (size_t)ctx->world, KM_NCCL_INT64, KM_NCCL_SUM, ctx->nccl_comm, ctx->stream);
        if (rc != 0) { cleanup(); return fail(ctx, KM_ERR_NCCL, std::string("ncclAllReduce: ") + g_nccl.get_error_string(rc)); }
        KM_PP_CUDA(cudaMemcpyAsync(totals.data(), small + 1, (size_t)ctx->world * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
        KM_PP_CUDA(cudaStreamSynchronize(ctx->stream));
      }
      unsigned __int128 T = 0;
      for (int q = 0; q < ctx->world; q++) T += (unsigned __int128)(unsigned long long)totals[(size_t)q];
      if (T == 0) {
        by_global_index = true;   // every point already sits on a centre
      } else {
        long long target = (long long)(((unsigned __int128)(r >> 11) * T) >> 53);
        int owner = ctx->world - 1;
        for (int q = 0; q < ctx->world; q++) {
          if (target < totals[(size_t)q]) { owner = q; break; }
          target -= totals[(size_t)q];
        }
        if (owner == ctx->rank) {
          int64_t b = 0;
          while (b + 1 < tiles && target >= h_partial[(size_t)b]) { target -= h_partial[(size_t)b]; b++; }
          const int64_t start = b * KM_PP_TILE;
          const int count = (int)std::min<int64_t>(KM_PP_TILE, n - start);
          KM_PP_CUDA(km_launch_pp_find(D, start, count, scale, target, small, ctx->stream, &ctx->launches));
          KM_PP_CUDA(cudaMemcpyAsync(h_small.data(), small, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
          KM_PP_CUDA(cudaStreamSynchronize(ctx->stream));
          pick_local = h_small[0];
        }
      }
    }
    if (by_global_index) {
      const int64_t g = (int64_t)floor(u * (double)ctx->n_total);
      if (g >= ctx->i0 && g < ctx->i0 + n) pick_local = g - ctx->i0;
    }
    for (int32_t t = 0; t < d; t++) crow[(size_t)t] = 0.0;
    if (pick_local >= 0) {
      KM_PP_CUDA(cudaMemcpyAsync(tmp.data(), (const char *)ctx->x + (size_t)pick_local * row, row, cudaMemcpyDeviceToHost, ctx->stream));
      KM_PP_CUDA(cudaStreamSynchronize(ctx->stream));
      for (int32_t t = 0; t < d; t++)
        crow[(size_t)t] = ctx->dtype == KM_DTYPE_F32 ? (double)((const float *)tmp.data())[t] : ((const double *)tmp.data())[t];
    }
    if (ctx->world > 1) {   // exactly one rank owns the row; the others contribute exact zeros
      KM_PP_CUDA(cudaMemcpyAsync(c_dev, crow.data(), (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
      int rc = g_nccl.all_reduce(c_dev, c_dev, (size_t)d, KM_NCCL_FLOAT64, KM_NCCL_SUM, ctx->nccl_comm, ctx->stream);
      if (rc != 0) { cleanup(); return fail(ctx, KM_ERR_NCCL, std::string("ncclAllReduce: ") + g_nccl.get_error_string(rc)); }
      KM_PP_CUDA(cudaMemcpyAsync(crow.data(), c_dev, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
      KM_PP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    for (int32_t t = 0; t < d; t++) c[(size_t)s * d + t] = crow[(size_t)t];
  }
#undef KM_PP_CUDA
  cleanup();
  return install_centroids(ctx, c.data(), k);
}

// ------------------------------------------------------------------ a3/a4/a5: the iteration
static km_status ready(km_ctx *ctx) {
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "no points: call km_upload_points / km_attach_points / km_generate_blobs");
  if (!ctx->have_centroids) return fail(ctx, KM_ERR_STATE, "no centroids: call km_set_centroids / km_init_forgy");
  return KM_OK;
}


static void fill_assign_args(km_ctx *ctx, KmAssignArgs &a) {
  a.x = ctx->x; a.dtype = ctx->dtype; a.n = ctx->n; a.d = ctx->d; a.k = ctx->k;
  a.labels = ctx->labels; a.c64 = ctx->c64; a.chi = ctx->chi; a.clo = ctx->clo;
  a.acc = ctx->acc; a.state = ctx->state;
  a.incremental = ctx->cfg.update_mode == KM_UPDATE_INCREMENTAL;
  a.all_push = (!a.incremental || ctx->labels_fresh) ? 1 : 0;
  a.exact_inertia = ctx->cfg.exact_inertia;
  a.sm_count = ctx->sm_count; a.max_smem_optin = ctx->max_smem_optin; a.stream = ctx->stream;
  a.hilo = ctx->use_hilo ? 1 : 0;
  a.scale_a = ldexp(1.0, ctx->qa);
  a.scale_b = ldexp(1.0, ctx->QB);
  {
    // scales that depend on max|c| (same formulas as km_scales_from_cmax, from the host bound)
    const double M = std::max(ctx->cbound, ctx->maxabs);
    const int EM = M > 0.0 ? ilogb(M) + 1 : 0;
    int dbits = 0;
    while ((1 << dbits) <= ctx->d) dbits++;
    a.h_cbound = M;
    a.h_qi = ctx->QB - (2 * EM + 2 + dbits);
    a.h_inertia_scale = ldexp(1.0, a.h_qi);
    const double c2 = M * M;
    a.h_tau_hilo = (float)(2.5 * (4.2 * ctx->d * 0x1.0p-72 * c2) + ctx->d * 0x1.0p-140) * 1.000001f;
    a.h_tau_hi = (float)(2.5 * (1.01 * ctx->d * 0x1.0p-24 * c2) + ctx->d * 0x1.0p-140) * 1.000001f;
  }
  memset(&a.ps, 0, sizeof(a.ps));
  memset(&a.pr, 0, sizeof(a.pr));
  a.t_cn = nullptr; a.t_cbf = nullptr; a.t_hist = nullptr; a.t_work_count = nullptr; a.t_order = nullptr; a.t_work = nullptr;
  a.t_labels_fresh = 0; a.t_order_reuse = 0; a.t_xnmax = 0.f; a.t_xn = nullptr; a.t_xbf = nullptr;
  a.prune = 0;
  if (ctx->have_bbox && ctx->cfg.no_prune == 0 &&
      km_small_prune_setup(ctx->d, ctx->k, ctx->n, ctx->cfg.prune_min_points, ctx->sm_count, ctx->bbox, ctx->perm ? 1 : 0, &a.pr))
    a.prune = 1;
  a.sorted = (a.prune && ctx->perm && ctx->cells_g && memcmp(&a.pr, &ctx->sort_pr, sizeof(KmPrune2)) == 0) ? 1 : 0;
  a.cells_g = ctx->cells_g;
  a.work_counter = ctx->cells_g ? reinterpret_cast<unsigned *>(ctx->cells_g + (size_t)KM_PRUNE_MAX_GP * KM_PRUNE_MAX_GP) : nullptr;
}

// Store layout for the pruned d = 2 path: once a job has shown it iterates (third pass on the
// same upload, or km_run asked for >= 3), reorder the owned points by candidate-grid cell so
// a warp's points share their candidates (km_sort.cu).  Results do not depend on the order.
static km_status maybe_sort_points(km_ctx *ctx) {
  if (ctx->perm || !ctx->own_x || ctx->d != 2 || ctx->dtype != KM_DTYPE_F32 || !ctx->have_bbox || ctx->cfg.no_prune ||
      ctx->passes_since_upload < 2 || !ctx->p1)
    return KM_OK;
  KmPrune2 pr;
  memset(&pr, 0, sizeof(pr));
  if (!km_small_prune_setup(ctx->d, ctx->k, ctx->n, ctx->cfg.prune_min_points, ctx->sm_count, ctx->bbox, 1, &pr)) return KM_OK;
  float *x = (float *)ctx->x;
  KM_CUDA(ctx, km_sort_points_by_cell(&x, &ctx->labels, ctx->n, pr, &ctx->perm, ctx->sm_count, ctx->stream, &ctx->launches));
  ctx->x = x;
  ctx->sort_pr = pr;
  if (!ctx->unperm_buf) KM_CUDA(ctx, cudaMalloc(&ctx->unperm_buf, std::max<size_t>((size_t)ctx->n * sizeof(int32_t), 16)));
  if (!ctx->cells_g) KM_CUDA(ctx, cudaMalloc(&ctx->cells_g, ((size_t)KM_PRUNE_MAX_GP * KM_PRUNE_MAX_GP + 1) * sizeof(unsigned long long)));   // words + the work counter
  return KM_OK;
}

// ---- general path: assignment kernel, [NCCL all-reduce of the packed table], finalize kernel
static km_status enqueue_pass(km_ctx *ctx, int do_stats, int do_update, int needs_change) {
  KmAssignArgs a;
  fill_assign_args(ctx, a);
  const KmAccLayout L = km_acc_layout(ctx->k, ctx->d);
  if (!a.incremental)  // full mode: sums and counts are rebuilt from nothing every pass
    KM_CUDA(ctx, cudaMemsetAsync(ctx->acc, 0, (size_t)L.scal * sizeof(long long), ctx->stream));
  KmStatSlot *slot = nullptr;
  if (do_stats) {
    ctx->host_it += 1;
    slot = ctx->ring_dev + (ctx->host_it % KM_RING);
  }
  const bool prof = ctx->cfg.profile_kernels != 0 && ctx->prof_n < KM_PROF_RING;
  if (prof) KM_CUDA(ctx, cudaEventRecord(ctx->prof_ev[ctx->prof_n][0], ctx->stream));
  // beyond ~1e15 the binary32 scores of the coarse stage overflow: such a pass runs on the binary64 kernel
  const double mag = sqrt((double)ctx->d) * std::max(ctx->maxabs, ctx->cbound);
  if (ctx->path_now == KM_PATH_TENSOR && mag < 1e15) {
    if (ctx->t_work_n < ctx->n || !ctx->t_work) {
      if (ctx->t_work) cudaFree(ctx->t_work);
      if (ctx->t_order) cudaFree(ctx->t_order);
      if (ctx->t_xn) cudaFree(ctx->t_xn);
      ctx->t_work = nullptr; ctx->t_order = nullptr; ctx->t_xn = nullptr;
      KM_CUDA(ctx, cudaMalloc(&ctx->t_work, km_tensor_work_bytes(ctx->n)));
      KM_CUDA(ctx, cudaMalloc(&ctx->t_order, std::max<size_t>((size_t)ctx->n * sizeof(unsigned), 16)));
      KM_CUDA(ctx, cudaMalloc(&ctx->t_xn, std::max<size_t>((size_t)ctx->n * sizeof(float), 16)));
      ctx->t_work_n = ctx->n;
      ctx->t_xn_valid = false;
      ctx->t_order_valid = false;
    }
    if (ctx->t_xbf_bytes < km_tensor_xbf_bytes(ctx->n, ctx->d)) {
      if (ctx->t_xbf) cudaFree(ctx->t_xbf);
      ctx->t_xbf = nullptr; ctx->t_xbf_bytes = 0;
      KM_CUDA(ctx, cudaMalloc(&ctx->t_xbf, km_tensor_xbf_bytes(ctx->n, ctx->d)));
      ctx->t_xbf_bytes = km_tensor_xbf_bytes(ctx->n, ctx->d);
      ctx->t_xn_valid = false;
    }
    if (!ctx->t_xn_valid) {   // once per upload: the rows' norms and their BF16 copy
      KM_CUDA(ctx, km_launch_row_norms((const float *)ctx->x, ctx->n, ctx->d, ctx->t_xn, ctx->sm_count, ctx->stream, &ctx->launches));
      KM_CUDA(ctx, km_launch_xbf((const float *)ctx->x, ctx->n, ctx->d, ctx->t_xbf, ctx->sm_count, ctx->stream, &ctx->launches));
      ctx->t_xn_valid = true;
    }
    a.t_xn = ctx->t_xn;
    a.t_xbf = ctx->t_xbf;
    a.t_cn = ctx->t_cn; a.t_cbf = ctx->t_cbf; a.t_hist = ctx->t_hist; a.t_work_count = ctx->t_work_count;
    a.t_order = ctx->t_order; a.t_work = ctx->t_work;
    a.t_labels_fresh = ctx->labels_fresh ? 1 : 0;
    // the first passes re-label many points: rebuild the list every time; later the list ages gracefully
    // (a fraction of a percent of the labels change per pass), so the interval grows with the iteration
    // count: every fourth pass at first, every 16th at most
    const int max_age = std::min(16, std::max(3, ctx->host_it / 2));
    a.t_order_reuse = (ctx->t_order_valid && ctx->host_it >= 6 && ctx->t_order_age < max_age) ? 1 : 0;
    if (!a.t_labels_fresh) {
      if (a.t_order_reuse) ctx->t_order_age++;
      else { ctx->t_order_valid = true; ctx->t_order_age = 0; }
    }
    a.t_xnmax = (float)(sqrt((double)ctx->d) * ctx->maxabs * (1.0 + 1e-6));
    KM_CUDA(ctx, km_launch_assign_tensor(a, &ctx->launches));
    ctx->last_path = KM_PATH_TENSOR;
  } else {
    KM_CUDA(ctx, km_launch_assign_direct(a, &ctx->launches));
    ctx->last_path = KM_PATH_DIRECT_F64;
  }
  if (prof) {
    KM_CUDA(ctx, cudaEventRecord(ctx->prof_ev[ctx->prof_n][1], ctx->stream));
    ctx->prof_n++;
  }
  ctx->labels_fresh = false;
  if (ctx->world > 1 && ctx->peer_ok && (size_t)ctx->acc_len * sizeof(long long) <= KM_PEER_SNAP_BYTES) {
    // the sum across ranks over peer memory (km_peer.cu): this rank's running table is copied into one of two
    // alternating snapshots inside the region its peers have mapped (the table itself keeps changing — the next
    // pass, the finalize kernel's scalars — while a slower peer may still be reading), then one kernel does the rest
    const unsigned long long e = ++ctx->peer_epoch2;
    const size_t off = KM_PEER_SNAP_OFF + (size_t)(e & 1) * KM_PEER_SNAP_BYTES;
    KM_CUDA(ctx, cudaMemcpyAsync(static_cast<unsigned char *>(ctx->peer_base[ctx->rank]) + off, ctx->acc,
                                 (size_t)ctx->acc_len * sizeof(long long), cudaMemcpyDeviceToDevice, ctx->stream));
    KM_CUDA(ctx, km_launch_peer_allreduce(ctx->peer_base, ctx->world, ctx->rank, off, KM_PEER_FLAGS2_OFF, ctx->acc_len, e,
                                          ctx->acc_global, ctx->peer_err, ctx->sm_count, ctx->stream, &ctx->launches));
  } else if (ctx->world > 1)
    KM_NCCL(ctx, g_nccl.all_reduce(ctx->acc, ctx->acc_global, (size_t)ctx->acc_len, KM_NCCL_INT64, KM_NCCL_SUM,
                                   ctx->nccl_comm, ctx->stream));
  KM_CUDA(ctx, km_launch_finalize(make_fin(ctx, slot, do_stats, do_update, needs_change), ctx->stream, &ctx->launches));
  if (do_stats) KM_CUDA(ctx, cudaEventRecord(ctx->ring_ev[ctx->host_it % KM_RING], ctx->stream));
  return KM_OK;
}

// ---- small-d path, prologue-finalize protocol.  One launch per pass:
// the kernel's prologue merges the pending pass's deltas into the sums, derives the
// centroids, publishes the pending pass's stats; then it assigns this rank's points and
// leaves its own deltas pending.  `assign` = false runs the prologue alone (n = 0).
static km_status p1_launch(km_ctx *ctx, bool assign, int update, int nc) {
  KmAssignArgs a;
  fill_assign_args(ctx, a);
  if (!assign) a.n = 0;
  KmPass &ps = a.ps;
  const int s = ctx->p1_s;
  ps.S_in = ctx->p1S[s]; ps.S_out = ctx->p1S[s ^ 1];
  ps.C_in = ctx->p1C[s]; ps.C_out = ctx->p1C[s ^ 1];
  // D_in: the job-wide table — the peer all-reduce leaves it in p1R, NCCL (and one rank) in place
  ps.D_in = ctx->p1_pending ? (ctx->p1D_in_peer ? ctx->p1R[ctx->p1_dpend] : ctx->p1D[ctx->p1_dpend]) : nullptr;
  ps.D_out = ctx->p1D[ctx->p1_dw];
  ps.D_zero = assign ? ctx->p1D[(ctx->p1_dw + 1) % 3] : nullptr;
  ps.slot = ctx->p1_pending ? ctx->ring_dev + (ctx->p1_pend_it % KM_RING) : nullptr;
  ps.it_pending = ctx->p1_pend_it;
  ps.full_mode = a.incremental ? 0 : 1;
  ps.apply_update = update;
  ps.needs_change = nc;
  ps.QB = ctx->QB; ps.qa = ctx->qa; ps.maxabs_x = ctx->maxabs;
  ps.Ta = ctx->Ta; ps.Tb = ctx->Tb; ps.qt = ctx->qt; ps.QBt = ctx->QBt;
  inertia_grid(ctx, &ps.qg, &ps.QBg);
  const bool prof = assign && ctx->cfg.profile_kernels != 0 && ctx->prof_n < KM_PROF_RING;
  if (prof) KM_CUDA(ctx, cudaEventRecord(ctx->prof_ev[ctx->prof_n][0], ctx->stream));
  KM_CUDA(ctx, km_launch_assign_small(a, &ctx->launches));
  if (prof) {
    KM_CUDA(ctx, cudaEventRecord(ctx->prof_ev[ctx->prof_n][1], ctx->stream));
    ctx->prof_n++;
  }
  if (ctx->p1_pending) KM_CUDA(ctx, cudaEventRecord(ctx->ring_ev[ctx->p1_pend_it % KM_RING], ctx->stream));
  ctx->p1_s = s ^ 1;
  ctx->p1_pending = false;
  ctx->last_path = KM_PATH_FILTER_F32;
  if (assign) {
    if (ctx->p1D_in_peer) {  // the one collective of an iteration: one small kernel over peer memory
      ctx->peer_epoch++;
      KM_CUDA(ctx, km_launch_peer_allreduce(ctx->peer_base, ctx->world, ctx->rank, (size_t)ctx->p1_dw * KM_PEER_TABLE_BYTES,
                                            KM_PEER_FLAGS_OFF, ctx->acc_len, ctx->peer_epoch, ctx->p1R[ctx->p1_dw], ctx->peer_err,
                                            ctx->sm_count, ctx->stream, &ctx->launches));
    } else if (ctx->world > 1)  // ... or NCCL, in place
      KM_NCCL(ctx, g_nccl.all_reduce(ps.D_out, ps.D_out, (size_t)ctx->acc_len, KM_NCCL_INT64, KM_NCCL_SUM,
                                     ctx->nccl_comm, ctx->stream));
    ctx->host_it += 1;
    ctx->p1_pending = true;
    ctx->p1_dpend = ctx->p1_dw;
    ctx->p1_dw = (ctx->p1_dw + 1) % 3;
    ctx->p1_pend_it = ctx->host_it;
    ctx->labels_fresh = false;
  }
  return KM_OK;
}

// an assignment pass; the pass pending before it is merged with the intent it was issued with
static km_status p1_assign(km_ctx *ctx, int intent_update, int intent_nc) {
  km_status s = maybe_sort_points(ctx);
  if (s != KM_OK) return s;
  ctx->passes_since_upload++;
  s = p1_launch(ctx, true, ctx->p1_pending ? ctx->p1_pend_update : 0, ctx->p1_pending ? ctx->p1_pend_nc : 0);
  ctx->p1_pend_update = intent_update;
  ctx->p1_pend_nc = intent_nc;
  return s;
}

// merge the pending pass now (prologue-only launch); -1 keeps the intent it was issued with
static km_status p1_flush(km_ctx *ctx, int override_update, int override_nc) {
  if (!ctx->p1 || !ctx->p1_pending) return KM_OK;
  return p1_launch(ctx, false, override_update < 0 ? ctx->p1_pend_update : override_update,
                   override_nc < 0 ? ctx->p1_pend_nc : override_nc);
}

static void fill_stats(km_ctx *ctx, const KmStatSlot &s, km_stats *out, float ms) {
  if (!out) return;
  out->iterations = (int32_t)s.it;
  out->converged = s.changed == 0;
  out->changed = s.changed;
  out->inertia = s.inertia;
  out->rescored = s.rescored;
  out->n_total = ctx->n_total;
  out->device_ms = ms;
  out->path_used = ctx->last_path;
}

static km_status wait_slot(km_ctx *ctx, int32_t it, KmStatSlot *out) {
  KM_CUDA(ctx, cudaEventSynchronize(ctx->ring_ev[it % KM_RING]));
  *out = ctx->ring[it % KM_RING];
  if (ctx->peer_err && *reinterpret_cast<volatile int *>(ctx->peer_err))
    return fail(ctx, KM_ERR_NCCL, "peer all-reduce: a rank did not arrive within the time limit");
  if (out->it != it) return fail(ctx, KM_ERR_STATE, "stats ring overrun");
  // filter escalation: hi-only binary32 centroids carry an absolute error ~ d*u*max|c|^2; when
  // that sends more than 2% of the job to the binary64 re-score, switch to hi/lo pairs
  if (!ctx->use_hilo && ctx->n_total > 0 && out->rescored * 50 > ctx->n_total) ctx->use_hilo = true;
  return KM_OK;
}

extern "C" km_status km_assign(km_ctx *ctx, km_stats *stats) {
  KM_BIND(ctx);
  km_status s = ready(ctx);
  if (s != KM_OK) return s;
  KM_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  if (ctx->p1) {
    if ((s = p1_assign(ctx, 0, 0)) != KM_OK) return s;
    if ((s = p1_flush(ctx, -1, -1)) != KM_OK) return s;
  } else if ((s = enqueue_pass(ctx, 1, 0, 0)) != KM_OK) return s;
  KM_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  KmStatSlot slot;
  if ((s = wait_slot(ctx, ctx->host_it, &slot)) != KM_OK) return s;
  KM_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  fill_stats(ctx, slot, stats, ms);
  return KM_OK;
}

extern "C" km_status km_update(km_ctx *ctx, km_stats *stats) {
  KM_BIND(ctx);
  km_status s = ready(ctx);
  if (s != KM_OK) return s;
  if (ctx->labels_fresh) return fail(ctx, KM_ERR_STATE, "km_update: nothing assigned yet");
  KM_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  if (ctx->p1) {
    if (ctx->p1_pending) s = p1_flush(ctx, 1, 0);
    else s = p1_launch(ctx, false, 1, 0);
    if (s != KM_OK) return s;
  } else {
    KM_CUDA(ctx, km_launch_finalize(make_fin(ctx, nullptr, 0, 1, 0), ctx->stream, &ctx->launches));
  }
  KM_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  KM_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  if (stats) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    KmStatSlot slot = ctx->ring[ctx->host_it % KM_RING];
    fill_stats(ctx, slot, stats, ms);
  }
  return KM_OK;
}

extern "C" km_status km_step_async(km_ctx *ctx) {
  KM_BIND(ctx);
  km_status s = ready(ctx);
  if (s != KM_OK) return s;
  return ctx->p1 ? p1_assign(ctx, 1, 1) : enqueue_pass(ctx, 1, 1, 1);
}

extern "C" km_status km_stats_fetch(km_ctx *ctx, km_stats *stats) {
  KM_BIND(ctx);
  if (ctx->host_it == 0) return fail(ctx, KM_ERR_STATE, "km_stats_fetch: no step issued");
  KmStatSlot slot;
  km_status s = p1_flush(ctx, -1, -1);
  if (s != KM_OK) return s;
  s = wait_slot(ctx, ctx->host_it, &slot);
  if (s != KM_OK) return s;
  fill_stats(ctx, slot, stats, 0.f);
  return KM_OK;
}

extern "C" km_status km_step(km_ctx *ctx, km_stats *stats) {
  KM_BIND(ctx);
  km_status s = ready(ctx);
  if (s != KM_OK) return s;
  KM_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  if ((s = km_step_async(ctx)) != KM_OK) return s;
  if ((s = p1_flush(ctx, -1, -1)) != KM_OK) return s;
  KM_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  KmStatSlot slot;
  if ((s = wait_slot(ctx, ctx->host_it, &slot)) != KM_OK) return s;
  KM_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  fill_stats(ctx, slot, stats, ms);
  return KM_OK;
}

extern "C" km_status km_set_update_mode(km_ctx *ctx, int32_t update_mode) {
  KM_BIND(ctx);
  if (update_mode != KM_UPDATE_INCREMENTAL && update_mode != KM_UPDATE_FULL)
    return fail(ctx, KM_ERR_INVALID_ARG, "km_set_update_mode: mode out of range");
  // the two modes keep different things in the tables: start both from nothing
  km_status s = p1_flush(ctx, -1, -1);
  if (s != KM_OK) return s;
  ctx->cfg.update_mode = update_mode;
  if (ctx->have_points && (s = reset_labels(ctx)) != KM_OK) return s;
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return KM_OK;
}

extern "C" km_status km_reset_labels(km_ctx *ctx) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "no points");
  km_status s = reset_labels(ctx);
  if (s != KM_OK) return s;
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return KM_OK;
}

// km_run of a job that fits in shared memory (km_resident.cu): the whole loop in one cooperative launch.
// Entered after reset_labels: labels -1, sums and delta tables zero, no pass pending.
static km_status run_resident(km_ctx *ctx, int32_t max_iter, km_stats *stats, int per_block, int grid, size_t smem) {
  KmResidentArgs a;
  a.x = (const float *)ctx->x; a.labels = ctx->labels; a.n = ctx->n; a.d = ctx->d; a.k = ctx->k;
  a.per_block = per_block; a.grid = grid; a.smem = smem; a.max_iter = max_iter;
  a.C_in = ctx->p1C[ctx->p1_s]; a.C_out = ctx->p1C[ctx->p1_s]; a.S_out = ctx->p1S[ctx->p1_s];
  for (int q = 0; q < 3; q++) a.Dt[q] = ctx->p1D[q];
  a.ring = ctx->ring_dev; a.state_it = &ctx->state->it;
  a.scale_a = ldexp(1.0, ctx->qa); a.scale_b = ldexp(1.0, ctx->QB); a.QB = ctx->QB; a.qa = ctx->qa;
  a.Ta = ctx->Ta; a.Tb = ctx->Tb; a.qt = ctx->qt; a.QBt = ctx->QBt;
  inertia_grid(ctx, &a.qg, &a.QBg);
  {
    KmAssignArgs f;   // the filter's error terms, from the same host-side bound as the stepwise kernels'
    fill_assign_args(ctx, f);
    a.tau2 = f.h_tau_hi;
    a.rho2 = 2.5f * (float)(ctx->d + 6) * 0x1.0p-24f;
  }
  KM_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  KM_CUDA(ctx, km_launch_resident(a, ctx->stream, &ctx->launches));   // leaves the three delta tables empty again
  KM_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  KM_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  const KmStatSlot last = ctx->ring[0];
  if (last.it < 1 || last.it > max_iter) return fail(ctx, KM_ERR_STATE, "resident run: no result");
  ctx->host_it = (int32_t)last.it;
  ctx->p1_pend_it = ctx->host_it;
  ctx->p1_pending = false;
  ctx->p1_dw = 0;
  ctx->labels_fresh = false;
  ctx->last_path = KM_PATH_FILTER_F32;
  ctx->passes_since_upload += ctx->host_it;
  ctx->ring[ctx->host_it % KM_RING] = last;
  fill_stats(ctx, last, stats, ms);
  return KM_OK;
}

// A5.  The device decides by itself whether an iteration's update happens (changed > 0),
// so the host may keep KM_RUN_LAG iterations in flight: passes issued after convergence
// change nothing, and the first slot with changed == 0 is the answer.
extern "C" km_status km_run(km_ctx *ctx, int32_t max_iter, km_stats *stats) {
  KM_BIND(ctx);
  km_status s = ready(ctx);
  if (s != KM_OK) return s;
  if (max_iter < 1) return fail(ctx, KM_ERR_INVALID_ARG, "km_run: max_iter must be >= 1");
  if ((s = reset_labels(ctx)) != KM_OK) return s;
  if (ctx->p1 && ctx->world == 1 && ctx->coop_ok && ctx->cfg.update_mode == KM_UPDATE_INCREMENTAL && !ctx->cfg.profile_kernels) {
    int per_block = 0, grid = 0;
    size_t smem = 0;
    if (km_resident_plan(ctx->n, ctx->d, ctx->k, ctx->dtype, ctx->sm_count, ctx->max_smem_optin, &per_block, &grid, &smem))
      return run_resident(ctx, max_iter, stats, per_block, grid, smem);
  }
  if (max_iter >= 3 && ctx->passes_since_upload < 2) ctx->passes_since_upload = 2;  // an iterating job: lay the store out for it now
  if ((s = maybe_sort_points(ctx)) != KM_OK) return s;
  KM_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  int32_t issued = 0, checked = 0;
  KmStatSlot last;
  memset(&last, 0, sizeof(last));
  bool done = false;
  while (!done) {
    while (issued < max_iter && issued - checked < KM_RUN_LAG + 1) {
      issued++;
      // the update after the max_iter-th assign never happens (A5)
      s = ctx->p1 ? p1_assign(ctx, issued < max_iter, 1) : enqueue_pass(ctx, 1, issued < max_iter, 1);
      if (s != KM_OK) return s;
    }
    // small-d path: pass t's stats are published by pass t+1's prologue; the newest pass
    // needs an explicit merge
    if (ctx->p1 && checked + 1 == issued && (s = p1_flush(ctx, -1, -1)) != KM_OK) return s;
    if ((s = wait_slot(ctx, checked + 1, &last)) != KM_OK) return s;
    checked++;
    if (last.changed == 0 || checked == max_iter) done = true;
  }
  if ((s = p1_flush(ctx, -1, -1)) != KM_OK) return s;  // speculative passes past the answer change nothing
  KM_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  KM_CUDA(ctx, cudaEventSynchronize(ctx->ev1));  // drains the speculative passes too
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  // speculative passes ran past the answer: rewind the iteration counters to it
  ctx->host_it = checked;
  ctx->p1_pend_it = checked;
  KM_CUDA(ctx, cudaMemcpyAsync(&ctx->state->it, &ctx->host_it, sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->ring[checked % KM_RING] = last;
  fill_stats(ctx, last, stats, ms);
  return KM_OK;
}

// ------------------------------------------------------------------ results
extern "C" km_status km_get_labels(km_ctx *ctx, int32_t *out, int64_t n) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "no points");
  if (n != ctx->n || (n > 0 && !out)) return fail(ctx, KM_ERR_INVALID_ARG, "km_get_labels: n must equal the shard size");
  if (ctx->perm && n > 0) {
    // the cell-sorted store keeps one n-label buffer for this (allocated with the sort): no allocation here
    if (!ctx->unperm_buf) KM_CUDA(ctx, cudaMalloc(&ctx->unperm_buf, (size_t)n * sizeof(int32_t)));
    int32_t *tmp = reinterpret_cast<int32_t *>(ctx->unperm_buf);
    KM_CUDA(ctx, km_unpermute_labels(ctx->labels, ctx->perm, n, tmp, ctx->sm_count, ctx->stream, &ctx->launches));
    return copy_d2h(ctx, out, tmp, (size_t)n * sizeof(int32_t));
  }
  return copy_d2h(ctx, out, ctx->labels, (size_t)n * sizeof(int32_t));
}

extern "C" km_status km_get_labels_u8(km_ctx *ctx, uint8_t *out, int64_t n) {
  KM_BIND(ctx);
  if (!ctx->have_points) return fail(ctx, KM_ERR_STATE, "no points");
  if (n != ctx->n || (n > 0 && !out)) return fail(ctx, KM_ERR_INVALID_ARG, "km_get_labels_u8: n must equal the shard size");
  if (ctx->have_centroids && ctx->k > 255) return fail(ctx, KM_ERR_UNSUPPORTED, "km_get_labels_u8: k must be <= 255");
  if (n == 0) return KM_OK;
  if (ctx->narrow_n < n) {
    if (ctx->narrow_buf) cudaFree(ctx->narrow_buf);
    ctx->narrow_buf = nullptr; ctx->narrow_n = 0;
    KM_CUDA(ctx, cudaMalloc(&ctx->narrow_buf, (size_t)n));
    ctx->narrow_n = n;
  }
  const int32_t *src = ctx->labels;
  if (ctx->perm) {
    // cell-sorted store: back to the caller's order as int32 first (a scatter of whole words; scattering single bytes
    // into shared sectors is several times slower), then narrow with coalesced reads and writes
    if (!ctx->unperm_buf) KM_CUDA(ctx, cudaMalloc(&ctx->unperm_buf, (size_t)n * sizeof(int32_t)));
    int32_t *tmp = reinterpret_cast<int32_t *>(ctx->unperm_buf);
    KM_CUDA(ctx, km_unpermute_labels(ctx->labels, ctx->perm, n, tmp, ctx->sm_count, ctx->stream, &ctx->launches));
    src = tmp;
  }
  KM_CUDA(ctx, km_narrow_labels(src, nullptr, n, ctx->narrow_buf, ctx->sm_count, ctx->stream, &ctx->launches));
  return copy_d2h(ctx, out, ctx->narrow_buf, (size_t)n);
}

extern "C" km_status km_get_centroids(km_ctx *ctx, double *out) {
  KM_BIND(ctx);
  if (!ctx->have_centroids || !out) return fail(ctx, KM_ERR_STATE, "no centroids");
  km_status s = p1_flush(ctx, -1, -1);
  if (s != KM_OK) return s;
  return copy_d2h(ctx, out, ctx->p1 ? ctx->p1C[ctx->p1_s] : ctx->c64, (size_t)ctx->k * ctx->d * sizeof(double));
}

static km_status fetch_acc(km_ctx *ctx, std::vector<long long> &h) {
  if (!ctx->have_centroids) return fail(ctx, KM_ERR_STATE, "no centroids");
  h.resize((size_t)ctx->acc_len);
  km_status s = p1_flush(ctx, -1, -1);
  if (s != KM_OK) return s;
  // whole-job sums as of the last assign (the all-reduced copy when world > 1)
  return copy_d2h(ctx, h.data(), ctx->p1 ? ctx->p1S[ctx->p1_s] : ctx->acc_global, h.size() * sizeof(long long));
}

extern "C" km_status km_get_counts(km_ctx *ctx, int64_t *out) {
  KM_BIND(ctx);
  if (!out) return fail(ctx, KM_ERR_INVALID_ARG, "out is NULL");
  std::vector<long long> h;
  km_status s = fetch_acc(ctx, h);
  if (s != KM_OK) return s;
  const KmAccLayout L = km_acc_layout(ctx->k, ctx->d);
  for (int32_t j = 0; j < ctx->k; j++) out[j] = h[(size_t)L.cnt + j];
  return KM_OK;
}

// (Sa * 2^QB + Sb) * 2^-shift rounded once to binary64, ties to even (host twin of km_fixed_to_double)
static double host_fixed_to_double(long long Sa, long long Sb, int QB, int shift) {
  const __int128 T = (__int128)Sa * ((__int128)1 << QB) + (__int128)Sb;
  if (T == 0) return 0.0;
  const bool neg = T < 0;
  const unsigned __int128 m = neg ? (unsigned __int128)(-T) : (unsigned __int128)T;
  int hb = 127;
  while (!((m >> hb) & 1)) hb--;
  double v;
  if (hb <= 52) {
    v = ldexp((double)(unsigned long long)m, -shift);
  } else {
    const int sh = hb - 52;
    unsigned __int128 q = m >> sh;
    const unsigned __int128 rem = m & ((((unsigned __int128)1) << sh) - 1), half = ((unsigned __int128)1) << (sh - 1);
    if (rem > half || (rem == half && (q & 1))) q++;
    v = ldexp((double)(unsigned long long)q, sh - shift);
  }
  return neg ? -v : v;
}

// f2: the numeric stand-in for the dashboard's per-cluster cohesion (/root/reference/app.mjs:462-475,
// 485-487): how tight each cluster is — the inertia of its members about its centroid.
extern "C" km_status km_get_cluster_inertia(km_ctx *ctx, double *out) {
  KM_BIND(ctx);
  km_status s = ready(ctx);
  if (s != KM_OK) return s;
  if (!out) return fail(ctx, KM_ERR_INVALID_ARG, "out is NULL");
  if (ctx->labels_fresh) return fail(ctx, KM_ERR_STATE, "km_get_cluster_inertia: nothing assigned yet");
  if ((s = p1_flush(ctx, -1, -1)) != KM_OK) return s;
  KmAssignArgs a;
  fill_assign_args(ctx, a);   // for the inertia scale (host bound on max|c|)
  const int32_t k = ctx->k;
  long long *table = nullptr;
  KM_CUDA(ctx, cudaMalloc(&table, (size_t)2 * k * sizeof(long long)));
  std::vector<long long> h((size_t)2 * k);
  cudaError_t e = km_launch_cluster_inertia(ctx->x, ctx->dtype, ctx->labels, ctx->n, ctx->d, k,
                                            ctx->p1 ? ctx->p1C[ctx->p1_s] : ctx->c64, a.h_inertia_scale, a.scale_b, table,
                                            ctx->sm_count, ctx->stream, &ctx->launches);
  int rc = 0;
  if (e == cudaSuccess && ctx->world > 1)
    rc = g_nccl.all_reduce(table, table, (size_t)2 * k, KM_NCCL_INT64, KM_NCCL_SUM, ctx->nccl_comm, ctx->stream);
  if (e == cudaSuccess && rc == 0) e = cudaMemcpyAsync(h.data(), table, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess && rc == 0) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(table);
  if (rc != 0) return fail(ctx, KM_ERR_NCCL, std::string("ncclAllReduce: ") + g_nccl.get_error_string(rc));
  KM_CUDA(ctx, e);
  for (int32_t j = 0; j < k; j++)
    out[j] = host_fixed_to_double(h[(size_t)j], h[(size_t)k + j], ctx->QB, a.h_qi + ctx->QB);
  return KM_OK;
}

extern "C" km_status km_get_balance(km_ctx *ctx, int64_t *max_out, int64_t *min_out, int64_t *gap_out, double *ratio_out) {
  KM_BIND(ctx);
  std::vector<long long> h;
  km_status s = fetch_acc(ctx, h);
  if (s != KM_OK) return s;
  const KmAccLayout L = km_acc_layout(ctx->k, ctx->d);
  long long mx = h[(size_t)L.cnt], mn = h[(size_t)L.cnt];
  for (int32_t j = 1; j < ctx->k; j++) { mx = std::max(mx, h[(size_t)L.cnt + j]); mn = std::min(mn, h[(size_t)L.cnt + j]); }
  if (max_out) *max_out = mx;
  if (min_out) *min_out = mn;
  if (gap_out) *gap_out = mx - mn;
  // /root/reference/app.mjs:493 — ratio = min ? max/min : (max ? Infinity : 1)
  if (ratio_out) *ratio_out = mn ? (double)mx / (double)mn : (mx ? INFINITY : 1.0);
  return KM_OK;
}

extern "C" km_status km_get_sums(km_ctx *ctx, int64_t *sa, int64_t *sb, int32_t params4[4]) {
  KM_BIND(ctx);
  std::vector<long long> h;
  km_status s = fetch_acc(ctx, h);
  if (s != KM_OK) return s;
  const KmAccLayout L = km_acc_layout(ctx->k, ctx->d);
  const size_t kd = (size_t)ctx->k * ctx->d;
  if (sa) for (size_t e = 0; e < kd; e++) sa[e] = h[(size_t)L.sa + e];
  if (sb) for (size_t e = 0; e < kd; e++) sb[e] = h[(size_t)L.sb + e];
  if (params4) { params4[0] = ctx->E; params4[1] = ctx->NB; params4[2] = ctx->QB; params4[3] = ctx->qa; }
  return KM_OK;
}

extern "C" km_status km_get_shape(const km_ctx *ctx, int64_t *n_local, int64_t *n_total, int32_t *d, int32_t *k) {
  if (!ctx) return KM_ERR_INVALID_ARG;
  if (n_local) *n_local = ctx->have_points ? ctx->n : 0;
  if (n_total) *n_total = ctx->have_points ? ctx->n_total : 0;
  if (d) *d = ctx->have_points ? ctx->d : 0;
  if (k) *k = ctx->have_centroids ? ctx->k : 0;
  return KM_OK;
}

extern "C" km_status km_sync(km_ctx *ctx) {
  KM_BIND(ctx);
  KM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return KM_OK;
}

extern "C" km_status km_kernel_time(km_ctx *ctx, double *ms_sum, int32_t *launches, int32_t reset) {
  KM_BIND(ctx);
  double sum = 0.0;
  for (int q = 0; q < ctx->prof_n; q++) {
    KM_CUDA(ctx, cudaEventSynchronize(ctx->prof_ev[q][1]));
    float ms = 0.f;
    KM_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->prof_ev[q][0], ctx->prof_ev[q][1]));
    sum += ms;
  }
  if (ms_sum) *ms_sum = sum;
  if (launches) *launches = ctx->prof_n;
  if (reset) ctx->prof_n = 0;
  return KM_OK;
}

extern "C" km_status km_set_profile(km_ctx *ctx, int32_t on) {
  KM_BIND(ctx);
  ctx->cfg.profile_kernels = on ? 1 : 0;
  return KM_OK;
}

extern "C" int64_t km_launch_count(const km_ctx *ctx) { return ctx ? ctx->launches : 0; }
