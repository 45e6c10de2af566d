/*
 * kmeans.h — C ABI of libkmeans.so: the Lloyd iteration (initialisation, assignment,
 * centroid update, run-to-convergence) on NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for the one path BASELINE.json's north_star names.  The
 * reference (schusto/k-means-demo) has no FFI, no `export` and no numerical Lloyd code
 * (SURVEY.md §0 F1, §8b); each entry point below therefore cites the *manual analog* in
 * /root/reference/app.mjs that it stands in for, and the semantics are SURVEY.md
 * Appendix A ("OUR SPEC").  A maintainer binds these from the JS side through an N-API
 * addon (INTEGRATION.md shows the stub); every buffer is a plain pointer + size that maps
 * 1:1 onto a Float32Array / Float64Array / Int32Array / BigInt64Array backing store.
 *
 * Conventions
 *   - every call returns km_status (0 = OK, negative = error); text via km_last_error().
 *     Nothing throws or aborts across the boundary.
 *   - host pointers passed in are borrowed for the duration of the call only; results are
 *     copied into caller-allocated buffers.  The library owns all device and pinned memory.
 *   - one km_ctx = one GPU = one owner thread at a time.  Multi-GPU = one ctx per process
 *     per GPU, joined with km_comm_init() (NCCL over NVLink/NVSwitch).
 *   - there is no CPU fallback: without a usable CUDA device km_create() fails.
 */
#ifndef KMEANS_B200_H
#define KMEANS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KM_ABI_VERSION 4

#if defined(__GNUC__)
#define KM_API __attribute__((visibility("default")))
#else
#define KM_API
#endif

typedef struct km_ctx km_ctx;
typedef int32_t km_status;

enum {
  KM_OK = 0,
  KM_ERR_INVALID_ARG = -1,
  KM_ERR_CUDA = -2,
  KM_ERR_NCCL = -3,
  KM_ERR_OOM = -4,
  KM_ERR_STATE = -5,      /* call made in the wrong order (no points, no centroids, ...) */
  KM_ERR_NONFINITE = -6,  /* NaN / Inf among the points or centroids */
  KM_ERR_UNSUPPORTED = -7
};

/* storage type of the points handed to km_upload_points / km_attach_points */
enum { KM_DTYPE_F32 = 0, KM_DTYPE_F64 = 1 };

/* which assignment kernel family runs (km_config.path) */
enum {
  KM_PATH_AUTO = 0,       /* pick by (d, k) */
  KM_PATH_DIRECT_F64 = 1, /* every distance in binary64, the oracle's exact op order (d up to ~28 000:
                             one centroid row must fit shared memory; km_set_centroids says so otherwise) */
  KM_PATH_FILTER_F32 = 2, /* binary32 distances + error-bound test + binary64 re-score */
  KM_PATH_TENSOR = 3      /* tcgen05 GEMM-form distances + error-bound test + re-score */
};

/* how the per-cluster sums are maintained (results are bit-identical either way) */
enum {
  KM_UPDATE_INCREMENTAL = 0, /* only points whose label changed move between sums */
  KM_UPDATE_FULL = 1         /* every point is re-accumulated every iteration */
};

typedef struct km_config {
  int32_t struct_size; /* = sizeof(km_config); ABI guard */
  int32_t device;      /* CUDA ordinal */
  int32_t path;        /* KM_PATH_* */
  int32_t update_mode; /* KM_UPDATE_* */
  int32_t exact_inertia; /* kept for layout compatibility and ignored: the inertia always comes from the
                            order-free sums (exact to ~1e-15 relative, identical at any GPU count) */
  int32_t profile_kernels; /* 1: bracket each assignment kernel with CUDA events (km_kernel_time) */
  void *stream;        /* cudaStream_t to run on, or NULL for a private stream */
  int32_t no_prune;    /* 1: never use the d = 2 candidate grid (brute force over all k) */
  int32_t prune_min_points; /* 0: default threshold; else prune once a rank holds this many points */
} km_config;

typedef struct km_stats {
  int32_t iterations;  /* assign steps executed since the labels were last reset */
  int32_t converged;   /* 1 when the last assign changed no label */
  int64_t changed;     /* labels changed by the last assign (whole job, all ranks) */
  double inertia;      /* sum of winner distances of the last assign (whole job) */
  int64_t rescored;    /* points the last assign re-scored in binary64 (whole job) */
  int64_t n_total;     /* points in the whole job */
  float device_ms;     /* device time of this call, CUDA events on the ctx stream */
  int32_t path_used;   /* KM_PATH_* the last assign actually ran */
} km_stats;

/* ---- life cycle ------------------------------------------------------------------ */
KM_API int32_t km_abi_version(void);
/* replaces: nothing — the reference keeps its state in module-scope Yjs globals
 * (app.mjs:29-33).  A ctx is that state made explicit. */
KM_API km_status km_create(const km_config *cfg, km_ctx **out);
KM_API km_status km_destroy(km_ctx *ctx);
/* last error text of `ctx` (or of the calling thread's last failed km_create if NULL) */
KM_API const char *km_last_error(const km_ctx *ctx);

/* page-locked host memory for callers that want km_upload_points / km_get_* to copy
 * straight from / into their buffer (an N-API host wraps it in an external ArrayBuffer) */
KM_API km_status km_host_alloc(void **out, uint64_t bytes);
KM_API km_status km_host_free(void *p);

/* ---- multi-GPU (one process per GPU) -------------------------------- */
/* 128-byte NCCL unique id: rank 0 produces it, the host layer broadcasts it */
KM_API km_status km_comm_unique_id(void *out128);
/* join a communicator of `world` ranks.  Must precede km_upload_points/km_generate_blobs.
 * From then on km_upload_points / km_generate_blobs / km_attach_points, the km_init_* calls, every
 * assign / step / run and km_get_cluster_inertia are COLLECTIVE: every rank must make the same
 * calls in the same order.  A rank whose call fails locally (CUDA error, out of memory) returns
 * its error before joining the collective, and the other ranks then wait in NCCL: treat any
 * non-zero status on one rank as fatal for the job and tear all ranks down.
 * replaces: the P2PT/WebRTC gossip of UI state (app.mjs:70-118) only in the sense that it
 * is the reference's sole inter-process link; it is not a compute collective there. */
KM_API km_status km_comm_init(km_ctx *ctx, const void *id128, int32_t rank, int32_t world);
/* how the per-iteration sum travels.  km_comm_init also tries to map one small region of every
 * peer (CUDA IPC; all ranks on one node with peer access).  Where that worked on every rank, the
 * per-iteration table (2kd + k + 4 int64) is summed by one small kernel over peer memory instead of an
 * NCCL all-reduce (same integers, same bits) — up to 64 KB on the small-d path, up to 896 KB on the
 * direct and tensor paths; larger tables stay on NCCL.  With it, km_set_centroids,
 * km_reset_labels and km_set_update_mode are collective too (they end with a barrier across the
 * ranks), and a rank that waits more than two minutes for a peer returns KM_ERR_NCCL.
 * *peer_allreduce: 0 = NCCL only, 1 = peer regions mapped, 2 = the tables now installed use them.
 * Any pointer may be NULL.  No counterpart in the reference. */
KM_API km_status km_comm_info(const km_ctx *ctx, int32_t *rank, int32_t *world, int32_t *peer_allreduce);

/* ---- a1: point / centroid store (analog: cards app.mjs:31,143-145; centroids :32,128) */
/* copy this rank's shard (n x d, row-major) from host memory through pinned staging.
 * Also scans for NaN/Inf and max|x| (needed by the order-free sums). */
KM_API km_status km_upload_points(km_ctx *ctx, const void *host, int64_t n, int32_t d, int32_t dtype);
/* use points already resident in device memory (borrowed; must outlive the ctx's use) */
KM_API km_status km_attach_points(km_ctx *ctx, const void *device, int64_t n, int32_t d, int32_t dtype);
/* device-side seeded blob generator: rows [i0, i0+n) of the job's matrix */
KM_API km_status km_generate_blobs(km_ctx *ctx, int64_t n, int64_t i0, int32_t d, int32_t k_true,
                            uint64_t seed, double sigma);
/* copy rows [i0, i0+n) of this rank's shard back to the host (float32 or float64 as stored) */
KM_API km_status km_get_points(km_ctx *ctx, void *host, int64_t i0, int64_t n);

/* ---- a2: initialisation (analog: addCentroid app.mjs:126-129; shuffle idiom :163) --- */
KM_API km_status km_set_centroids(km_ctx *ctx, const double *c, int32_t k);
/* seeded Forgy: first k picks of a partial Fisher-Yates over the job's point indices */
KM_API km_status km_init_forgy(km_ctx *ctx, int32_t k, uint64_t seed);
/* seeded k-means++: first centre = point floor(u*N), every further one drawn
 * with probability proportional to its squared distance to the nearest centre so far.  The
 * distances are quantised to integers before they are summed, so the draw does not depend on
 * how the points are split over blocks or GPUs (oracle: ko_kmeanspp_indices).  No counterpart in
 * the reference beyond its seeded-draw idiom (app.mjs:163). */
KM_API km_status km_init_kmeanspp(km_ctx *ctx, int32_t k, uint64_t seed);

/* ---- a3: assignment step (analog: updateCardAssign app.mjs:146-156, drop :358-372) -- */
/* label[i] = argmin_j sum_t (x_it - c_jt)^2, strict <, lowest j wins ties; also counts
 * changed labels and moves each re-labelled point between the per-cluster sums (the update's
 * partial sums are fused into this pass); the sum of the winner distances (km_stats.inertia)
 * follows from those sums.  Kernel family by shape (KM_PATH_AUTO): d <= 8 the binary32 filter,
 * 9 <= d <= 128 with k >= 16 the tcgen05 tensor path, anything else binary64 brute force. */
KM_API km_status km_assign(km_ctx *ctx, km_stats *stats);

/* ---- a4: centroid-update step (analog: traitCountsFor/suggestionFromCounts/
 *          applySuggestedName app.mjs:450-461,476-480,571-573; counts :485-486) -------- */
/* c_j = sum_j / count_j; an empty cluster keeps its centroid (app.mjs:322,484-488) */
KM_API km_status km_update(km_ctx *ctx, km_stats *stats);

/* one Lloyd iteration: assign; update unless nothing changed */
KM_API km_status km_step(km_ctx *ctx, km_stats *stats);
/* same, enqueued without waiting for the device; read results with km_stats_fetch */
KM_API km_status km_step_async(km_ctx *ctx);
KM_API km_status km_stats_fetch(km_ctx *ctx, km_stats *stats);

/* ---- a5: run to convergence (analog: #iter -> yMeta.iteration app.mjs:286-288,498-508) */
/* labels reset to -1; repeat { assign; it++; if (changed==0 || it==max_iter) break; update }.
 * A single-rank job on the small-d path whose points fit in shared memory (n * (4d + 4) bytes over the
 * SMs, k * d <= 32, KM_UPDATE_INCREMENTAL, profile_kernels off) runs the whole loop in ONE cooperative
 * launch with the points resident on the chip; every other job is issued iteration by iteration.  Same
 * results either way. */
KM_API km_status km_run(km_ctx *ctx, int32_t max_iter, km_stats *stats);
/* forget labels and sums (the next assign sees every label as changed) */
KM_API km_status km_reset_labels(km_ctx *ctx);
/* switch km_config.update_mode for the iterations that follow; labels and sums are reset, as
 * by km_reset_labels (the two modes keep different things in the tables).  No counterpart in
 * the reference: both modes produce the same bits, the switch exists to measure one against
 * the other on resident points. */
KM_API km_status km_set_update_mode(km_ctx *ctx, int32_t update_mode);

/* ---- results ----------------------------------------------------------------------- */
/* sizes a binding needs to check a caller's output buffers before the copies below write into
 * them: rows of this rank's shard, rows of the whole job, d, k (0 while not yet known; any
 * pointer may be NULL).  No counterpart in the reference. */
KM_API km_status km_get_shape(const km_ctx *ctx, int64_t *n_local, int64_t *n_total, int32_t *d, int32_t *k);
KM_API km_status km_get_labels(km_ctx *ctx, int32_t *out, int64_t n);   /* this rank's shard */
/* the same labels as bytes, for k <= 255 (255 = never assigned): a quarter of the device-to-host
 * traffic of km_get_labels.  The first call per shard size allocates an n-byte device buffer.
 * No counterpart in the reference (card.cluster is a JS number, app.mjs:143-156). */
KM_API km_status km_get_labels_u8(km_ctx *ctx, uint8_t *out, int64_t n);
KM_API km_status km_get_centroids(km_ctx *ctx, double *out);            /* k*d, row-major */
/* a6: per-cluster member counts of the whole job (analog: snapshotMetrics counts
 * app.mjs:481-496) */
KM_API km_status km_get_counts(km_ctx *ctx, int64_t *out);
/* a6: balance{max,min,gap,ratio} with the reference's formulas (app.mjs:489-493):
 * ratio = min ? max/min : (max ? Infinity : 1) */
KM_API km_status km_get_balance(km_ctx *ctx, int64_t *max_out, int64_t *min_out, int64_t *gap_out,
                         double *ratio_out);
/* per-cluster inertia: out[j] = sum over the members of cluster j of their squared distance to
 * centroid j, for the current labels and centroids — the numeric stand-in for the dashboard's
 * per-cluster cohesion (app.mjs:462-475, 485-487); sums to km_stats.inertia of the last assign */
KM_API km_status km_get_cluster_inertia(km_ctx *ctx, double *out /* k */);
/* raw order-free sums of the whole job: k*d `a` limbs, k*d `b` limbs, then the fixed-point
 * parameters {E, NB, QB, qa} (for bit-level parity tests) */
KM_API km_status km_get_sums(km_ctx *ctx, int64_t *sa, int64_t *sb, int32_t params4[4]);

/* ---- for hosts that may not allocate --------------------------------------------------- */
/* No counterpart in the reference (nor in any reference interface): an addition of this
 * library for callers that are not allowed to touch the device allocator once they are
 * running — a process whose device memory is owned by its own pool, a control loop that
 * cannot take the device-wide stall of a cudaMalloc/cudaFree, a sandboxed plugin capped at
 * the footprint it declared.  One Lloyd iteration on buffers the CALLER owns: assign every
 * point of x (n x d float32, device) to its nearest of centroids_in (k x d float64, device;
 * same distance and tie rules as km_assign's binary64 path), write the labels (in/out:
 * previous labels or -1), the new centroids (cluster means; an empty cluster keeps its
 * centroid; may alias centroids_in), the member counts and {changed, inertia} — all device
 * pointers, enqueued on `stream` (a cudaStream_t, NULL = default stream) without waiting.
 * The library's own working memory for this call is a fixed 256 KiB array built into the
 * library image: it is held by every GPU that loads libkmeans.so until the process exits,
 * it limits the shape to k*(2d+1)+4 <= 32768 (larger shapes fail with KM_ERR_UNSUPPORTED),
 * and it is shared by every call on that GPU — the call clears what it needs first, and two
 * calls whose GPU work overlaps would corrupt each other, so issue them one after another.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress);
 * make that first call before any stream capture.  Calls no allocating routine.  The first
 * call on a device is also when the CUDA runtime itself may still allocate (primary context,
 * lazy loading of this library's kernels and of the array): a host with a no-allocation phase
 * makes one warm-up call per device at start-up and enters that phase afterwards.  Runs on
 * `device` and leaves the caller's current device as it found it.  result->status is
 * KM_ERR_NONFINITE when a point or a centroid was NaN/Inf. */
typedef struct km_noalloc_result {
  int64_t changed;  /* labels this call changed */
  double inertia;   /* sum of winner distances */
  int64_t n;
  int32_t status;   /* KM_OK, or KM_ERR_NONFINITE if a point or centroid was NaN/Inf (outputs then undefined) */
  int32_t pad;
} km_noalloc_result;
KM_API km_status km_lloyd_step_noalloc(int32_t device, void *stream, const float *x, int64_t n, int32_t d,
                                       const double *centroids_in, int32_t k, int32_t *labels,
                                       double *centroids_out, int64_t *counts_out, km_noalloc_result *result);

KM_API km_status km_sync(km_ctx *ctx);
/* with km_config.profile_kernels: summed device time of the assignment kernels launched
 * since the last reset (up to 256 are kept), and how many there were */
KM_API km_status km_kernel_time(km_ctx *ctx, double *ms_sum, int32_t *launches, int32_t reset);
/* switch km_config.profile_kernels on or off for the calls that follow */
KM_API km_status km_set_profile(km_ctx *ctx, int32_t on);
/* number of kernels this ctx has launched since creation (bench.py's gpu_launches) */
KM_API int64_t km_launch_count(const km_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* KMEANS_B200_H */
