// km_internal.h — host-side context and kernel launcher prototypes (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

#include "../../include/kmeans.h"
#include "km_common.cuh"
#include "km_finalize.cuh"

// peer all-reduce region: three rotating tables, then one epoch flag per rank
#define KM_PEER_MAX_WORLD 16
#define KM_PEER_TABLE_BYTES (64 * 1024)
#define KM_PEER_FLAGS_OFF (3 * KM_PEER_TABLE_BYTES)
#define KM_PEER_REGION_BYTES (2u << 20)
// general path (direct / tensor kernels): two alternating snapshots of this rank's running table, copied in before
// each sum across ranks; their epoch flags sit 128 bytes after the small-d path's
#define KM_PEER_FLAGS2_OFF (KM_PEER_FLAGS_OFF + 128)
#define KM_PEER_SNAP_OFF (KM_PEER_FLAGS_OFF + 4096)
#define KM_PEER_SNAP_BYTES (896 * 1024)
#define KM_RING 8      // stats slots in pinned host memory
#define KM_RUN_LAG 2   // iterations km_run keeps in flight ahead of the host's check
#define KM_PROF_RING 256  // assign-kernel event pairs kept when km_config.profile_kernels is set

struct km_ctx {
  km_config cfg;
  int device = 0;
  int sm_count = 148;
  int max_smem_optin = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  std::string err;

  // ---- a1: store
  void *x = nullptr;          // device, n x d row-major
  bool own_x = false;
  int dtype = KM_DTYPE_F32;
  int64_t n = 0;              // this rank's shard
  int64_t i0 = 0;             // global index of row 0 of the shard
  int64_t n_total = 0;
  int32_t d = 0;
  double maxabs = 0.0;        // over the whole job
  double bbox[4] = {0, 0, 0, 0};  // d == 2, float32: job-wide xmin, xmax, ymin, ymax
  bool have_bbox = false;
  bool have_points = false;
  int32_t *labels = nullptr;  // device, n
  unsigned *perm = nullptr;   // device, n: original row of each stored row once the store was cell-sorted (km_sort.cu)
  unsigned long long *cells_g = nullptr;   // device: candidate word per grid cell of the sorted store (rebuilt every pass)
  void *unperm_buf = nullptr; // device, n labels: where km_get_labels / small km_get_points put rows back in the caller's order
  KmPrune2 sort_pr;           // the grid the stored order follows
  int passes_since_upload = 0;
  void *pinned = nullptr;     // staging for uploads / downloads
  size_t pinned_bytes = 0;

  // ---- centroids
  int32_t k = 0;
  bool have_centroids = false;
  double *c64 = nullptr;      // device k x d, binary64 master
  float *chi = nullptr;       // device k x d, fl32(c)
  float *clo = nullptr;       // device k x d, fl32(c - chi)

  // ---- accumulators / state
  long long *acc = nullptr;         // device, local table
  long long *acc_global = nullptr;  // device, all-reduced table (world > 1), else == acc
  int64_t acc_len = 0;
  KmState *state = nullptr;         // device
  KmStatSlot *ring = nullptr;       // pinned host, KM_RING slots (device-mapped)
  KmStatSlot *ring_dev = nullptr;   // device alias of `ring`
  cudaEvent_t ring_ev[KM_RING];
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int64_t launches = 0;
  unsigned char *narrow_buf = nullptr;   // [n] labels as bytes (km_get_labels_u8), allocated by its first call per upload shape
  int64_t narrow_n = 0;
  int coop_ok = 0;                  // device supports cooperative launches (km_resident.cu)
  int32_t host_it = 0;              // assign steps issued since the last label reset
  int32_t last_path = 0;
  int32_t QB = 0, qa = 0, E = 0, NB = 0;
  int32_t QBt = 0, qt = 0;          // grid of T = sum |x|^2 (inertia from the sums)
  long long Ta = 0, Tb = 0;         // T of the whole job
  long long *scan_t = nullptr;      // device scratch: this rank's T limbs
  double cbound = 0.0;              // max(max|x|, max|c0|): bounds max|c| for every pass until centroids are set again
  bool use_hilo = false;            // small-d filter escalated to hi/lo centroid pairs
  bool labels_fresh = true;         // labels == -1 everywhere, sums zero

  double *scan_max = nullptr;       // device scratch of the upload scan
  int *scan_bad = nullptr;
  unsigned *scan_bbox = nullptr;    // device, 4 ordered-uint encodings
  cudaEvent_t prof_ev[KM_PROF_RING][2];
  int prof_n = 0;

  // ---- small-d path, prologue-finalize protocol
  bool p1 = false;
  long long *p1S[2] = {nullptr, nullptr};  // merged sums + counts (ping-pong)
  double *p1C[2] = {nullptr, nullptr};     // centroids (ping-pong)
  long long *p1D[3] = {nullptr, nullptr, nullptr};  // per-pass delta tables (+ scalars), rotating
  int p1_s = 0;            // current S / C index
  int p1_dw = 0;           // delta table the next pass writes (zero by invariant)
  bool p1_pending = false; // a pass's deltas await merging (and its stats publishing)
  int p1_dpend = 0;        // ... in this table
  int p1_pend_update = 0, p1_pend_nc = 0;  // ... to be applied with this intent
  int32_t p1_pend_it = 0;

  // ---- peer all-reduce of the small-d tables (km_peer.cu); set up in km_comm_init
  bool peer_ok = false;             // every rank mapped every peer's region
  void *peer_base[KM_PEER_MAX_WORLD] = {};   // regions as mapped in this process; [rank] is this rank's own allocation
  bool p1D_in_peer = false;         // p1D[] point into peer_base[rank] (not separately allocated)
  long long *p1R[3] = {nullptr, nullptr, nullptr};  // the summed tables the next pass reads
  unsigned long long peer_epoch = 0;
  unsigned long long peer_epoch2 = 0;  // ... of the general path's snapshots
  int *peer_err = nullptr;          // mapped pinned host word: a peer never arrived

  // ---- tensor path scratch
  float *t_cn = nullptr;            // -|chi_j|^2 / 2 per (padded) centroid
  void *t_cbf = nullptr;            // BF16 copy of the centroids
  unsigned *t_hist = nullptr;       // k + 1 bucket cursors of the label-ordered index list
  unsigned *t_work_count = nullptr;
  unsigned *t_order = nullptr;      // [n] point indices ordered by previous label
  int *t_work = nullptr;            // [n] worklist entries of the exact stage
  bool t_order_valid = false;       // t_order holds a list built from this upload's labels
  int t_order_age = 0;              // passes since it was built
  float *t_xn = nullptr;            // [n] |x_i|, computed on the first tensor pass after an upload
  bool t_xn_valid = false;
  void *t_xbf = nullptr;            // [n][64 * kblocks] BF16 copy of x (same lifetime as t_xn)
  size_t t_xbf_bytes = 0;
  int64_t t_work_n = 0;
  int path_now = 0;                 // KM_PATH_* chosen for the current (points, centroids)

  // ---- multi-GPU
  int rank = 0, world = 1;
  void *nccl_comm = nullptr;
};

struct KmAssignArgs {
  const void *x;
  int dtype;
  int64_t n;
  int32_t d, k;
  int32_t *labels;
  const double *c64;
  const float *chi;
  const float *clo;
  long long *acc;
  KmState *state;
  int incremental;     // 1: move only re-labelled points; 0: add every point (acc S zeroed)
  int all_push;        // small-d path: every point will be accumulated in this pass (labels all -1, or full-update mode)
  int exact_inertia;
  int sm_count;
  int max_smem_optin;
  double scale_a, scale_b;  // 2^qa, 2^QB (host copies of KmState's)
  double h_cbound, h_inertia_scale;  // small-d path: host-side bound on max|c| and the scales derived from it
  int h_qi;
  float h_tau_hilo, h_tau_hi;
  int hilo;            // small-d filter: 1 = hi/lo binary32 centroid pairs (tighter bound, +1 op)
  KmPass ps;           // small-d path: the pass's sums / centroids / delta tables
  float *t_cn;         // tensor path: -|chi_j|^2 / 2 (padded with -inf)
  void *t_cbf;         // tensor path: BF16 copy of the centroids, [kpad][kblocks*64]
  unsigned *t_hist, *t_work_count, *t_order;
  int *t_work;         // tensor path: worklist of the exact stage
  const float *t_xn;   // tensor path: |x_i| per point
  const void *t_xbf;   // tensor path: BF16 copy of x, rows padded to 64 * kblocks
  int t_labels_fresh;  // every label is still -1: no point ordering the points by it
  int t_order_reuse;   // keep the index list of an earlier pass (rebuilt every few passes once the run has settled)
  float t_xnmax;       // tensor path: upper bound of |x_i|
  int prune;           // small-d path, d == 2: use the candidate grid `pr`
  KmPrune2 pr;
  int sorted;          // ... and the points are stored in ascending cell order of `pr`
  unsigned long long *cells_g;   // ... with room for one candidate word per cell (km_prune_grid_kernel)
  unsigned *work_counter;        // ... and a counter the assignment kernel's warps take their units of work from
  cudaStream_t stream;
};

// each returns cudaError_t and bumps *launches by the number of kernels enqueued
cudaError_t km_launch_assign_direct(const KmAssignArgs &a, int64_t *launches);
cudaError_t km_launch_assign_small(const KmAssignArgs &a, int64_t *launches);
cudaError_t km_launch_assign_tensor(const KmAssignArgs &a, int64_t *launches);
bool km_tensor_path_supported(int32_t d, int32_t k, int dtype, int max_smem_optin);
size_t km_tensor_cbf_bytes(int32_t d, int32_t k);
size_t km_tensor_h_bytes(int32_t k);
size_t km_tensor_work_bytes(int64_t n);
size_t km_tensor_xbf_bytes(int64_t n, int32_t d);
cudaError_t km_launch_xbf(const float *x, int64_t n, int32_t d, void *xbf, int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_row_norms(const float *x, int64_t n, int32_t d, float *xn, int sm_count, cudaStream_t s, int64_t *launches);
bool km_small_prune_setup(int32_t d, int32_t k, int64_t n_local, int64_t min_points, int sm_count, const double bbox[4],
                          int fine, KmPrune2 *out);
cudaError_t km_narrow_labels(const int32_t *labels, const unsigned *perm, int64_t n, unsigned char *out_dev, int sm_count,
                             cudaStream_t s, int64_t *launches);
cudaError_t km_launch_peer_allreduce(void *const *bases, int world, int rank, size_t table_off, size_t flags_off, int64_t len,
                                     unsigned long long epoch, long long *out, int *err, int sm_count, cudaStream_t s, int64_t *launches);
// km_resident.cu: the whole km_run loop in one cooperative launch for jobs that fit in shared memory
#define KM_RESIDENT_MAX_KD 32    // binary64 argmin over all k, no filter: only worth it while k * d is small
struct KmResidentArgs {
  const float *x; int32_t *labels; int64_t n; int d, k, per_block, grid, max_iter; size_t smem;
  const double *C_in; double *C_out; long long *S_out; long long *Dt[3];
  KmStatSlot *ring; int32_t *state_it;
  double scale_a, scale_b; float tau2, rho2; int QB, qa; long long Ta, Tb; int qt, QBt, qg, QBg;
};
bool km_resident_plan(int64_t n, int32_t d, int32_t k, int dtype, int sm_count, int max_smem_optin, int *per_block, int *grid,
                      size_t *smem);
cudaError_t km_launch_resident(const KmResidentArgs &a, cudaStream_t stream, int64_t *launches);
#define KM_PRUNE_MAX_GP 260   // finest grid side (256 cells + margins)
bool km_small_path_supported(int32_t d, int32_t k, int dtype, int max_smem_optin);

cudaError_t km_launch_scan(const void *x, int dtype, int64_t count, double *maxabs_out /*device*/,
                           int *nonfinite_out /*device*/, int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_sumsq(const void *x, int dtype, int64_t count, double scale_t, double scale_tb, long long *out2 /*device*/,
                            int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_bbox2(const float *x, int64_t n, unsigned *out4, int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_gen_blobs(float *x, int64_t n, int64_t i0, int32_t d, int32_t k_true, uint64_t seed,
                                double sigma, int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_fill_i32(int32_t *p, int64_t n, int32_t v, int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_finalize(const KmFinalizeArgs &f, cudaStream_t s, int64_t *launches);

// k-means++ passes (km_kernels_misc.cu)
#define KM_PP_TILE 4096
cudaError_t km_launch_pp_update(const void *x, int dtype, int64_t n, int32_t d, const double *c_dev, double *D,
                                double scale, int first, long long *partial, cudaStream_t s, int64_t *launches);
cudaError_t km_launch_pp_find(const double *D, int64_t start, int count, double scale, long long target,
                              long long *out_idx, cudaStream_t s, int64_t *launches);

cudaError_t km_launch_cluster_inertia(const void *x, int dtype, const int32_t *labels, int64_t n, int32_t d, int32_t k,
                                      const double *c_dev, double scale_i, double scale_b, long long *table,
                                      int sm_count, cudaStream_t s, int64_t *launches);

// km_sort.cu: one-time reordering of owned d = 2 points by candidate-grid cell
cudaError_t km_sort_points_by_cell(float **x, int32_t **labels, int64_t n, const KmPrune2 &pr, unsigned **perm_out,
                                   int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_unpermute_labels(const int32_t *labels_sorted, const unsigned *perm, int64_t n, int32_t *out_dev,
                                int sm_count, cudaStream_t s, int64_t *launches);
cudaError_t km_unpermute_rows2(const float *x_sorted, const unsigned *perm, int64_t n, int64_t i0, int64_t cnt,
                               float *out_dev, int sm_count, cudaStream_t s, int64_t *launches);
