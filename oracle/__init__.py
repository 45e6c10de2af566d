"""ctypes front end of the CPU oracle — TEST INFRASTRUCTURE, NOT PRODUCT.

PARITY UNPINNED: the reference holds no Lloyd code, tests or golden vectors
(see kmeans_oracle.c's header and DESIGN.md §3).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this package.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

F32, F64 = 0, 1
SUM_SEQUENTIAL, SUM_EXACT = 0, 1

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libkmeans_oracle.so")
        if not os.path.exists(path) or (
                os.path.exists(_build.SRC) and os.path.getmtime(path) < os.path.getmtime(_build.SRC)):
            _build.build(force=True)
        L = C.CDLL(path)
        vp, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
        L.ko_splitmix64.restype = u64
        L.ko_splitmix64.argtypes = [u64]
        L.ko_hash3.restype = u64
        L.ko_hash3.argtypes = [u64, u64, u64]
        L.ko_blob_centres.restype = None
        L.ko_blob_centres.argtypes = [u64, i32, i32, vp]
        L.ko_gen_blobs.restype = None
        L.ko_gen_blobs.argtypes = [u64, i64, i64, i32, i32, dbl, vp]
        L.ko_forgy_indices.restype = i32
        L.ko_forgy_indices.argtypes = [i64, i32, u64, vp]
        L.ko_kmeanspp_indices.restype = i32
        L.ko_kmeanspp_indices.argtypes = [vp, C.c_int, i64, i32, i32, u64, C.c_double, i64, vp]
        L.ko_cluster_inertia.restype = None
        L.ko_cluster_inertia.argtypes = [vp, C.c_int, i64, i32, vp, vp, i32, vp]
        L.ko_assign.restype = i64
        L.ko_assign.argtypes = [vp, C.c_int, i64, i32, vp, i32, vp, vp]
        L.ko_assign_mt.restype = i64
        L.ko_assign_mt.argtypes = [vp, C.c_int, i64, i32, vp, i32, vp, vp, i32]
        L.ko_update_seq.restype = None
        L.ko_update_seq.argtypes = [vp, C.c_int, i64, i32, vp, i32, vp, vp]
        L.ko_fixed_params.restype = None
        L.ko_fixed_params.argtypes = [dbl, i64, vp, vp, vp, vp]
        L.ko_fixed_split.restype = None
        L.ko_fixed_split.argtypes = [dbl, i32, i32, vp, vp]
        L.ko_fixed_to_double.restype = dbl
        L.ko_fixed_to_double.argtypes = [i64, i64, i32, i32]
        L.ko_update_exact.restype = None
        L.ko_update_exact.argtypes = [vp, C.c_int, i64, i32, vp, i32, dbl, i64, vp, vp, vp, vp]
        L.ko_inertia_exact.restype = dbl
        L.ko_inertia_exact.argtypes = [vp, i64, i64, i32]
        L.ko_best_distances.restype = None
        L.ko_best_distances.argtypes = [vp, C.c_int, i64, i32, vp, i32, vp, vp]
        L.ko_maxabs.restype = dbl
        L.ko_maxabs.argtypes = [vp, C.c_int, i64]
        L.ko_run.restype = i32
        L.ko_run.argtypes = [vp, C.c_int, i64, i32, i32, vp, vp, i32, i32, dbl, i64, i32, vp, vp, vp]
        L.ko_balance.restype = None
        L.ko_balance.argtypes = [vp, i32, vp, vp, vp, vp]
        L.ko_num_threads.restype = i32
        L.ko_num_threads.argtypes = []
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _dtype_code(x):
    if x.dtype == np.float32:
        return F32
    if x.dtype == np.float64:
        return F64
    raise TypeError("points must be float32 or float64")


def _points(x):
    x = np.ascontiguousarray(x)
    if x.ndim != 2:
        raise ValueError("points must be N x d")
    return x, _dtype_code(x)


def gen_blobs(seed, n, d, k_true, sigma=1.0, i0=0):
    """A8 generator: rows [i0, i0+n) of the seeded blob matrix, float32."""
    x = np.empty((n, d), dtype=np.float32)
    lib().ko_gen_blobs(seed, i0, n, d, k_true, float(sigma), _p(x))
    return x


def blob_centres(seed, k_true, d):
    mu = np.empty((k_true, d), dtype=np.float64)
    lib().ko_blob_centres(seed, k_true, d, _p(mu))
    return mu


def forgy_indices(n, k, seed):
    """A7: indices of the k points a seeded partial Fisher-Yates picks."""
    out = np.empty(k, dtype=np.int64)
    rc = lib().ko_forgy_indices(n, k, seed, _p(out))
    if rc != 0:
        raise ValueError("forgy: need 0 <= k <= n")
    return out


def forgy_init(x, k, seed):
    idx = forgy_indices(x.shape[0], k, seed)
    return np.ascontiguousarray(x[idx].astype(np.float64))


def kmeanspp_indices(x, k, seed, maxabs_=None, n_total=None):
    """A7b: the k rows a seeded k-means++ (D^2 sampling on integer-quantised distances) picks."""
    x, dt = _points(x)
    n, d = x.shape
    out = np.empty(k, dtype=np.int64)
    m = maxabs(x) if maxabs_ is None else maxabs_
    rc = lib().ko_kmeanspp_indices(_p(x), dt, n, d, k, seed, float(m), n if n_total is None else n_total, _p(out))
    if rc != 0:
        raise ValueError("kmeans++: need 1 <= k <= n")
    return out


def kmeanspp_init(x, k, seed):
    idx = kmeanspp_indices(x, k, seed)
    return np.ascontiguousarray(np.asarray(x)[idx].astype(np.float64))


def cluster_inertia(x, labels, c):
    """f2: per-cluster sum of the members' squared distances to their centroid (index order)."""
    x, dt = _points(x)
    c = np.ascontiguousarray(c, dtype=np.float64)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    out = np.empty(c.shape[0], dtype=np.float64)
    lib().ko_cluster_inertia(_p(x), dt, x.shape[0], x.shape[1], _p(labels), _p(c), c.shape[0], _p(out))
    return out


def maxabs(x):
    x, dt = _points(x)
    return lib().ko_maxabs(_p(x), dt, x.size)


def assign(x, c, labels=None, nthreads=1):
    """A1/A2/A6.  Returns (labels, changed, inertia); `labels` in = previous labels."""
    x, dt = _points(x)
    c = np.ascontiguousarray(c, dtype=np.float64)
    n, d = x.shape
    k = c.shape[0]
    lab = np.full(n, -1, dtype=np.int32) if labels is None else np.array(labels, dtype=np.int32)
    inertia = C.c_double(0.0)
    if nthreads == 1:
        changed = lib().ko_assign(_p(x), dt, n, d, _p(c), k, _p(lab), C.byref(inertia))
    else:
        changed = lib().ko_assign_mt(_p(x), dt, n, d, _p(c), k, _p(lab), C.byref(inertia), nthreads)
    return lab, int(changed), inertia.value


def update(x, labels, c, sum_mode=SUM_SEQUENTIAL, maxabs_=None, n_total=None):
    """A3/A4.  Returns (new centroids, counts); empty clusters keep their centroid."""
    x, dt = _points(x)
    n, d = x.shape
    c = np.array(c, dtype=np.float64, order="C")
    k = c.shape[0]
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    counts = np.zeros(k, dtype=np.int64)
    if sum_mode == SUM_SEQUENTIAL:
        lib().ko_update_seq(_p(x), dt, n, d, _p(labels), k, _p(c), _p(counts))
    else:
        m = maxabs(x) if maxabs_ is None else maxabs_
        nt = n if n_total is None else n_total
        lib().ko_update_exact(_p(x), dt, n, d, _p(labels), k, float(m), nt, _p(c), _p(counts), None, None)
    return c, counts


def partial_sums_exact(x, labels, k, maxabs_, n_total):
    """The two int64 limbs per (cluster, dim) plus counts for one shard of the points."""
    x, dt = _points(x)
    n, d = x.shape
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    counts = np.zeros(k, dtype=np.int64)
    sa = np.zeros((k, d), dtype=np.int64)
    sb = np.zeros((k, d), dtype=np.int64)
    lib().ko_update_exact(_p(x), dt, n, d, _p(labels), k, float(maxabs_), n_total, None, _p(counts), _p(sa), _p(sb))
    return sa, sb, counts


def fixed_params(maxabs_, n_total):
    v = [C.c_int32(0) for _ in range(4)]
    lib().ko_fixed_params(float(maxabs_), n_total, *[C.byref(t) for t in v])
    return dict(E=v[0].value, NB=v[1].value, QB=v[2].value, qa=v[3].value)


def fixed_to_double(sa, sb, QB, shift):
    return lib().ko_fixed_to_double(int(sa), int(sb), QB, shift)


def centroids_from_limbs(sa, sb, counts, c_prev, maxabs_, n_total):
    p = fixed_params(maxabs_, n_total)
    c = np.array(c_prev, dtype=np.float64, order="C")
    k, d = c.shape
    for j in range(k):
        if counts[j] > 0:
            for t in range(d):
                c[j, t] = fixed_to_double(sa[j, t], sb[j, t], p["QB"], p["qa"] + p["QB"]) / float(counts[j])
    return c


def best_distances(x, c):
    x, dt = _points(x)
    c = np.ascontiguousarray(c, dtype=np.float64)
    n, d = x.shape
    b1 = np.empty(n, dtype=np.float64)
    b2 = np.empty(n, dtype=np.float64)
    lib().ko_best_distances(_p(x), dt, n, d, _p(c), c.shape[0], _p(b1), _p(b2))
    return b1, b2


def run(x, c0, max_iter=300, sum_mode=SUM_SEQUENTIAL, nthreads=1, maxabs_=None, n_total=None):
    """A5.  Returns dict(labels, centroids, iterations, inertia, counts, changed)."""
    x, dt = _points(x)
    n, d = x.shape
    c = np.array(c0, dtype=np.float64, order="C")
    k = c.shape[0]
    labels = np.empty(n, dtype=np.int32)
    counts = np.zeros(k, dtype=np.int64)
    inertia = C.c_double(0.0)
    changed = C.c_int64(0)
    m = 0.0
    if sum_mode == SUM_EXACT:
        m = maxabs(x) if maxabs_ is None else maxabs_
    nt = n if n_total is None else n_total
    it = lib().ko_run(_p(x), dt, n, d, k, _p(c), _p(labels), max_iter, sum_mode, float(m), nt,
                      nthreads, C.byref(inertia), _p(counts), C.byref(changed))
    return dict(labels=labels, centroids=c, iterations=int(it), inertia=inertia.value,
                counts=counts, changed=int(changed.value))


def balance(counts):
    counts = np.ascontiguousarray(counts, dtype=np.int64)
    mx, mn, gap = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    ratio = C.c_double(0.0)
    lib().ko_balance(_p(counts), counts.size, C.byref(mx), C.byref(mn), C.byref(gap), C.byref(ratio))
    return dict(max=mx.value, min=mn.value, gap=gap.value, ratio=ratio.value)


def num_threads():
    return lib().ko_num_threads()
