"""Host-side mirror of the JS facade the north-star asks for (kmeans.mjs):

    init(points, k, {seed} | {centroids}) -> state
    assign(state) / update(state) / step(state)
    run(state, {maxIter}) -> {labels, centroids, iterations, inertia}

Same names, argument meaning and error behaviour (a non-zero km_status becomes a raised
KMeansError, as the N-API shim turns it into a thrown JS Error).  Everything computes on
the GPU through the C ABI in include/kmeans.h; there is no CPU path in this module.

The reference has no such module to mirror (its "k-means" is people dragging cards:
/root/reference/app.mjs:126-129 add a centroid, :146-156 assign a card, :450-461/:571-573
summarise and rename, :286-288 bump the iteration counter); the entry points keep that
vocabulary.
"""
import ctypes as C

import numpy as np

from . import _lib as L
from . import board_json


class KMeansError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"libkmeans status {status}: {message}")
        self.status = status


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class KMeansState:
    """One km_ctx: one GPU's shard of the points, the centroids, labels and sums."""

    def __init__(self, device=0, path=L.PATH_AUTO, update_mode=L.UPDATE_INCREMENTAL, exact_inertia=True,
                 stream=None, profile_kernels=False, no_prune=False, prune_min_points=0):
        self._lib = L.load()
        cfg = L.Config(struct_size=C.sizeof(L.Config), device=device, path=path, update_mode=update_mode,
                       exact_inertia=1 if exact_inertia else 0,
                       profile_kernels=1 if profile_kernels else 0, stream=stream,
                       no_prune=1 if no_prune else 0, prune_min_points=prune_min_points)
        h = C.c_void_p()
        st = self._lib.km_create(C.byref(cfg), C.byref(h))
        if st != L.OK:
            raise KMeansError(st, (self._lib.km_last_error(None) or b"").decode())
        self._h = h
        self.n = 0
        self.d = 0
        self.k = 0
        self.dtype = np.float32
        self.rank, self.world = 0, 1

    # -- plumbing
    def _check(self, st):
        if st != L.OK:
            raise KMeansError(st, (self._lib.km_last_error(self._h) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._lib.km_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- multi-GPU
    @staticmethod
    def comm_unique_id():
        lib = L.load()
        buf = (C.c_char * 128)()
        st = lib.km_comm_unique_id(buf)
        if st != L.OK:
            raise KMeansError(st, (lib.km_last_error(None) or b"").decode())
        return bytes(buf)

    def comm_init(self, unique_id, rank, world):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        self._check(self._lib.km_comm_init(self._h, buf, rank, world))
        self.rank, self.world = rank, world

    def comm_info(self):
        """(rank, world, peer_allreduce): 0 = NCCL only, 1 = peer regions mapped, 2 = in use by the installed tables."""
        r, w, p = C.c_int32(0), C.c_int32(1), C.c_int32(0)
        self._check(self._lib.km_comm_info(self._h, C.byref(r), C.byref(w), C.byref(p)))
        return r.value, w.value, p.value

    # -- a1: store
    def upload_points(self, points):
        x = np.ascontiguousarray(points)
        if x.ndim != 2:
            raise ValueError("points must be N x d")
        if x.dtype == np.float32:
            dt = L.DTYPE_F32
        elif x.dtype == np.float64:
            dt = L.DTYPE_F64
        else:
            raise TypeError("points must be float32 or float64")
        self._check(self._lib.km_upload_points(self._h, _ptr(x), x.shape[0], x.shape[1], dt))
        self.n, self.d, self.dtype = x.shape[0], x.shape[1], x.dtype

    def upload_points_raw(self, host_ptr, n, d, dtype=np.float32):
        """Upload from a raw host address (e.g. page-locked memory from host_alloc)."""
        dt = L.DTYPE_F32 if np.dtype(dtype) == np.float32 else L.DTYPE_F64
        self._check(self._lib.km_upload_points(self._h, C.c_void_p(host_ptr), n, d, dt))
        self.n, self.d, self.dtype = n, d, np.dtype(dtype)

    def attach_points(self, device_ptr, n, d, dtype=np.float32):
        dt = L.DTYPE_F32 if np.dtype(dtype) == np.float32 else L.DTYPE_F64
        self._check(self._lib.km_attach_points(self._h, C.c_void_p(device_ptr), n, d, dt))
        self.n, self.d, self.dtype = n, d, np.dtype(dtype)

    def generate_blobs(self, n, d, k_true, seed, sigma=1.0, i0=0):
        self._check(self._lib.km_generate_blobs(self._h, n, i0, d, k_true, seed, float(sigma)))
        self.n, self.d, self.dtype = n, d, np.dtype(np.float32)

    @staticmethod
    def _check_out(out, count, dtype, what):
        """A caller-supplied result buffer must be exactly what the C call will write."""
        if not isinstance(out, np.ndarray) or out.dtype != np.dtype(dtype) or not out.flags["C_CONTIGUOUS"] \
                or out.size != count or not out.flags["WRITEABLE"]:
            raise ValueError(f"{what}: out must be a writable C-contiguous {np.dtype(dtype).name} array of {count} elements")

    def get_points(self, i0=0, n=None, out=None):
        n = self.n - i0 if n is None else n
        if out is None:
            out = np.empty((n, self.d), dtype=self.dtype)
        else:
            self._check_out(out, n * self.d, self.dtype, "get_points")
        self._check(self._lib.km_get_points(self._h, _ptr(out), i0, n))
        return out

    # -- a2: initialisation
    def set_centroids(self, centroids):
        c = np.ascontiguousarray(centroids, dtype=np.float64)
        if c.ndim != 2 or (self.d and c.shape[1] != self.d):
            raise ValueError("centroids must be k x d")
        self._check(self._lib.km_set_centroids(self._h, _ptr(c), c.shape[0]))
        self.k = c.shape[0]

    def init_forgy(self, k, seed):
        self._check(self._lib.km_init_forgy(self._h, k, seed))
        self.k = k

    def init_kmeanspp(self, k, seed):
        self._check(self._lib.km_init_kmeanspp(self._h, k, seed))
        self.k = k

    # -- a3 / a4 / a5
    def _stats_call(self, fn, *args):
        s = L.Stats()
        self._check(fn(self._h, *args, C.byref(s)))
        return s.as_dict()

    def assign(self):
        return self._stats_call(self._lib.km_assign)

    def update(self):
        return self._stats_call(self._lib.km_update)

    def step(self):
        return self._stats_call(self._lib.km_step)

    def step_async(self):
        self._check(self._lib.km_step_async(self._h))

    def stats_fetch(self):
        return self._stats_call(self._lib.km_stats_fetch)

    def run(self, max_iter=300):
        return self._stats_call(self._lib.km_run, max_iter)

    def reset_labels(self):
        self._check(self._lib.km_reset_labels(self._h))

    def set_update_mode(self, mode):
        """Switch between UPDATE_INCREMENTAL and UPDATE_FULL (labels and sums start over)."""
        self._check(self._lib.km_set_update_mode(self._h, mode))

    # -- results
    def labels(self, out=None, dtype=np.int32):
        """This rank's labels in the caller's row order; dtype=np.uint8 (k <= 255, 255 = never assigned)
        brings them home as bytes."""
        dtype = np.dtype(dtype) if out is None else out.dtype
        if dtype not in (np.dtype(np.int32), np.dtype(np.uint8)):
            raise TypeError("labels are int32 or uint8")
        if out is None:
            out = np.empty(self.n, dtype=dtype)
        else:
            self._check_out(out, self.n, dtype, "labels")
        if dtype == np.dtype(np.uint8):
            self._check(self._lib.km_get_labels_u8(self._h, _ptr(out), self.n))
        else:
            self._check(self._lib.km_get_labels(self._h, _ptr(out), self.n))
        return out

    def centroids(self, out=None):
        if out is None:
            out = np.empty((self.k, self.d), dtype=np.float64)
        else:
            self._check_out(out, self.k * self.d, np.float64, "centroids")
        self._check(self._lib.km_get_centroids(self._h, _ptr(out)))
        return out

    def counts(self):
        out = np.empty(self.k, dtype=np.int64)
        self._check(self._lib.km_get_counts(self._h, _ptr(out)))
        return out

    def balance(self):
        mx, mn, gap, ratio = C.c_int64(), C.c_int64(), C.c_int64(), C.c_double()
        self._check(self._lib.km_get_balance(self._h, C.byref(mx), C.byref(mn), C.byref(gap), C.byref(ratio)))
        return dict(max=mx.value, min=mn.value, gap=gap.value, ratio=ratio.value)

    def cluster_inertia(self):
        out = np.empty(self.k, dtype=np.float64)
        self._check(self._lib.km_get_cluster_inertia(self._h, _ptr(out)))
        return out

    def snapshot(self):
        """The numeric form of the dashboard's snapshotMetrics() (/root/reference/app.mjs:481-496):
        `counts`, `balance{max,min,gap,ratio}` with the same formulas, and — where the demo has a
        per-cluster cohesion and its average — `inertia` per cluster and its mean over clusters."""
        inertia = self.cluster_inertia()
        return dict(counts=self.counts(), inertia=inertia, balance=self.balance(),
                    avg_inertia=float(inertia.mean()) if inertia.size else 0.0, total_inertia=float(inertia.sum()))

    @staticmethod
    def deltas(prev, now):
        """What the dashboard prints next to its chips when a previous snapshot exists
        (app.mjs:516-521): change of the balance gap (<= 0 is "tighter") and of the average
        tightness; here also per-cluster count and inertia changes."""
        return dict(gap=int(now["balance"]["gap"] - prev["balance"]["gap"]),
                    avg_inertia=float(now["avg_inertia"] - prev["avg_inertia"]),
                    counts=now["counts"] - prev["counts"], inertia=now["inertia"] - prev["inertia"])

    def sums(self):
        sa = np.empty((self.k, self.d), dtype=np.int64)
        sb = np.empty((self.k, self.d), dtype=np.int64)
        p = (C.c_int32 * 4)()
        self._check(self._lib.km_get_sums(self._h, _ptr(sa), _ptr(sb), p))
        return sa, sb, dict(E=p[0], NB=p[1], QB=p[2], qa=p[3])

    def sync(self):
        self._check(self._lib.km_sync(self._h))

    def kernel_time(self, reset=True):
        """(summed ms, count) of the assignment kernels since the last reset (profile_kernels=True)."""
        ms, n = C.c_double(), C.c_int32()
        self._check(self._lib.km_kernel_time(self._h, C.byref(ms), C.byref(n), 1 if reset else 0))
        return ms.value, n.value

    def set_profile(self, on):
        self._check(self._lib.km_set_profile(self._h, 1 if on else 0))

    def launch_count(self):
        return int(self._lib.km_launch_count(self._h))


def host_alloc(nbytes):
    """Page-locked host buffer as a uint8 NumPy array (freed with host_free)."""
    lib = L.load()
    p = C.c_void_p()
    st = lib.km_host_alloc(C.byref(p), nbytes)
    if st != L.OK:
        raise KMeansError(st, (lib.km_last_error(None) or b"").decode())
    arr = np.ctypeslib.as_array((C.c_uint8 * max(nbytes, 1)).from_address(p.value))
    return arr, p.value


def host_free(addr):
    L.load().km_host_free(C.c_void_p(addr))


# ---- the facade proper: same four entry points as kmeans.mjs ---------------------------
def init(points, k, seed=None, centroids=None, method="forgy", **ctx_kwargs):
    """init(points, k, {seed} | {centroids}) -> state.  `method`: "forgy" or "kmeans++" (seeded)."""
    if method not in ("forgy", "kmeans++"):
        raise ValueError("init: method must be 'forgy' or 'kmeans++'")
    if (seed is None) == (centroids is None):
        raise ValueError("init: pass exactly one of seed= or centroids=")
    st = KMeansState(**ctx_kwargs)
    try:
        st.upload_points(points)
        if centroids is not None:
            c = np.asarray(centroids, dtype=np.float64)
            if c.shape[0] != k:
                raise ValueError("init: centroids must have k rows")
            st.set_centroids(c)
        elif method == "kmeans++":
            st.init_kmeanspp(k, seed)
        else:
            st.init_forgy(k, seed)
    except Exception:
        st.close()
        raise
    return st


def assign(state):
    return state.assign()


def update(state):
    return state.update()


def step(state):
    return state.step()


def run(state, max_iter=300):
    """run(state, {maxIter}) -> {labels, centroids, iterations, inertia} (+ stats)."""
    s = state.run(max_iter)
    return dict(labels=state.labels(), centroids=state.centroids(), iterations=s["iterations"],
                inertia=s["inertia"], converged=bool(s["converged"]), changed=s["changed"],
                rescored=s["rescored"], counts=state.counts(), device_ms=s["device_ms"])


# ---- the reference's on-disk format ------------------------------------------
def export_board(state, iteration=None, inertia=None):
    """The state as the `{cards, centroids, meta}` object the demo's Import button accepts
    (/root/reference/app.mjs:263-283); see board_json.py for the mapping.  Meant for boards a
    person can look at: one JSON object per point."""
    return board_json.export_board(state.get_points(), state.labels(), state.centroids(),
                                   iteration=0 if iteration is None else iteration, inertia=inertia,
                                   counts=state.counts())


def import_board(board, **ctx_kwargs):
    """A board exported by export_board -> (state, info): points uploaded, centroids installed.
    Labels are recomputed by the next assign() (a context's labels are its own)."""
    info = board_json.import_board(board)
    if info["points"] is None or info["centroids"] is None:
        raise ValueError("import_board: the board carries no numeric points / centroid coordinates")
    st = KMeansState(**ctx_kwargs)
    try:
        st.upload_points(info["points"])
        st.set_centroids(info["centroids"])
    except Exception:
        st.close()
        raise
    return st, info
