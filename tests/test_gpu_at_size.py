"""GPU: the named BASELINE.json configurations at (or near) their full per-GPU size, every label
checked against the oracle, plus an input built to attack the BF16 coarse stage's error bound.

Pattern: generate on the device, run 3 Lloyd iterations so the steady-state
machinery is live (cell-sorted store, label-ordered gather, seeded running minimum, survivor
lists, worklists), then compare EVERY label of one more km_assign with the oracle's argmin under
the centroids in use.  The oracle pass takes seconds to a minute on all host threads.
"""
import os

import numpy as np
import pytest

import oracle

gpu = pytest.mark.gpu


def _threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def _labels_after_iterations(km, n, d, k, k_true, seed, iters=3, **kw):
    with km.KMeansState(**kw) as st:
        st.generate_blobs(n, d, k_true, seed)
        st.init_forgy(k, seed)
        for _ in range(iters):
            s = st.step()
        c_used = st.centroids()
        s = st.assign()
        return st.get_points(), st.labels(), c_used, s


def _check(km, n, d, k, k_true, seed, path_name, **kw):
    x, lab, c_used, s = _labels_after_iterations(km, n, d, k, k_true, seed, **kw)
    assert s["path_used"] == getattr(km._lib, path_name)
    assert np.array_equal(x[:50000], oracle.gen_blobs(seed, 50000, d, k_true))    # the rows come back in generation order
    olab, _, oin = oracle.assign(x, c_used, nthreads=_threads())
    bad = int(np.count_nonzero(lab != olab))
    assert bad == 0, f"{bad} of {n} labels differ from the oracle"
    assert s["inertia"] == pytest.approx(oin, rel=1e-9)


@gpu
def test_configs4_shard_full_size(km):
    """configs[4] per-GPU shard: 1.25e8 x d2 x k64 (cell-sorted store, pruned small-d kernel)."""
    _check(km, 125_000_000, 2, 64, 64, 5, "PATH_FILTER_F32")


@gpu
def test_configs2_full_size(km):
    """configs[2]: 1e7 x d32 x k256 (tensor path)."""
    _check(km, 10_000_000, 32, 256, 256, 3, "PATH_TENSOR")


@gpu
def test_configs3_shard_2e6(km):
    """configs[3] per-GPU shape at 2e6 points: d128 x k1024 (tensor path, 8 centroid tiles)."""
    _check(km, 2_000_000, 128, 1024, 1024, 4, "PATH_TENSOR")


# ---------------------------------------------------------------- BF16 bound, attacked
def _bf16_rn(a):
    """float32 -> nearest bfloat16 (ties to even), returned as float32."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return r.astype(np.uint32).view(np.float32)


def _adversarial_case():
    """x is (almost) orthogonal to c1 and c2 = -c1, so both are (almost) equally far in exact
    arithmetic — c1 nearer by 4 * 2^-20 — while every coordinate of x and c1 sits on a BF16
    rounding midpoint arranged so that all rounding errors of x~.c~1 have the same sign: the
    coarse distances of c1 and c2 come out 2^-5 * sum|x_t c_t| apart, the wrong way round."""
    d, k = 64, 160
    a, b = np.float32(1 + 2.0 ** -8), np.float32(1 + 3 * 2.0 ** -8)   # round down to 1, up to 1 + 2^-6
    x = np.zeros(d, np.float32)
    c1 = np.zeros(d, np.float64)
    for t in range(0, 62, 2):
        x[t], c1[t] = a, a          # true +a^2, coarse 1            -> error < 0
        x[t + 1], c1[t + 1] = -b, b  # true -b^2, coarse -(1+2^-6)^2 -> error < 0
    x[62], c1[62] = np.float32(31 * 2.0 ** -6), 1.0
    x[63], c1[63] = np.float32(31 * 2.0 ** -13 + 2.0 ** -20), 1.0
    dot = float(np.dot(x.astype(np.float64), c1))
    assert dot == 2.0 ** -20
    rng = np.random.default_rng(5)
    c = np.empty((k, d), np.float64)
    for j in range(k):   # everything else lies well behind x, with a smaller norm than c1
        c[j] = -0.9 * x.astype(np.float64) + rng.uniform(-0.02, 0.02, d)
    c[3], c[131] = c1, -c1
    c[140] = c1 * (1 + 2.0 ** -30)        # near-duplicates of both, after them
    c[141] = -c1 * (1 + 2.0 ** -30)
    pts = [x]
    filler = rng.uniform(-1, 1, (400, d)).astype(np.float32)
    xs = np.concatenate([np.stack(pts), filler]).astype(np.float32)
    return xs, c


def test_adversarial_case_violates_the_old_window():
    """CPU self-check of the construction (no GPU): under BF16 rounding the two equally distant
    centroids come out further apart than the round-1 window 2 * 2^-6.5 |x|max |c|max, and
    inside the rigorous one."""
    xs, c = _adversarial_case()
    x, c1 = xs[0], c[3]
    xt, ct = _bf16_rn(x).astype(np.float64), _bf16_rn(c1.astype(np.float32)).astype(np.float64)
    gap = 4 * abs(float(np.dot(xt, ct)) - float(np.dot(x.astype(np.float64), c1)))   # dist~(c1) - dist~(c2)
    xnmax = np.sqrt(xs.shape[1]) * float(np.abs(xs).max())
    cnmax = float(np.sqrt((c.astype(np.float32).astype(np.float64) ** 2).sum(1)).max())
    assert gap > 2 * 2.0 ** -6.5 * xnmax * cnmax * 1.02
    assert gap < 2 * (2.0 ** -6 + 2.0 ** -15) * xnmax * cnmax



@gpu
def test_tensor_bf16_bound_adversarial(km):
    xs, c = _adversarial_case()
    olab, _, oin = oracle.assign(xs, c)
    assert olab[0] == 3          # c1 is nearer by 4 * 2^-20; its mirror image 131 is what a too-small window picks
    with km.KMeansState(path=km._lib.PATH_TENSOR) as st:
        st.upload_points(xs)
        st.set_centroids(c)
        s = st.assign()
        lab = st.labels()
        assert s["path_used"] == km._lib.PATH_TENSOR
    assert np.array_equal(lab, olab), np.flatnonzero(lab != olab)[:10]
    assert s["inertia"] == pytest.approx(oin, rel=1e-9)
