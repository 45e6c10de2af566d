"""CPU: the oracle against the golden vectors, the closed-form cases of SURVEY.md A9 and
an independent NumPy statement of the same spec.  The reference has no tests of its own
to mirror; these are ours."""
import glob
import json
import os

import numpy as np
import pytest

import oracle
from oracle import numpy_ref as nr


def golden_cases(golden_dir=os.path.join(os.path.dirname(__file__), "golden")):
    return sorted(glob.glob(os.path.join(golden_dir, "*.npz")))


@pytest.mark.parametrize("path", golden_cases(), ids=lambda p: os.path.basename(p)[:-4])
def test_oracle_matches_golden(path):
    g = np.load(path)
    r = oracle.run(g["x"], g["c0"], max_iter=int(g["max_iter"]))
    assert r["iterations"] == int(g["iterations"])
    assert np.array_equal(r["labels"], g["labels"])
    assert np.array_equal(r["centroids"], g["centroids"])  # same sequential sums: bit-exact
    assert r["inertia"] == float(g["inertia"])
    assert np.array_equal(r["counts"], g["counts"])
    lab, changed, inertia = oracle.assign(g["x"], g["c0"])
    assert np.array_equal(lab, g["first_labels"]) and changed == g["x"].shape[0]
    assert inertia == float(g["first_inertia"])
    c1, cnt = oracle.update(g["x"], lab, g["c0"])
    assert np.array_equal(c1, g["first_centroids"]) and np.array_equal(cnt, g["first_counts"])


@pytest.mark.parametrize("path", golden_cases(), ids=lambda p: os.path.basename(p)[:-4])
def test_generator_and_forgy_match_golden(path):
    g = np.load(path)
    n, d = g["x"].shape
    assert np.array_equal(oracle.gen_blobs(int(g["seed"]), n, d, int(g["k_true"])), g["x"])
    k = g["c0"].shape[0]
    assert np.array_equal(oracle.forgy_indices(n, k, int(g["init_seed"])), g["forgy_idx"])
    assert np.array_equal(oracle.kmeanspp_indices(g["x"], k, int(g["init_seed"])), g["kpp_idx"])
    # a slice generated from the middle equals the same rows of the whole
    assert np.array_equal(oracle.gen_blobs(int(g["seed"]), 50, d, int(g["k_true"]), i0=100), g["x"][100:150])


def test_a9_closed_form(golden_dir):
    for case in json.load(open(os.path.join(golden_dir, "a9.json"))):
        x = np.array(case["x"], dtype=np.float32)
        for mode in (oracle.SUM_SEQUENTIAL, oracle.SUM_EXACT):
            r = oracle.run(x, case["c0"], sum_mode=mode)
            assert r["labels"].tolist() == case["labels"], case["name"]
            assert r["centroids"].tolist() == [[float(v) for v in row] for row in case["centroids"]]
            assert r["inertia"] == case["inertia"] and r["iterations"] == case["iterations"]


def test_ties_lowest_index_and_nan_never_wins():
    x = np.array([[0.0, 0.0]], dtype=np.float32)
    c = np.array([[1, 0], [0, 1], [-1, 0], [0, -1]], dtype=np.float64)
    lab, _, inertia = oracle.assign(x, c)
    assert lab[0] == 0 and inertia == 1.0
    c[0, 0] = np.nan
    lab, _, _ = oracle.assign(x, c)
    assert lab[0] == 1
    lab, _, inertia = oracle.assign(x, np.full((3, 2), np.nan))
    assert lab[0] == 0 and inertia == np.inf  # nothing wins: label 0, best stays +Infinity


def test_float64_points_and_threads_agree():
    x = oracle.gen_blobs(5, 5000, 3, 4).astype(np.float64) * 1.000000001
    c0 = x[:4].copy()
    a = oracle.run(x, c0, nthreads=1)
    b = oracle.run(x, c0, nthreads=0)
    assert np.array_equal(a["labels"], b["labels"]) and a["iterations"] == b["iterations"]
    assert np.array_equal(a["centroids"], b["centroids"])
    assert a["inertia"] == pytest.approx(b["inertia"], rel=1e-12)
    n = nr.run(x, c0)
    assert np.array_equal(a["labels"], n["labels"]) and np.array_equal(a["centroids"], n["centroids"])


def test_exact_sums_equal_rational_arithmetic():
    rng = np.random.default_rng(0)
    x = (rng.standard_normal((400, 3)) * np.array([1e-3, 1.0, 300.0])).astype(np.float32)
    labels = rng.integers(0, 5, size=400).astype(np.int32)
    c_prev = np.zeros((5, 3))
    ce, cnt = oracle.update(x, labels, c_prev, sum_mode=oracle.SUM_EXACT)
    cf = nr.update_exact_fraction(x, labels, c_prev)
    assert np.array_equal(ce, cf)
    assert cnt.sum() == 400


def test_exact_sums_are_order_and_partition_free():
    x = oracle.gen_blobs(21, 3000, 2, 5)
    labels = (np.arange(3000) * 7 % 5).astype(np.int32)
    m, n = oracle.maxabs(x), 3000
    sa, sb, cnt = oracle.partial_sums_exact(x, labels, 5, m, n)
    perm = np.random.default_rng(1).permutation(3000)
    sa2, sb2, cnt2 = oracle.partial_sums_exact(x[perm], labels[perm], 5, m, n)
    assert np.array_equal(sa, sa2) and np.array_equal(sb, sb2) and np.array_equal(cnt, cnt2)
    parts = [oracle.partial_sums_exact(x[lo:hi], labels[lo:hi], 5, m, n) for lo, hi in ((0, 1), (1, 1700), (1700, 3000))]
    assert np.array_equal(sum(p[0] for p in parts), sa) and np.array_equal(sum(p[1] for p in parts), sb)
    # and the means they give sit within the sequential sum's own rounding of it
    ce = oracle.centroids_from_limbs(sa, sb, cnt, np.zeros((5, 2)), m, n)
    cs, _ = oracle.update(x, labels, np.zeros((5, 2)))
    assert np.allclose(ce, cs, rtol=1e-12, atol=1e-13)


def test_fixed_point_round_trip_edges():
    p = oracle.fixed_params(16.0, 10 ** 9)
    assert p["NB"] == 30 and p["E"] == 5 and p["QB"] == 32 and p["qa"] == 27
    for v in (0.0, 1.0, -1.0, 15.999999, -15.999999, 2.0 ** -30, -(2.0 ** -40), 1e-30, 3.1415927):
        v = float(np.float32(v))
        x = np.array([[v]], dtype=np.float32)
        c, _ = oracle.update(x, np.zeros(1, np.int32), np.zeros((1, 1)), sum_mode=oracle.SUM_EXACT, maxabs_=16.0,
                             n_total=10 ** 9)
        if abs(v) >= 2.0 ** -35 or v == 0.0:
            assert c[0, 0] == v
        else:
            assert abs(c[0, 0] - v) <= 2.0 ** -60


def test_inertia_never_increases_and_fixed_point():
    x = oracle.gen_blobs(3, 4000, 2, 6)
    c = oracle.forgy_init(x, 6, 1)
    labels, prev = None, np.inf
    for _ in range(40):
        labels, changed, inertia = oracle.assign(x, c, labels)
        assert inertia <= prev * (1 + 1e-15)
        prev = inertia
        if changed == 0:
            break
        c, counts = oracle.update(x, labels, c)
        assert counts.sum() == 4000
    l2, ch2, _ = oracle.assign(x, c, labels)
    assert ch2 == 0 and np.array_equal(l2, labels)


def test_balance_formulas():
    # /root/reference/app.mjs:489-493
    assert oracle.balance([4, 4, 4]) == dict(max=4, min=4, gap=0, ratio=1.0)
    assert oracle.balance([5, 0, 2])["ratio"] == np.inf
    assert oracle.balance([0, 0])["ratio"] == 1.0
    assert oracle.balance([6, 3]) == dict(max=6, min=3, gap=3, ratio=2.0)


def test_kmeanspp_matches_independent_numpy_statement():
    """A7b: the C oracle's seeded k-means++ against arbitrary-precision Python integers."""
    from oracle import numpy_ref as nr
    for (n, d, k, seed) in [(500, 2, 8, 1), (300, 5, 20, 7), (64, 1, 64, 3), (400, 3, 5, 11)]:
        x = oracle.gen_blobs(seed, n, d, max(2, k // 2))
        got = oracle.kmeanspp_indices(x, k, seed)
        assert np.array_equal(got, nr.kmeanspp_indices(x, k, seed))
        assert len(set(got.tolist())) == k          # distinct points: a chosen row has D = 0 and cannot be drawn again
    x = np.ones((10, 2), np.float32)                # all distances zero -> uniform fallback
    assert np.array_equal(oracle.kmeanspp_indices(x, 4, 5), nr.kmeanspp_indices(x, 4, 5))
    with pytest.raises(ValueError):
        oracle.kmeanspp_indices(x, 11, 1)


def test_cluster_inertia_sums_to_inertia():
    x = oracle.gen_blobs(9, 3000, 3, 6)
    c0 = oracle.forgy_init(x, 6, 2)
    r = oracle.run(x, c0, max_iter=50)
    ci = oracle.cluster_inertia(x, r["labels"], r["centroids"])
    assert ci.shape == (6,) and ci.sum() == pytest.approx(r["inertia"], rel=1e-12)
    d = nr.distances(x, r["centroids"])
    want = np.array([d[r["labels"] == j, j].sum() for j in range(6)])
    assert np.allclose(ci, want, rtol=1e-12)
