"""GPU: the CUDA path (through the C ABI) against the CPU oracle and the golden vectors.

Bar (north-star): labels bit-exact and iteration count identical; centroids and inertia
within 1e-9 relative of the oracle's sequential binary64 sums — and *bit-identical* to the
oracle's order-free (exact) sums, which is the stronger statement this design allows.
"""
import glob
import json
import os

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

REL = 1e-9  # north-star tolerance for centroids and inertia in binary64 mode
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def close(a, b, rel=REL):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    scale = max(1.0, float(np.max(np.abs(b))) if b.size else 1.0)
    return bool(np.all(np.abs(a - b) <= rel * scale))


def paths(km):
    return {"direct": km._lib.PATH_DIRECT_F64, "filter": km._lib.PATH_FILTER_F32, "auto": km._lib.PATH_AUTO}


def gpu_run(km, x, c0, max_iter=300, path="auto", update_mode=0, **kw):
    with km.KMeansState(path=paths(km)[path], update_mode=update_mode, **kw) as st:
        st.upload_points(x)
        st.set_centroids(c0)
        s = st.run(max_iter)
        sa, sb, p = st.sums()
        return dict(labels=st.labels(), centroids=st.centroids(), iterations=s["iterations"], inertia=s["inertia"],
                    counts=st.counts(), changed=s["changed"], rescored=s["rescored"], sums=(sa, sb, p),
                    path_used=s["path_used"], balance=st.balance())


def check_against_oracle(km, x, c0, max_iter=300, path="auto", update_mode=0):
    g = gpu_run(km, x, c0, max_iter, path, update_mode)
    o = oracle.run(x, c0, max_iter=max_iter)
    assert g["iterations"] == o["iterations"]
    assert np.array_equal(g["labels"], o["labels"])
    assert g["changed"] == o["changed"]
    assert np.array_equal(g["counts"], o["counts"])
    assert close(g["centroids"], o["centroids"])
    assert g["inertia"] == pytest.approx(o["inertia"], rel=REL)
    # order-free sums: the oracle's exact mode must give the very same bits
    e = oracle.run(x, c0, max_iter=max_iter, sum_mode=oracle.SUM_EXACT)
    assert e["iterations"] == g["iterations"] and np.array_equal(e["labels"], g["labels"])
    assert np.array_equal(e["centroids"], g["centroids"])
    assert g["balance"] == oracle.balance(o["counts"])
    return g, o


# ---------------------------------------------------------------- golden vectors
@pytest.mark.parametrize("path", ["direct", "auto"])
@pytest.mark.parametrize("case", GOLDEN, ids=lambda p: os.path.basename(p)[:-4])
def test_golden(km, case, path):
    g = np.load(case)
    r = gpu_run(km, g["x"], g["c0"], int(g["max_iter"]), path)
    assert r["iterations"] == int(g["iterations"])
    assert np.array_equal(r["labels"], g["labels"])
    assert np.array_equal(r["counts"], g["counts"])
    assert close(r["centroids"], g["centroids"])
    assert r["inertia"] == pytest.approx(float(g["inertia"]), rel=REL)
    if path == "auto":   # the seeded initialisations pick the golden rows
        with km.KMeansState() as st:
            st.upload_points(g["x"])
            st.init_forgy(g["c0"].shape[0], int(g["init_seed"]))
            assert np.array_equal(st.centroids(), g["x"][g["forgy_idx"]].astype(np.float64))
            st.init_kmeanspp(g["c0"].shape[0], int(g["init_seed"]))
            assert np.array_equal(st.centroids(), g["x"][g["kpp_idx"]].astype(np.float64))


def test_a9_closed_form(km, golden_dir):
    for case in json.load(open(os.path.join(golden_dir, "a9.json"))):
        x = np.array(case["x"], dtype=np.float32)
        for path in ("direct", "auto"):
            for mode in (0, 1):
                r = gpu_run(km, x, case["c0"], path=path, update_mode=mode)
                assert r["labels"].tolist() == case["labels"], case["name"]
                assert r["centroids"].tolist() == [[float(v) for v in row] for row in case["centroids"]]
                assert r["inertia"] == case["inertia"] and r["iterations"] == case["iterations"]


# ---------------------------------------------------------------- seeded blobs vs oracle
SHAPES = [  # n, d, k_true, k
    (5000, 2, 8, 8), (4099, 1, 3, 4), (20000, 3, 5, 5), (7777, 4, 6, 9), (3001, 5, 4, 4), (2500, 8, 7, 7),
    (12000, 2, 64, 64), (1000, 2, 3, 200), (3000, 32, 16, 16), (1500, 128, 8, 32), (700, 17, 5, 11),
]


@pytest.mark.parametrize("mode", [0, 1], ids=["incremental", "full"])
@pytest.mark.parametrize("path", ["direct", "auto"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "n%d_d%d_kt%d_k%d" % s)
def test_blobs_match_oracle(km, shape, path, mode):
    n, d, k_true, k = shape
    x = oracle.gen_blobs(100 + d, n, d, k_true)
    c0 = oracle.forgy_init(x, k, 7)
    g, _ = check_against_oracle(km, x, c0, max_iter=60, path=path, update_mode=mode)
    if path == "auto" and d <= 8:
        assert g["path_used"] == km._lib.PATH_FILTER_F32


def test_incremental_equals_full_bitwise(km):
    x = oracle.gen_blobs(5, 60000, 2, 16)
    c0 = oracle.forgy_init(x, 16, 3)
    a = gpu_run(km, x, c0, 25, "auto", 0)
    b = gpu_run(km, x, c0, 25, "auto", 1)
    c = gpu_run(km, x, c0, 25, "direct", 0)
    for other in (b, c):
        assert np.array_equal(a["labels"], other["labels"]) and np.array_equal(a["centroids"], other["centroids"])
        assert a["iterations"] == other["iterations"]
        assert np.array_equal(a["sums"][0], other["sums"][0]) and np.array_equal(a["sums"][1], other["sums"][1])
    # same kernel, same thread->point map: the inertia is the same bits; across kernels the
    # per-thread binary64 partials differ in their last bits only
    assert a["inertia"] == b["inertia"]
    assert a["inertia"] == pytest.approx(c["inertia"], rel=1e-12)


def test_sums_equal_oracle_limbs(km):
    x = oracle.gen_blobs(9, 30000, 3, 6)
    c0 = oracle.forgy_init(x, 6, 2)
    g = gpu_run(km, x, c0, 40)
    sa, sb, p = g["sums"]
    assert p == oracle.fixed_params(oracle.maxabs(x), x.shape[0])
    osa, osb, ocnt = oracle.partial_sums_exact(x, g["labels"], 6, oracle.maxabs(x), x.shape[0])
    assert np.array_equal(sa, osa) and np.array_equal(sb, osb) and np.array_equal(g["counts"], ocnt)


def test_float64_points(km):
    x = oracle.gen_blobs(31, 6000, 3, 5).astype(np.float64) * (1.0 + 2.0 ** -40)
    c0 = x[:5].copy()
    g, _ = check_against_oracle(km, x, c0, path="auto")
    assert g["path_used"] == km._lib.PATH_DIRECT_F64


# ---------------------------------------------------------------- edge cases
def test_exact_ties_take_lowest_index(km):
    # integer lattice points on perpendicular bisectors of integer centroids: every sum and
    # distance is exact, so ties are real ties and must break to the lowest index
    xs = np.array([[x, y] for x in range(-6, 7) for y in range(-6, 7)], dtype=np.float32)
    c0 = np.array([[2, 0], [-2, 0], [0, 2], [0, -2], [2, 0]], dtype=np.float64)  # incl. a duplicate centroid
    for path in ("direct", "filter"):
        with km.KMeansState(path=paths(km)[path]) as st:
            st.upload_points(xs)
            st.set_centroids(c0)
            s = st.assign()
            lab, _, inertia = oracle.assign(xs, c0)
            assert np.array_equal(st.labels(), lab)
            assert s["inertia"] == inertia
            assert not np.any(st.labels() == 4)  # the duplicate never wins
            if path == "filter":
                assert s["rescored"] > 0  # the ties went through the binary64 re-score
    check_against_oracle(km, xs, c0, path="filter")


def test_near_ties_one_ulp_off_the_bisector(km):
    # points within an ulp of the bisector between two centroids: the binary32 filter cannot
    # separate them, the re-score must, and the labels must still be the oracle's
    rng = np.random.default_rng(4)
    c0 = np.array([[1.0, 3.0], [5.0, -1.0], [40.0, 40.0]], dtype=np.float64)
    mid = (c0[0] + c0[1]) / 2
    dirv = np.array([1.0, 1.0]) / np.sqrt(2.0)  # along the bisector of c0[0], c0[1]
    t = rng.uniform(-3, 3, size=4000)
    pts = (mid[None, :] + t[:, None] * dirv[None, :]).astype(np.float32)
    pts[::3, 0] = np.nextafter(pts[::3, 0], np.float32(np.inf))
    pts[1::3, 1] = np.nextafter(pts[1::3, 1], np.float32(-np.inf))
    with km.KMeansState(path=paths(km)["filter"]) as st:
        st.upload_points(pts)
        st.set_centroids(c0)
        s = st.assign()
        lab, _, inertia = oracle.assign(pts, c0)
        assert np.array_equal(st.labels(), lab)
        assert s["rescored"] > 1000
        assert s["inertia"] == pytest.approx(inertia, rel=1e-12)
    check_against_oracle(km, pts, c0, path="filter")


def test_empty_cluster_keeps_centroid_and_duplicates(km):
    x = np.repeat(np.array([[1.0, 1.0], [2.0, 2.0], [1.0, 1.0]], dtype=np.float32), 50, axis=0)
    c0 = np.array([[0.0, 0.0], [3.0, 3.0], [100.0, 100.0], [-50.0, 7.0]])
    g, o = check_against_oracle(km, x, c0)
    assert g["counts"].tolist() == [100, 50, 0, 0]
    assert g["centroids"][2].tolist() == [100.0, 100.0] and g["centroids"][3].tolist() == [-50.0, 7.0]
    assert g["balance"]["ratio"] == np.inf and g["balance"]["gap"] == 100


def test_k1_kN_and_ragged_sizes(km):
    for n in (1, 2, 31, 33, 255, 257, 511, 513, 1023):
        x = oracle.gen_blobs(n, n, 2, 3)
        check_against_oracle(km, x, x[:1].astype(np.float64))            # k = 1
        check_against_oracle(km, x, oracle.forgy_init(x, min(n, 5), 1))   # ragged tail
    x = oracle.gen_blobs(77, 40, 3, 40)
    g, _ = check_against_oracle(km, x, x.astype(np.float64))             # k = N, C0 = X
    assert g["iterations"] == 2 and g["inertia"] == 0.0


def test_empty_input(km):
    with km.KMeansState() as st:
        st.upload_points(np.zeros((0, 2), dtype=np.float32))
        st.set_centroids(np.array([[0.0, 0.0], [1.0, 1.0]]))
        s = st.run(10)
        assert s["iterations"] == 1 and s["changed"] == 0 and s["inertia"] == 0.0
        assert st.labels().shape == (0,) and st.counts().tolist() == [0, 0]
        assert st.centroids().tolist() == [[0.0, 0.0], [1.0, 1.0]]


def test_wide_dynamic_range(km):
    rng = np.random.default_rng(8)
    x = (rng.standard_normal((5000, 3)) * np.array([1e-6, 1.0, 1e5])).astype(np.float32)
    c0 = x[[0, 10, 20, 30]].astype(np.float64)
    check_against_oracle(km, x, c0, path="auto")
    check_against_oracle(km, x * np.float32(1e20), c0 * 1e20, path="auto")  # binary32 squares overflow -> re-score


def test_max_iter_cap(km):
    x = oracle.gen_blobs(18, 3000, 2, 3)
    c0 = oracle.forgy_init(x, 8, 10)
    for cap in (1, 2, 3, 5):
        g, o = check_against_oracle(km, x, c0, max_iter=cap)
        assert g["iterations"] == cap and g["changed"] > 0


# ---------------------------------------------------------------- stepwise API, generator, init
def test_stepwise_assign_update_equals_oracle(km):
    x = oracle.gen_blobs(2, 9000, 2, 8)
    c = oracle.forgy_init(x, 8, 2)
    with km.KMeansState() as st:
        st.upload_points(x)
        st.set_centroids(c)
        labels = None
        for it in range(1, 8):
            s = st.assign()
            labels, changed, inertia = oracle.assign(x, c, labels)
            assert s["iterations"] == it and s["changed"] == changed
            assert np.array_equal(st.labels(), labels)
            assert s["inertia"] == pytest.approx(inertia, rel=REL)
            st.update()
            c_exact, counts = oracle.update(x, labels, c, sum_mode=oracle.SUM_EXACT)
            c_seq, _ = oracle.update(x, labels, c)
            assert np.array_equal(st.centroids(), c_exact) and close(st.centroids(), c_seq)
            assert np.array_equal(st.counts(), counts)
            c = c_exact


def test_step_and_async_step_equal_run(km):
    x = oracle.gen_blobs(6, 50000, 2, 8)
    c0 = oracle.forgy_init(x, 8, 6)
    ref = gpu_run(km, x, c0, 12)
    n_it = ref["iterations"]
    with km.KMeansState() as st:
        st.upload_points(x)
        st.set_centroids(c0)
        # km_run never updates after its last assign (A5): n_it-1 full steps, then one assign
        for _ in range(n_it - 1):
            st.step()
        s = st.assign()
        assert s["iterations"] == n_it and np.array_equal(st.labels(), ref["labels"])
        assert np.array_equal(st.centroids(), ref["centroids"]) and s["inertia"] == ref["inertia"]
        st.reset_labels()
        st.set_centroids(c0)
        for _ in range(n_it - 1):
            st.step_async()
        s = st.stats_fetch()
        assert s["iterations"] == n_it - 1
        s = st.assign()
        assert s["iterations"] == n_it and np.array_equal(st.labels(), ref["labels"])
        assert np.array_equal(st.centroids(), ref["centroids"]) and s["inertia"] == ref["inertia"]
        assert st.launch_count() > 0


def test_device_generator_is_the_oracle_generator(km):
    for (n, d, kt, seed, i0) in ((5000, 2, 8, 2, 0), (1234, 7, 5, 99, 0), (600, 32, 16, 3, 4000)):
        with km.KMeansState() as st:
            st.generate_blobs(n, d, kt, seed, sigma=1.0, i0=i0)
            assert np.array_equal(st.get_points(), oracle.gen_blobs(seed, n, d, kt, i0=i0))


def test_forgy_init_is_the_oracle_forgy(km):
    x = oracle.gen_blobs(41, 10000, 3, 6)
    for k, seed in ((6, 1), (50, 2), (1, 3)):
        with km.KMeansState() as st:
            st.upload_points(x)
            st.init_forgy(k, seed)
            assert np.array_equal(st.centroids(), oracle.forgy_init(x, k, seed))


def test_facade_entry_points(km):
    x = oracle.gen_blobs(2, 20000, 2, 8)
    st = km.init(x, 8, seed=5)
    out = km.run(st, max_iter=50)
    o = oracle.run(x, oracle.forgy_init(x, 8, 5), max_iter=50)
    assert out["iterations"] == o["iterations"] and np.array_equal(out["labels"], o["labels"])
    assert close(out["centroids"], o["centroids"]) and out["inertia"] == pytest.approx(o["inertia"], rel=REL)
    st.close()
    with pytest.raises(ValueError):
        km.init(x, 8)


# ---------------------------------------------------------------- error behaviour
@pytest.mark.parametrize("shape", [(5000, 2, 8, np.float32), (3000, 5, 20, np.float32), (2000, 32, 16, np.float32),
                                   (1500, 3, 7, np.float64), (64, 1, 64, np.float32), (20000, 2, 64, np.float32)],
                         ids=lambda s: "n%d_d%d_k%d_%s" % (s[0], s[1], s[2], np.dtype(s[3]).name))
def test_kmeanspp_picks_the_oracles_rows(km, shape):
    """f1: seeded k-means++ (D^2 sampling on integer-quantised distances, A7b) — the very rows the
    oracle picks, hence the very same run."""
    n, d, k, dt = shape
    x = oracle.gen_blobs(50 + d, n, d, max(2, k // 2)).astype(dt)
    want = oracle.kmeanspp_init(x, k, 21)
    with km.KMeansState() as st:
        st.upload_points(x)
        st.init_kmeanspp(k, 21)
        assert np.array_equal(st.centroids(), want)
        s = st.run(30)
        o = oracle.run(x, want, max_iter=30)
        assert s["iterations"] == o["iterations"] and np.array_equal(st.labels(), o["labels"])


def test_kmeanspp_edge_cases(km):
    # every point on one spot: all distances 0 -> the uniform fallback draw
    x = np.ones((10, 2), dtype=np.float32)
    with km.KMeansState() as st:
        st.upload_points(x)
        st.init_kmeanspp(4, 5)
        assert np.array_equal(st.centroids(), oracle.kmeanspp_init(x, 4, 5))
    # after the store was cell-sorted the draw still walks the caller's row order
    x = oracle.gen_blobs(8, 40000, 2, 32)
    with km.KMeansState(prune_min_points=1) as st:
        st.upload_points(x)
        st.init_forgy(32, 1)
        st.run(10)                       # sorts the store
        st.init_kmeanspp(32, 9)
        assert np.array_equal(st.centroids(), oracle.kmeanspp_init(x, 32, 9))
        assert np.array_equal(st.get_points(), x)
        s = st.run(40)
        o = oracle.run(x, oracle.kmeanspp_init(x, 32, 9), max_iter=40)
        assert s["iterations"] == o["iterations"] and np.array_equal(st.labels(), o["labels"])
    st = km.init(x[:5000], 8, seed=3, method="kmeans++")
    assert np.array_equal(st.centroids(), oracle.kmeanspp_init(x[:5000], 8, 3))
    with pytest.raises(km.KMeansError):
        st.init_kmeanspp(0, 1)
    st.close()


@pytest.mark.parametrize("shape", [(20000, 2, 8, {}), (6000, 5, 12, {}), (4000, 32, 16, {}), (40000, 2, 64, {"prune_min_points": 1})],
                         ids=lambda s: "n%d_d%d_k%d" % s[:3])
def test_cluster_inertia_and_snapshot_deltas(km, shape):
    """f2: per-cluster inertia (the numeric stand-in for the dashboard's cohesion) and the
    snapshot / delta pair of app.mjs:481-521."""
    n, d, k, kw = shape
    x = oracle.gen_blobs(70 + d, n, d, k)
    c0 = oracle.forgy_init(x, k, 4)
    with km.KMeansState(**kw) as st:
        st.upload_points(x)
        st.set_centroids(c0)
        s1 = st.assign()
        snap1 = st.snapshot()
        lab1, _, _ = oracle.assign(x, c0)
        assert np.allclose(snap1["inertia"], oracle.cluster_inertia(x, lab1, c0), rtol=1e-10, atol=0)
        assert snap1["total_inertia"] == pytest.approx(s1["inertia"], rel=1e-12)
        assert snap1["balance"] == oracle.balance(snap1["counts"])
        s = st.run(40)   # (sorts the store where the pruned path is on)
        snap2 = st.snapshot()
        o = oracle.run(x, c0, max_iter=40)
        assert np.allclose(snap2["inertia"], oracle.cluster_inertia(x, o["labels"], o["centroids"]), rtol=1e-10, atol=0)
        assert snap2["total_inertia"] == pytest.approx(s["inertia"], rel=1e-12)
        dl = st.deltas(snap1, snap2)
        assert dl["gap"] == snap2["balance"]["gap"] - snap1["balance"]["gap"]
        assert dl["avg_inertia"] <= 0.0 and int(dl["counts"].sum()) == 0   # Lloyd never loosens the fit
    with km.KMeansState() as st:
        st.upload_points(x)
        st.set_centroids(c0)
        with pytest.raises(km.KMeansError):
            st.cluster_inertia()   # nothing assigned yet


def test_board_export_import_round_trip(km):
    """f3: a finished run written in the demo's {cards, centroids, meta} schema and read back."""
    x = oracle.gen_blobs(6, 300, 2, 3)
    st = km.init(x, 3, seed=1)
    out = km.run(st, max_iter=50)
    board = km.board_json.loads(km.board_json.dumps(km.export_board(st, iteration=out["iterations"], inertia=out["inertia"])))
    assert [c["count"] for c in board["centroids"]] == out["counts"].tolist()
    assert board["meta"]["iteration"] == out["iterations"]
    st.close()
    st2, info = km.import_board(board)
    assert np.array_equal(info["points"], x) and np.array_equal(info["labels"], out["labels"])
    s = st2.assign()   # the imported centroids are the converged ones: nothing is left to move
    assert np.array_equal(st2.labels(), out["labels"]) and s["inertia"] == out["inertia"]
    st2.close()


def test_labels_as_bytes(km):
    # km_get_labels_u8: the same labels, a byte each — plain store, cell-sorted store, ragged n, fresh labels, k > 255
    for n, d, k, kw in [(10_003, 3, 7, {}), (40_001, 2, 64, {"prune_min_points": 1}), (5, 2, 2, {})]:
        x = oracle.gen_blobs(21, n, d, k)
        c0 = oracle.forgy_init(x, k, 3)
        with km.KMeansState(**kw) as st:
            st.upload_points(x)
            st.set_centroids(c0)
            assert np.all(st.labels(dtype=np.uint8) == 255)          # never assigned
            st.run(12)
            a, b = st.labels(), st.labels(dtype=np.uint8)
            assert b.dtype == np.uint8 and np.array_equal(a, b.astype(np.int32))
            out = np.empty(n, dtype=np.uint8)
            assert st.labels(out=out) is out and np.array_equal(out, b)
    x = oracle.gen_blobs(22, 2000, 2, 300)
    with km.KMeansState() as st:
        st.upload_points(x)
        st.set_centroids(oracle.forgy_init(x, 300, 3))
        with pytest.raises(km.KMeansError):
            st.labels(dtype=np.uint8)


def test_errors(km):
    L = km._lib
    with km.KMeansState() as st:
        with pytest.raises(km.KMeansError) as e:
            st.set_centroids(np.zeros((2, 2)))
        assert e.value.status == L.ERR_STATE
        with pytest.raises(km.KMeansError) as e:
            st.run(5)
        assert e.value.status == L.ERR_STATE
        bad = np.ones((10, 2), dtype=np.float32)
        bad[3, 1] = np.nan
        with pytest.raises(km.KMeansError) as e:
            st.upload_points(bad)
        assert e.value.status == L.ERR_NONFINITE
        st.upload_points(np.ones((10, 2), dtype=np.float32))
        with pytest.raises(km.KMeansError) as e:
            st.set_centroids(np.array([[np.inf, 0.0]]))
        assert e.value.status == L.ERR_NONFINITE
        with pytest.raises(km.KMeansError) as e:
            st.init_forgy(11, 1)
        assert e.value.status == L.ERR_INVALID_ARG
        st.set_centroids(np.zeros((2, 2)))
        with pytest.raises(km.KMeansError) as e:
            st.update()  # nothing assigned yet
        assert e.value.status == L.ERR_STATE
        with pytest.raises(km.KMeansError) as e:
            st.run(0)
        assert e.value.status == L.ERR_INVALID_ARG
    with pytest.raises(km.KMeansError):
        km.KMeansState(device=99)


# ---------------------------------------------------------------- BASELINE.json sizes
def test_config2_full_size_bit_exact(km):
    """configs[1]: N=1e6, d=2, k=8 — full size, the oracle needs ~1 s."""
    x = oracle.gen_blobs(2, 1_000_000, 2, 8)
    c0 = oracle.forgy_init(x, 8, 2)
    g = gpu_run(km, x, c0, 30)
    o = oracle.run(x, c0, max_iter=30)
    assert g["iterations"] == o["iterations"] and np.array_equal(g["labels"], o["labels"])
    assert close(g["centroids"], o["centroids"]) and g["inertia"] == pytest.approx(o["inertia"], rel=REL)
    e = oracle.run(x, c0, max_iter=30, sum_mode=oracle.SUM_EXACT, nthreads=0)
    assert np.array_equal(e["centroids"], g["centroids"]) and np.array_equal(e["labels"], g["labels"])


def test_large_properties_d2_k64(km):
    """configs[4] per-GPU shape scaled to what one test may take: size-independent
    properties — inertia never rises, counts sum to N, convergence is a fixed point,
    incremental == full — plus label parity of every point after 13 iterations
    (the store is cell-sorted and dealt out in round-robin chunks by then)."""
    n = 8_000_000
    with km.KMeansState() as st, km.KMeansState(update_mode=1) as full:
        for s in (st, full):
            s.generate_blobs(n, 2, 64, 5)
            s.init_forgy(64, 5)
        prev = np.inf
        for it in range(12):
            a, b = st.step(), full.step()
            assert a["inertia"] == b["inertia"] and a["changed"] == b["changed"]
            assert a["inertia"] <= prev
            prev = a["inertia"]
            assert int(st.counts().sum()) == n
        assert np.array_equal(st.centroids(), full.centroids())
        # the labels of a prefix equal the oracle's argmin under the centroids that made them
        c_before = st.centroids()
        st.assign()
        lab = st.labels()
        xs = st.get_points()
        assert np.array_equal(xs[:100000], oracle.gen_blobs(5, 100000, 2, 64))   # rows come back in generation order
        olab, _, _ = oracle.assign(xs, c_before, nthreads=0)
        assert np.array_equal(lab, olab)


@pytest.mark.parametrize("case", ["k8_d2", "k5_d3", "grid_on_borrowed_buffer"])
def test_many_rounds_per_block(km, case):
    """Shapes large enough that every block walks several rounds (the software pipeline, ragged last
    rounds of the last blocks) in the kernel variants the small tests reach with one round."""
    if case == "grid_on_borrowed_buffer":
        import torch
        x = oracle.gen_blobs(3, 6_000_000, 2, 64)
        c0 = oracle.forgy_init(x, 64, 4)
        xt = torch.from_numpy(x).cuda()
        with km.KMeansState(prune_min_points=1) as st:
            st.attach_points(xt.data_ptr(), x.shape[0], 2)   # borrowed: never re-ordered -> grid over caller order
            st.set_centroids(c0)
            s = st.run(5)
            lab = st.labels()
    else:
        n, d, k = (4_000_001, 2, 8) if case == "k8_d2" else (3_000_003, 3, 5)
        x = oracle.gen_blobs(3, n, d, k)
        c0 = oracle.forgy_init(x, k, 4)
        g = gpu_run(km, x, c0, 5)
        s, lab = g, g["labels"]
    o = oracle.run(x, c0, max_iter=5, nthreads=0)
    assert s["iterations"] == o["iterations"] and np.array_equal(lab, o["labels"])


# ---------------------------------------------------------------- d = 2 candidate-grid pruning
def pruned_run(km, x, c0, max_iter=60, update_mode=0):
    return gpu_run(km, x, c0, max_iter, "auto", update_mode, prune_min_points=1)


@pytest.mark.parametrize("mode", [0, 1], ids=["incremental", "full"])
@pytest.mark.parametrize("shape", [(12000, 64, 64), (30000, 16, 16), (9000, 5, 200), (20000, 256, 256), (4000, 3, 40)],
                         ids=lambda s: "n%d_kt%d_k%d" % s)
def test_pruned_matches_oracle_and_brute_force(km, shape, mode):
    n, k_true, k = shape
    x = oracle.gen_blobs(300 + k, n, 2, k_true)
    c0 = oracle.forgy_init(x, k, 5)
    g = pruned_run(km, x, c0, 50, mode)
    b = gpu_run(km, x, c0, 50, "auto", mode, no_prune=True)
    o = oracle.run(x, c0, max_iter=50)
    e = oracle.run(x, c0, max_iter=50, sum_mode=oracle.SUM_EXACT)
    assert g["iterations"] == o["iterations"] == b["iterations"]
    assert np.array_equal(g["labels"], o["labels"]) and np.array_equal(b["labels"], o["labels"])
    assert np.array_equal(g["centroids"], e["centroids"]) and np.array_equal(b["centroids"], e["centroids"])
    assert g["inertia"] == pytest.approx(o["inertia"], rel=REL)
    assert g["rescored"] <= max(b["rescored"], 0) + n // 50  # overflowed cells add a few full scans, not many


def test_pruned_adversarial_geometry(km):
    rng = np.random.default_rng(12)
    # duplicate centroids, a far outlier centroid, points on cell boundaries and exact lattice ties
    xs = np.array([[x, y] for x in np.arange(-8, 8.25, 0.25) for y in np.arange(-8, 8.25, 0.25)], dtype=np.float32)
    xs = np.concatenate([xs, (rng.standard_normal((5000, 2)) * 3).astype(np.float32)])
    c0 = np.array([[x, y] for x in (-6, -2, 2, 6) for y in (-6, -2, 2, 6)] + [[2, 2], [500, 500], [-6, -6]], dtype=np.float64)
    g = pruned_run(km, xs, c0, 30)
    o = oracle.run(xs, c0, max_iter=30)
    assert g["iterations"] == o["iterations"] and np.array_equal(g["labels"], o["labels"])
    assert close(g["centroids"], o["centroids"]) and g["inertia"] == pytest.approx(o["inertia"], rel=REL)
    # all k centroids on one spot: every cell overflows its 7-candidate word -> full scans, still exact
    c_same = np.tile(np.array([[1.0, 1.0]]), (20, 1))
    g = pruned_run(km, xs, c_same, 5)
    o = oracle.run(xs, c_same, max_iter=5)
    assert np.array_equal(g["labels"], o["labels"]) and g["iterations"] == o["iterations"]
    # degenerate box (all x equal): pruning must switch itself off, not divide by zero
    xl = np.stack([np.full(3000, 2.5, np.float32), rng.standard_normal(3000).astype(np.float32)], axis=1)
    cl = oracle.forgy_init(xl, 16, 1)
    g = pruned_run(km, xl, cl, 20)
    o = oracle.run(xl, cl, max_iter=20)
    assert np.array_equal(g["labels"], o["labels"]) and g["iterations"] == o["iterations"]
    # data far from the origin relative to its spread
    xf = (rng.standard_normal((20000, 2)) + np.array([4000.0, -9000.0])).astype(np.float32)
    cf = oracle.forgy_init(xf, 32, 2)
    g = pruned_run(km, xf, cf, 25)
    o = oracle.run(xf, cf, max_iter=25)
    assert np.array_equal(g["labels"], o["labels"]) and g["iterations"] == o["iterations"]
    assert close(g["centroids"], o["centroids"])


def test_sorted_store_is_invisible(km):
    """The pruned path re-orders its copy of the points by grid cell before the third pass on an
    upload (km_sort.cu).  Labels, rows and Forgy picks must keep coming back in the caller's order,
    before and after, and a borrowed (attached) buffer is never re-ordered."""
    n, k = 50000, 64
    x = oracle.gen_blobs(41, n, 2, k)
    c0 = oracle.forgy_init(x, k, 9)
    with km.KMeansState(prune_min_points=1) as st:
        st.upload_points(x)
        st.set_centroids(c0)
        c = c0.copy()
        lab_prev = np.full(n, -1, np.int32)
        for it in range(6):   # passes 1-2 run on the uploaded order, 3+ on the sorted store
            s = st.assign()
            olab, _, _ = oracle.assign(x, c)
            assert np.array_equal(st.labels(), olab), "pass %d" % it
            assert s["changed"] == int(np.sum(olab != lab_prev))
            lab_prev = olab
            st.update()
            c, _ = oracle.update(x, olab, c, sum_mode=oracle.SUM_EXACT)
            assert np.array_equal(st.centroids(), c)
        assert np.array_equal(st.get_points(), x)
        assert np.array_equal(st.get_points(1234, 777), x[1234:2011])
        st.init_forgy(k, 9)   # picks rows by their ORIGINAL index
        assert np.array_equal(st.centroids(), c0)
        # a second upload of the same shape starts over in the caller's order
        x2 = oracle.gen_blobs(42, n, 2, k)
        st.upload_points(x2)
        st.set_centroids(c0)
        r = st.run(40)
        o = oracle.run(x2, c0, max_iter=40)
        assert r["iterations"] == o["iterations"] and np.array_equal(st.labels(), o["labels"])
        assert np.array_equal(st.get_points(), x2)


def test_pruned_attached_points_stay_unsorted(km):
    import torch
    n, k = 40000, 32
    x = oracle.gen_blobs(77, n, 2, k)
    c0 = oracle.forgy_init(x, k, 2)
    xt = torch.from_numpy(x).cuda()
    with km.KMeansState(prune_min_points=1) as st:
        st.attach_points(xt.data_ptr(), n, 2)
        st.set_centroids(c0)
        r = st.run(40)
        o = oracle.run(x, c0, max_iter=40)
        assert r["iterations"] == o["iterations"] and np.array_equal(st.labels(), o["labels"])
        assert close(st.centroids(), o["centroids"])
    assert np.array_equal(xt.cpu().numpy(), x)


# ---------------------------------------------------------------- tensor path (tcgen05 coarse stage + exact stage)
@pytest.mark.parametrize("mode", [0, 1], ids=["incremental", "full"])
@pytest.mark.parametrize("shape", [(3000, 32, 16, 16), (1500, 128, 8, 32), (20000, 64, 40, 200), (4099, 32, 3, 129),
                                   (9000, 96, 500, 1024), (257, 128, 2, 2), (6000, 16, 12, 40), (5003, 24, 9, 130),
                                   (3001, 9, 5, 17), (2500, 77, 6, 300), (70000, 48, 64, 64),
                                   # rows without spare columns: h rides in the extra no-swizzle K-step (d = 49..64, 113..128)
                                   (2500, 120, 5, 140), (3000, 50, 7, 130), (1031, 113, 3, 300), (777, 64, 4, 129)],
                         ids=lambda s: "n%d_d%d_kt%d_k%d" % s)
def test_tensor_path_matches_oracle(km, shape, mode):
    n, d, k_true, k = shape
    x = oracle.gen_blobs(700 + d, n, d, k_true)
    c0 = oracle.forgy_init(x, k, 9)
    with km.KMeansState(path=km._lib.PATH_TENSOR, update_mode=mode) as st:
        st.upload_points(x)
        st.set_centroids(c0)
        s = st.run(25)
        assert s["path_used"] == km._lib.PATH_TENSOR
        lab, cen, counts = st.labels(), st.centroids(), st.counts()
    o = oracle.run(x, c0, max_iter=25, nthreads=0)
    e = oracle.run(x, c0, max_iter=25, sum_mode=oracle.SUM_EXACT, nthreads=0)
    assert s["iterations"] == o["iterations"] and np.array_equal(lab, o["labels"])
    assert np.array_equal(cen, e["centroids"]) and close(cen, o["centroids"])
    assert np.array_equal(counts, o["counts"])
    assert s["inertia"] == pytest.approx(o["inertia"], rel=REL)


def test_tensor_path_near_ties_and_duplicates(km):
    # many centroids per blob (k >> k_true) makes within-blob near-ties the common case, the
    # regime where the BF16 coarse stage cannot decide and the exact stage must; plus exact duplicates
    rng = np.random.default_rng(3)
    x = oracle.gen_blobs(91, 6000, 64, 4)
    c0 = oracle.forgy_init(x, 96, 4)
    c0[50] = c0[7]                                  # duplicate centroid: index 7 must win every tie
    x[100] = ((c0[3] + c0[11]) / 2).astype(np.float32)   # a point (almost) on a bisector
    with km.KMeansState(path=km._lib.PATH_TENSOR) as st:
        st.upload_points(x)
        st.set_centroids(c0)
        s = st.assign()
        lab, _, inertia = oracle.assign(x, c0)
        assert np.array_equal(st.labels(), lab) and not np.any(st.labels() == 50)
        assert s["inertia"] == pytest.approx(inertia, rel=REL)
    check_against_oracle(km, x, c0, max_iter=20, path="auto")


def test_tensor_path_rejects_unsupported_shapes(km):
    x = oracle.gen_blobs(1, 500, 200, 3)   # d beyond the 128 columns of one operand row
    with km.KMeansState(path=km._lib.PATH_TENSOR) as st:
        st.upload_points(x)
        with pytest.raises(km.KMeansError) as e:
            st.set_centroids(x[:4].astype(np.float64))
        assert e.value.status == km._lib.ERR_UNSUPPORTED
