"""GPU: km_run of a job that fits in shared memory runs as ONE cooperative launch (km_resident.cu).
It must be indistinguishable from the stepwise path and from the oracle: labels, iteration count,
counts, centroids (bit-identical to the exact-sum oracle), inertia — and the state it leaves behind
must carry on under km_step exactly as if the iterations had been issued one by one."""
import numpy as np
import pytest

import oracle

gpu = pytest.mark.gpu

SHAPES = [(100_003, 2, 8), (50_000, 3, 5), (5_000, 8, 4), (1_000_000, 2, 8), (777, 1, 3), (31, 2, 8), (200_000, 4, 8)]


@gpu
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "n%d_d%d_k%d" % s)
def test_resident_run_equals_oracle_and_stepwise(km, shape):
    n, d, k = shape
    x = oracle.gen_blobs(7 + d, n, d, k)
    c0 = oracle.forgy_init(x, k, 5)
    o = oracle.run(x, c0, max_iter=25)
    e = oracle.run(x, c0, max_iter=25, sum_mode=oracle.SUM_EXACT, nthreads=0)
    res = {}
    for mode in (0, 1):   # 0: incremental -> resident launch; 1: full -> one launch per iteration
        with km.KMeansState(update_mode=mode) as st:
            st.upload_points(x)
            st.set_centroids(c0)
            l0 = st.launch_count()
            s = st.run(25)
            res[mode] = dict(s=s, labels=st.labels(), cent=st.centroids(), counts=st.counts(), sums=st.sums(),
                             launches=st.launch_count() - l0)
    r, w = res[0], res[1]
    assert r["launches"] <= 2, "the resident path is the label reset and one launch"
    assert w["launches"] >= r["s"]["iterations"]
    for g in (r, w):
        assert g["s"]["iterations"] == o["iterations"]
        assert g["s"]["changed"] == o["changed"]
        assert np.array_equal(g["labels"], o["labels"])
        assert np.array_equal(g["counts"], o["counts"])
        assert np.array_equal(g["cent"], e["centroids"])
        assert g["s"]["inertia"] == pytest.approx(o["inertia"], rel=1e-9)
    assert r["s"]["inertia"] == w["s"]["inertia"]
    for a, b in zip(r["sums"], w["sums"]):
        assert np.array_equal(a, b)


@gpu
def test_resident_state_carries_on_stepwise(km):
    n, d, k = 60_000, 2, 8
    x = oracle.gen_blobs(3, n, d, k)
    c0 = oracle.forgy_init(x, k, 9)
    outs = []
    for mode in (0, 1):
        with km.KMeansState(update_mode=mode) as st:
            st.upload_points(x)
            st.set_centroids(c0)
            st.run(4)                      # stops at max_iter, not converged
            seq = []
            for _ in range(3):
                s = st.step()
                seq.append((s["iterations"], s["changed"], s["inertia"]))
            outs.append((seq, st.labels(), st.centroids(), st.counts()))
    assert outs[0][0] == outs[1][0]
    assert np.array_equal(outs[0][1], outs[1][1])
    assert np.array_equal(outs[0][2], outs[1][2])
    assert np.array_equal(outs[0][3], outs[1][3])
    # and again from the top: a second km_run on the same state starts from unlabelled points
    with km.KMeansState() as st:
        st.upload_points(x)
        st.set_centroids(c0)
        a = st.run(30)
        ca = st.centroids()
        st.set_centroids(c0)
        b = st.run(30)
        assert a["iterations"] == b["iterations"] and a["inertia"] == b["inertia"]
        assert np.array_equal(ca, st.centroids())


@gpu
def test_comm_info_single_rank(km):
    with km.KMeansState() as st:
        assert st.comm_info() == (0, 1, 0)


@gpu
def test_resident_ties_empty_clusters_and_max_iter(km):
    # exact ties (lowest index wins), a centroid nobody picks (keeps its place), max_iter = 1
    x = np.array([[0, 0], [2, 0], [1, 0], [1, 0], [10, 10]], dtype=np.float32)
    c0 = np.array([[0, 0], [2, 0], [100, 100], [2, 0]], dtype=np.float64)
    for it in (1, 2, 10):
        o = oracle.run(x, c0, max_iter=it)
        with km.KMeansState() as st:
            st.upload_points(x)
            st.set_centroids(c0)
            s = st.run(it)
            assert s["iterations"] == o["iterations"]
            assert np.array_equal(st.labels(), o["labels"])
            assert np.array_equal(st.centroids(), o["centroids"])
            assert s["inertia"] == pytest.approx(o["inertia"], rel=1e-9, abs=1e-12)
