"""Development check (run by hand on a GPU box, not collected by pytest): multi-million-point
shapes of every small-d kernel variant against the oracle — many rounds per block, the
cell-sorted store, the caller-order grid on a borrowed buffer.

    python tests/dev_large_shapes_check.py [all|noprune|d3|grid|unsorted]
"""
import sys, time
import numpy as np
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import kmeans_b200 as km
import oracle

def run(n, d, k, k_true, **kw):
    t = time.time()
    x = oracle.gen_blobs(3, n, d, k_true)
    c0 = oracle.forgy_init(x, k, 4)
    with km.KMeansState(**kw) as st:
        st.upload_points(x)
        st.set_centroids(c0)
        s = st.run(5)
        lab = st.labels()
    o = oracle.run(x, c0, max_iter=5)
    print(n, d, k, kw, "ok" if np.array_equal(lab, o["labels"]) and s["iterations"] == o["iterations"] else "MISMATCH",
          "%.1fs" % (time.time() - t), flush=True)

which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "noprune"):
    run(4_000_000, 2, 8, 8)            # PRUNE=0, ~7 rounds per block
if which in ("all", "d3"):
    run(3_000_000, 3, 5, 5)
if which in ("all", "grid"):
    run(6_000_000, 2, 64, 64, prune_min_points=1)   # km_run(5) sorts at start -> PRUNE=2
if which in ("all", "unsorted"):
    import torch
    x = oracle.gen_blobs(3, 6_000_000, 2, 64)
    c0 = oracle.forgy_init(x, 64, 4)
    xt = torch.from_numpy(x).cuda()
    with km.KMeansState(prune_min_points=1) as st:
        st.attach_points(xt.data_ptr(), x.shape[0], 2)   # borrowed: never sorted -> PRUNE=1
        st.set_centroids(c0)
        s = st.run(5)
        lab = st.labels()
    o = oracle.run(x, c0, max_iter=5)
    print("unsorted grid", "ok" if np.array_equal(lab, o["labels"]) else "MISMATCH", flush=True)
