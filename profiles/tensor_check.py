import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import kmeans_b200 as km, oracle
L = km._lib
for (n, d, kt, k) in [(3000, 32, 16, 16), (1500, 128, 8, 32), (20000, 64, 40, 200), (50000, 128, 100, 300), (30000, 96, 1000, 1024), (4099, 32, 3, 129)]:
    x = oracle.gen_blobs(7 + d, n, d, kt)
    c0 = oracle.forgy_init(x, k, 3)
    with km.KMeansState(path=L.PATH_TENSOR) as st:
        st.upload_points(x); st.set_centroids(c0)
        t = time.time(); s = st.run(30); dt = time.time() - t
        lab = st.labels(); cen = st.centroids()
    o = oracle.run(x, c0, max_iter=30, nthreads=0)
    e = oracle.run(x, c0, max_iter=30, sum_mode=oracle.SUM_EXACT, nthreads=0)
    print(n, d, k, "path", s["path_used"], "it", s["iterations"], o["iterations"], "labels", np.array_equal(lab, o["labels"]),
          "centroids==exact", np.array_equal(cen, e["centroids"]), "inertia rel", abs(s["inertia"] - o["inertia"]) / o["inertia"],
          "rescored", s["rescored"], "ms", round(1e3 * dt, 1), "dev_ms", round(s["device_ms"], 2), flush=True)
