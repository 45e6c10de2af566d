"""Development aid: host wall time of the one-time cell sort and of km_run on the configs[4] shard."""
import sys, time
sys.path.insert(0, ".")
import kmeans_b200 as km

n = 125_000_000
with km.KMeansState() as st:
    st.generate_blobs(n, 2, 64, 5)
    st.init_forgy(64, 5)
    for it in (3, 50, 50):
        t = time.time(); s = st.run(it); dt = time.time() - t
        print("run(%d): %.1f ms wall, device_ms %.1f, iterations %d" % (it, dt * 1e3, s["device_ms"], s["iterations"]), flush=True)
    t = time.time(); lab = st.labels(); print("labels(): %.1f ms" % ((time.time() - t) * 1e3), flush=True)
    st.generate_blobs(n, 2, 64, 5); st.init_forgy(64, 5)
    t = time.time(); s = st.run(50); print("fresh run(50) incl. sort: %.1f ms" % ((time.time() - t) * 1e3), flush=True)
