import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import kmeans_b200 as km
with km.KMeansState() as st:
    st.generate_blobs(1000000, 2, 8, 4)
    st.init_forgy(8, 4)
    c0 = st.centroids()
    for cap in (40, 40):
        st.set_centroids(c0)
        s = st.run(cap)
        print(cap, s["iterations"], round(s["device_ms"] * 1000, 1), "us", flush=True)
