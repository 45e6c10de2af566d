"""km_run through the resident kernel at several iteration caps: python profiles/resident_prof.py n d k"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmeans_b200 as km
n, d, k = (int(v) for v in sys.argv[1:4])
for mode in (0, 1):
    with km.KMeansState(update_mode=mode) as st:
        st.generate_blobs(n, d, k, 4)
        st.init_forgy(k, 4)
        c0 = st.centroids()
        for cap in (1, 1, 2, 3, 5, 10, 20, 40):
            st.set_centroids(c0)
            s = st.run(cap)
            print(mode, cap, s["iterations"], round(s["device_ms"] * 1000, 1), "us", s["changed"])
