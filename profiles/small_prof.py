"""Small driver for ncu captures of the small-d path: python profiles/small_prof.py n d k iters [k_true]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmeans_b200 as km
n, d, k, iters = (int(v) for v in sys.argv[1:5])
k_true = int(sys.argv[5]) if len(sys.argv) > 5 else k
with km.KMeansState() as st:
    st.generate_blobs(n, d, k_true, 5)
    st.init_forgy(k, 5)
    for _ in range(iters):
        st.step_async()
    print(st.stats_fetch())
