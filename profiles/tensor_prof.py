"""Small driver for ncu captures of the tensor path: python profiles/tensor_prof.py n d k iters"""
import sys
sys.path.insert(0, "/root/repo")
import kmeans_b200 as km
n, d, k, iters = (int(v) for v in sys.argv[1:5])
with km.KMeansState(path=km._lib.PATH_TENSOR) as st:
    st.generate_blobs(n, d, k, 3)
    st.init_forgy(k, 3)
    for _ in range(iters):
        s = st.step()
    print(s)
