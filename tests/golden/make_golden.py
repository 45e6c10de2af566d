"""Regenerate tests/golden/*.npz.

    python tests/golden/make_golden.py

The reference (schusto/k-means-demo) has no golden vectors, known-answer tests or
fixtures for this path — it has no Lloyd code at all ("parity
unpinned").  These fixtures are therefore produced by oracle/numpy_ref.py, an independent
NumPy statement of SURVEY.md Appendix A, and pin (i) the C oracle and (ii) the CUDA path
against something neither of them shares code with.  appendix_b.json holds the only
reference-derived known answers there are: the dashboard metrics of app.mjs:436-496
executed by hand on the reference's own seed cards (app.mjs:188, 204-216).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import numpy_ref as nr  # noqa: E402

CASES = [
    # name, seed, n, d, k_true, k, init_seed, max_iter
    ("blobs_d2_k8", 11, 600, 2, 8, 8, 3, 100),
    ("blobs_d1_k3", 12, 257, 1, 3, 3, 4, 100),
    ("blobs_d3_k5", 13, 400, 3, 5, 5, 5, 100),
    ("blobs_d4_k6", 14, 333, 4, 4, 6, 6, 100),
    ("blobs_d8_k7", 15, 300, 8, 7, 7, 7, 100),
    ("blobs_d32_k16", 16, 300, 32, 16, 16, 8, 100),
    ("blobs_d2_k64", 17, 900, 2, 64, 64, 9, 60),
    ("capped_d2_k8", 18, 500, 2, 3, 8, 10, 3),   # stops on max_iter, not convergence
]


def main():
    for name, seed, n, d, k_true, k, init_seed, max_iter in CASES:
        x = nr.gen_blobs(seed, n, d, k_true)
        idx = nr.forgy_indices(n, k, init_seed)
        c0 = x[idx].astype(np.float64)
        r = nr.run(x, c0, max_iter=max_iter)
        lab1, changed1, inertia1 = nr.assign(x, c0)
        c1, counts1 = nr.update(x, lab1, c0)
        kpp = nr.kmeanspp_indices(x, k, init_seed)   # A7b: seeded k-means++ picks for the same data
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"), x=x, c0=c0, forgy_idx=idx, kpp_idx=kpp, labels=r["labels"],
            centroids=r["centroids"], iterations=r["iterations"], inertia=r["inertia"], counts=r["counts"],
            changed=r["changed"], max_iter=max_iter, seed=seed, k_true=k_true, init_seed=init_seed,
            first_labels=lab1, first_inertia=inertia1, first_centroids=c1, first_counts=counts1)
        print(name, "iterations", r["iterations"], "inertia", r["inertia"])
    # closed-form cases of SURVEY.md A9
    a9 = [
        dict(name="a9_1", x=[[0], [1], [10], [11]], c0=[[0], [10]], labels=[0, 0, 1, 1], centroids=[[0.5], [10.5]],
             inertia=1.0, iterations=2),
        dict(name="a9_2_tie", x=[[5]], c0=[[0], [10]], labels=[0], centroids=[[5], [10]], inertia=0.0, iterations=2),
        dict(name="a9_3_empty", x=[[0], [1]], c0=[[0.5], [100]], labels=[0, 0], centroids=[[0.5], [100]],
             inertia=0.5, iterations=2),
        dict(name="a9_4_k_eq_n", x=[[0, 0], [3, 4], [-2, 7]], c0=[[0, 0], [3, 4], [-2, 7]], labels=[0, 1, 2],
             centroids=[[0, 0], [3, 4], [-2, 7]], inertia=0.0, iterations=2),
    ]
    with open(os.path.join(HERE, "a9.json"), "w") as f:
        json.dump(a9, f, indent=1)


if __name__ == "__main__":
    main()
