"""Run under torchrun (one rank per GPU): the sharded run must equal the single-rank oracle
bit for bit — labels, iteration count, centroids (exact-sum mode) — whatever the world size.
Rank 0 prints MULTI_GPU_OK on success.  Used by tests/test_gpu_multi.py and by hand:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multi_gpu_worker.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kmeans_b200 as km  # noqa: E402
import oracle  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    uid = torch.zeros(128, dtype=torch.uint8, device=dev)
    cases = [  # n, d, k, path, update_mode
        (200_003, 2, 8, km._lib.PATH_AUTO, 0), (200_003, 2, 8, km._lib.PATH_AUTO, 1),
        (50_001, 3, 5, km._lib.PATH_DIRECT_F64, 0), (9_000, 32, 16, km._lib.PATH_AUTO, 0), (64_000, 2, 64, km._lib.PATH_AUTO, 0),
        (20_000, 8, 700, km._lib.PATH_AUTO, 0),   # small-d path with a 95 KB table: above the peer all-reduce's limit, NCCL
    ]
    for ci, (n, d, k, path, mode) in enumerate(cases):
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(km.KMeansState.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        lo, hi = km.shard_range(n, rank, world)
        with km.KMeansState(device=local, path=path, update_mode=mode) as st:
            st.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
            x_local = oracle.gen_blobs(40 + ci, hi - lo, d, k, i0=lo)
            st.upload_points(x_local)
            st.init_kmeanspp(k, 13)
            cpp = st.centroids()
            st.init_forgy(k, 11)
            c0 = st.centroids()
            s = st.run(40)
            how = st.comm_info()[2]
            # again from the same start (set_centroids clears the tables: a barrier across the ranks on the peer path),
            # then single steps with a download in between
            cent1 = st.centroids()
            st.set_centroids(c0)
            s2 = st.run(40)
            assert s2["iterations"] == s["iterations"] and s2["inertia"] == s["inertia"], (s, s2)
            assert np.array_equal(st.centroids(), cent1)
            st.set_centroids(c0)
            for _ in range(3):
                st.step()
                st.labels()
            st.set_centroids(c0)
            s = st.run(40)
            labels = torch.from_numpy(st.labels()).to(dev)
            cent = st.centroids()
            counts = st.counts()
        sizes = [km.shard_range(n, r, world)[1] - km.shard_range(n, r, world)[0] for r in range(world)]
        gathered = [torch.empty(sz, dtype=torch.int32, device=dev) for sz in sizes]
        dist.all_gather(gathered, labels)
        if rank == 0:
            x = oracle.gen_blobs(40 + ci, n, d, k)
            assert np.array_equal(c0, oracle.forgy_init(x, k, 11)), "forgy rows differ"
            assert np.array_equal(cpp, oracle.kmeanspp_init(x, k, 13)), "k-means++ rows differ"
            ref = oracle.run(x, c0, max_iter=40)
            ex = oracle.run(x, c0, max_iter=40, sum_mode=oracle.SUM_EXACT, nthreads=0)
            got = torch.cat(gathered).cpu().numpy()
            assert s["iterations"] == ref["iterations"], (s["iterations"], ref["iterations"])
            assert np.array_equal(got, ref["labels"]), "labels differ from the single-rank oracle"
            assert np.array_equal(cent, ex["centroids"]), "centroids differ from the exact-sum oracle"
            assert np.array_equal(counts, ref["counts"])
            assert abs(s["inertia"] - ref["inertia"]) <= 1e-9 * ref["inertia"]
            assert s["n_total"] == n
            print(f"case {ci}: world {world} n={n} d={d} k={k} iterations={s['iterations']} "
                  f"collective={['nccl', 'peer regions mapped, nccl for this table', 'peer kernel'][how]} ok", flush=True)
        dist.barrier()
    if rank == 0:
        print("MULTI_GPU_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
