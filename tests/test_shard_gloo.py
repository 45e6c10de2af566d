"""CPU, world_size 2 over gloo: the multi-rank path's host logic and arithmetic.

What a rank does on the GPU per iteration — assign its contiguous shard, reduce the packed
int64 table {coarse limbs, fine limbs, counts, changed} with ONE all-reduce(sum), derive the
centroids locally — is replayed here with the oracle standing in for the kernels, and must
reproduce the single-rank run bit for bit (labels, centroids, iteration count).  That is the
property SURVEY.md §8e asks for ("identical at 1/2/4/8 GPUs") and the reason the sums are
integers.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, d, k, seed, max_iter, out_dir):
    sys.path.insert(0, ROOT)
    import kmeans_b200 as km
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = km.shard_range(n, rank, world)
    x = oracle.gen_blobs(seed, hi - lo, d, k, i0=lo)          # this rank's rows only
    # job-wide scale: max over ranks of max|x| (the library all-gathers the same two numbers)
    m = torch.tensor([oracle.maxabs(x)], dtype=torch.float64)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    maxabs = float(m.item())
    # Forgy rows: each picked row is owned by exactly one rank, the others add exact zeros
    idx = oracle.forgy_indices(n, k, 7)
    c = torch.zeros((k, d), dtype=torch.float64)
    for s_, g in enumerate(idx):
        if lo <= g < hi:
            c[s_] = torch.from_numpy(x[g - lo].astype(np.float64))
    dist.all_reduce(c, op=dist.ReduceOp.SUM)
    c = c.numpy().copy()
    labels = np.full(hi - lo, -1, dtype=np.int32)
    it = 0
    while True:
        labels, changed, _ = oracle.assign(x, c, labels)
        sa, sb, cnt = oracle.partial_sums_exact(x, labels, k, maxabs, n)
        packed = torch.from_numpy(np.concatenate([sa.ravel(), sb.ravel(), cnt, np.array([changed], dtype=np.int64)]))
        dist.all_reduce(packed, op=dist.ReduceOp.SUM)          # the iteration's one collective
        p = packed.numpy()
        it += 1
        if p[-1] == 0 or it >= max_iter:
            break
        c = oracle.centroids_from_limbs(p[:k * d].reshape(k, d), p[k * d:2 * k * d].reshape(k, d),
                                        p[2 * k * d:2 * k * d + k], c, maxabs, n)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), labels=labels, centroids=c, it=it, lo=lo, hi=hi)
    dist.destroy_process_group()


@pytest.mark.parametrize("n,d,k", [(6001, 2, 8), (3000, 5, 4)])
def test_two_ranks_reproduce_one_rank_bit_for_bit(tmp_path, n, d, k):
    import oracle
    world, seed, max_iter = 2, 13, 40
    mp.spawn(_worker, args=(world, _free_port(), n, d, k, seed, max_iter, str(tmp_path)), nprocs=world, join=True)
    x = oracle.gen_blobs(seed, n, d, k)
    c0 = x[oracle.forgy_indices(n, k, 7)].astype(np.float64)
    ref = oracle.run(x, c0, max_iter=max_iter, sum_mode=oracle.SUM_EXACT)
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    labels = np.concatenate([p["labels"] for p in parts])
    assert [int(p["lo"]) for p in parts] == [0, -(-n // 2)] and int(parts[-1]["hi"]) == n
    assert np.array_equal(labels, ref["labels"])
    for p in parts:
        assert int(p["it"]) == ref["iterations"]
        assert np.array_equal(p["centroids"], ref["centroids"])
    # and the plain sequential-sum restatement agrees on labels and iteration count
    seq = oracle.run(x, c0, max_iter=max_iter)
    assert np.array_equal(labels, seq["labels"]) and seq["iterations"] == ref["iterations"]


def _kpp_worker(rank, world, port, n, d, k, seed, out_dir):
    """The sharded k-means++ draw as km_init_kmeanspp does it: per-rank integer totals of the
    quantised D^2, one all-reduce of the list of totals, the owner rank walks its own prefix."""
    import math
    sys.path.insert(0, ROOT)
    import kmeans_b200 as km
    import oracle
    from oracle import numpy_ref as nr
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = km.shard_range(n, rank, world)
    x = oracle.gen_blobs(seed, hi - lo, d, max(2, k // 2), i0=lo).astype(np.float64)
    m = torch.tensor([float(np.max(np.abs(x)))], dtype=torch.float64)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    E = math.frexp(float(m.item()))[1] if m.item() > 0 else 0
    qd = 62 - max(1, int(n).bit_length()) - (2 * E + 2 + int(d).bit_length())
    D = np.full(hi - lo, np.inf)
    state, picks, centre = 21, [], None
    for s_ in range(k):
        r = nr.splitmix64(state)
        state = (state + nr.GAMMA) & nr.MASK
        u = float(r >> 11) * 2.0 ** -53
        local = -1
        by_index = s_ == 0
        if s_ > 0:
            D = np.minimum(D, nr.distances(x, centre[None, :])[:, 0])
            q = [int(math.ldexp(float(v), qd)) for v in D]
            totals = torch.zeros(world, dtype=torch.int64)
            totals[rank] = sum(q)
            dist.all_reduce(totals, op=dist.ReduceOp.SUM)   # every rank fills its own slot
            tot = [int(v) for v in totals]
            if sum(tot) == 0:
                by_index = True
            else:
                target = ((r >> 11) * sum(tot)) >> 53
                owner = world - 1
                for rk in range(world):
                    if target < tot[rk]:
                        owner = rk
                        break
                    target -= tot[rk]
                if owner == rank:
                    run = 0
                    local = len(q) - 1
                    for i, v in enumerate(q):
                        run += v
                        if run > target:
                            local = i
                            break
        if by_index:
            g = int(math.floor(u * float(n)))
            if lo <= g < hi:
                local = g - lo
        row = torch.zeros(d + 1, dtype=torch.float64)        # the row and its global index; zeros elsewhere
        if local >= 0:
            row[:d] = torch.from_numpy(x[local])
            row[d] = float(lo + local)
        dist.all_reduce(row, op=dist.ReduceOp.SUM)
        centre = row[:d].numpy().copy()
        picks.append(int(row[d].item()))
    np.save(os.path.join(out_dir, f"kpp{rank}.npy"), np.array(picks, dtype=np.int64))
    dist.destroy_process_group()


@pytest.mark.parametrize("n,d,k", [(1201, 2, 9), (700, 3, 16)])
def test_two_ranks_draw_the_single_rank_kmeanspp_rows(tmp_path, n, d, k):
    import oracle
    world, seed = 2, 17
    mp.spawn(_kpp_worker, args=(world, _free_port(), n, d, k, seed, str(tmp_path)), nprocs=world, join=True)
    x = oracle.gen_blobs(seed, n, d, max(2, k // 2))
    want = oracle.kmeanspp_indices(x, k, 21)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"kpp{r}.npy"), want)


def test_unique_id_needs_nccl_only_when_asked(km):
    # loading the library and the binding must not require NCCL or a GPU
    lib = km._lib.load()
    assert hasattr(lib, "km_comm_init") and hasattr(lib, "km_comm_unique_id")
