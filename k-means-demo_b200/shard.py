"""Point sharding across ranks: contiguous index ranges, no point ever
moves after upload.  Rank g owns [g*ceil(N/G), min(N, (g+1)*ceil(N/G)))."""


def shard_range(n_total, rank, world):
    if world < 1 or not (0 <= rank < world) or n_total < 0:
        raise ValueError("need world >= 1, 0 <= rank < world, n_total >= 0")
    per = -(-n_total // world)
    lo = min(n_total, rank * per)
    hi = min(n_total, (rank + 1) * per)
    return lo, hi
