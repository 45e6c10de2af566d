"""GPU: km_lloyd_step_noalloc — the entry point for hosts that may not allocate.

Some callers hand a library a fixed set of device buffers and forbid it to allocate after
start-up (a process whose device memory belongs to its own pool, a loop that cannot take the
stall a cudaMalloc may cause, a sandbox capped at its declared footprint).  For them the
library offers one Lloyd iteration whose only working memory is an array built into the
library image.  This test drives it through the C ABI on buffers it allocates itself (torch,
standing in for the host's pool) and checks every output against the existing context entry
points (km_step on the binary64 path with the full update) on two calls in a row — the second
call reuses the shared array the first one dirtied.  It is named to be collected last, and it
must pass both alone in a fresh process and after every other test.
"""
import ctypes as C

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

REL = 1e-9  # same tolerance as tests/test_gpu_parity.py


def _call(lib, L, xt, n, d, c_in, k, labels, c_out, counts, res):
    st = lib.km_lloyd_step_noalloc(0, None, xt.data_ptr(), n, d, c_in.data_ptr(), k, labels.data_ptr(),
                                   c_out.data_ptr(), counts.data_ptr(), res.data_ptr())
    assert st == L.OK, (lib.km_last_error(None) or b"").decode()


@pytest.mark.parametrize("n,d,k", [(20000, 2, 8), (3001, 5, 7), (1500, 32, 16)])
def test_noalloc_step_matches_context_step_twice(km, n, d, k):
    import torch
    L = km._lib
    lib = L.load()
    x = oracle.gen_blobs(50 + d, n, d, k)
    c0 = oracle.forgy_init(x, k, 3)
    dev = torch.device("cuda", 0)
    xt = torch.from_numpy(x).to(dev)
    c_a = torch.from_numpy(c0).to(dev)
    c_b = torch.empty_like(c_a)
    labels = torch.full((n,), -1, dtype=torch.int32, device=dev)
    counts = torch.zeros(k, dtype=torch.int64, device=dev)
    res = torch.zeros(C.sizeof(L.NoAllocResult), dtype=torch.uint8, device=dev)

    with km.KMeansState(path=L.PATH_DIRECT_F64, update_mode=L.UPDATE_FULL) as ref:
        ref.upload_points(x)
        ref.set_centroids(c0)
        c_in, c_out = c_a, c_b
        for call in range(2):                        # two calls in a row share the built-in array
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(dev)[0]
            _call(lib, L, xt, n, d, c_in, k, labels, c_out, counts, res)
            torch.cuda.synchronize()
            if call == 1:
                # the property the entry point exists for: past the warm-up call (context and
                # module set-up), a call leaves the device's free memory exactly where it was
                assert torch.cuda.mem_get_info(dev)[0] == free_before
            assert torch.cuda.current_device() == 0
            want = ref.step()
            got = L.NoAllocResult.from_buffer_copy(res.cpu().numpy().tobytes())
            assert got.status == L.OK and got.n == n
            assert got.changed == want["changed"]
            assert got.inertia == pytest.approx(want["inertia"], rel=REL)
            assert np.array_equal(labels.cpu().numpy(), ref.labels())
            assert np.array_equal(counts.cpu().numpy(), ref.counts())
            wc = ref.centroids()
            gc = c_out.cpu().numpy()
            scale = max(1.0, float(np.abs(wc).max()))
            assert np.all(np.abs(gc - wc) <= REL * scale)
            c_in, c_out = c_out, c_in                # next call starts from this call's centroids

    # and against the CPU oracle directly: labels of the second call are the oracle's
    lab1, _, _ = oracle.assign(x, c0)
    c1, _ = oracle.update(x, lab1, c0, sum_mode=oracle.SUM_EXACT)
    lab2, _, _ = oracle.assign(x, c1, lab1)
    assert np.array_equal(labels.cpu().numpy(), lab2)


def test_noalloc_rejects_shapes_beyond_the_builtin_array(km):
    import torch
    L = km._lib
    lib = L.load()
    dev = torch.device("cuda", 0)
    n, d, k = 64, 128, 1024                          # k*(2d+1)+4 = 263172 > 32768
    xt = torch.zeros((n, d), dtype=torch.float32, device=dev)
    c = torch.zeros((k, d), dtype=torch.float64, device=dev)
    labels = torch.full((n,), -1, dtype=torch.int32, device=dev)
    counts = torch.zeros(k, dtype=torch.int64, device=dev)
    res = torch.zeros(C.sizeof(L.NoAllocResult), dtype=torch.uint8, device=dev)
    st = lib.km_lloyd_step_noalloc(0, None, xt.data_ptr(), n, d, c.data_ptr(), k, labels.data_ptr(), c.data_ptr(),
                                   counts.data_ptr(), res.data_ptr())
    assert st == L.ERR_UNSUPPORTED
    msg = (lib.km_last_error(None) or b"").decode()
    assert "32768" in msg and "scratch" in msg
    st = lib.km_lloyd_step_noalloc(0, None, xt.data_ptr(), n, 0, c.data_ptr(), k, labels.data_ptr(), c.data_ptr(),
                                   counts.data_ptr(), res.data_ptr())
    assert st == L.ERR_INVALID_ARG


def test_noalloc_flags_nonfinite_centroids_and_keeps_the_current_device(km):
    import torch
    L = km._lib
    lib = L.load()
    dev = torch.device("cuda", 0)
    n, d, k = 256, 2, 3
    xt = torch.zeros((n, d), dtype=torch.float32, device=dev)
    c = torch.zeros((k, d), dtype=torch.float64, device=dev)
    c[1, 0] = float("nan")
    labels = torch.full((n,), -1, dtype=torch.int32, device=dev)
    counts = torch.zeros(k, dtype=torch.int64, device=dev)
    res = torch.zeros(C.sizeof(L.NoAllocResult), dtype=torch.uint8, device=dev)
    out = torch.empty_like(c)
    _call(lib, L, xt, n, d, c, k, labels, out, counts, res)
    torch.cuda.synchronize()
    got = L.NoAllocResult.from_buffer_copy(res.cpu().numpy().tobytes())
    assert got.status == L.ERR_NONFINITE
    if torch.cuda.device_count() > 1:                # a call for another device leaves the caller's device current
        with torch.cuda.device(1):
            x1, c1 = xt.to("cuda:1"), torch.zeros((k, d), dtype=torch.float64, device="cuda:1")
            l1, n1 = labels.to("cuda:1"), counts.to("cuda:1")
            r1, o1 = res.to("cuda:1"), torch.empty((k, d), dtype=torch.float64, device="cuda:1")
        st = lib.km_lloyd_step_noalloc(1, None, x1.data_ptr(), n, d, c1.data_ptr(), k, l1.data_ptr(), o1.data_ptr(),
                                       n1.data_ptr(), r1.data_ptr())
        assert st == L.OK and torch.cuda.current_device() == 0
        torch.cuda.synchronize(1)
