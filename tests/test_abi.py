"""CPU: the C-ABI library loads without a GPU, exports every symbol include/kmeans.h
declares, the ctypes mirror agrees with the header, and compute fails loudly (no CPU
fallback) when there is no device."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "kmeans.h")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"KM_API[^;(]*?\b(km_\w+)\s*\(", src)))


def test_header_declares_the_path():
    names = declared_symbols()
    for must in ("km_create", "km_destroy", "km_upload_points", "km_set_centroids", "km_init_forgy", "km_assign",
                 "km_update", "km_step", "km_run", "km_get_labels", "km_get_centroids", "km_get_counts",
                 "km_comm_init", "km_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol(km):
    lib = km._lib.load()
    out = subprocess.run(["nm", "-D", "--defined-only", km._lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (km_\w+)", out))
    declared = declared_symbols()
    assert declared, "no KM_API prototypes found"
    assert set(declared) <= exported, sorted(set(declared) - exported)
    for name in declared:
        assert hasattr(lib, name)
    # nothing but the ABI leaks out
    assert exported == set(declared), sorted(exported - set(declared))


def test_ctypes_table_matches_header(km):
    assert sorted(km._lib.PROTOTYPES) == declared_symbols()
    assert km._lib.load().km_abi_version() == km._lib.ABI_VERSION
    hdr = open(HEADER).read()
    assert f"#define KM_ABI_VERSION {km._lib.ABI_VERSION}" in hdr
    assert C.sizeof(km._lib.Config) == 40 and C.sizeof(km._lib.Stats) == 48


def test_struct_layout_agrees_with_compiler(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "kmeans.h"\n'
                   'int main(){printf("%zu %zu %zu %zu %zu\\n", sizeof(km_config), sizeof(km_stats),'
                   ' offsetof(km_config, stream), offsetof(km_stats, inertia), offsetof(km_stats, device_ms));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    assert got == ["40", "48", "24", "16", "40"]


def test_no_cpu_fallback_without_device(km):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(km.KMeansError) as ei:
        km.KMeansState()
    assert ei.value.status == km._lib.ERR_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_product_never_imports_the_oracle():
    """The product path must not import, include, link or dlopen anything under oracle/
    (comments may mention the oracle; code may not reach it)."""
    pkg = os.path.join(ROOT, "k-means-demo_b200")
    bad = re.compile(r"(import\s+oracle|from\s+oracle|oracle/|oracle\\|libkmeans_oracle|kmeans_oracle|numpy_ref)")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(base, f)).read()
                assert not bad.search(text), f
    for f in ("kmeans_b200.py",):
        assert not bad.search(open(os.path.join(ROOT, f)).read()), f


def test_shard_ranges(km):
    for n, w in ((10, 4), (0, 3), (7, 8), (10 ** 9, 8), (5, 1)):
        spans = [km.shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert all(hi - lo <= -(-n // w) for lo, hi in spans)
    with pytest.raises(ValueError):
        km.shard_range(10, 4, 4)
