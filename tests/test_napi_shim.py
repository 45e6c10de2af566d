"""Row f4: the Node-API addon source is well-formed C and binds exactly the C ABI.

No Node.js exists on the image, so the addon is compiled against napi/node_api_min.h (self-declared
prototypes) and inspected, never executed: "built, unexecuted"."""
import os
import re
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _nm(path, flag):
    out = subprocess.run(["nm", "-D", flag, path], capture_output=True, text=True, check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


def test_addon_compiles_and_binds_only_declared_abi():
    with tempfile.TemporaryDirectory() as tmp:
        node = os.path.join(tmp, "kmeans_napi.node")
        subprocess.run(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-DKM_NAPI_SELF_DECLARED",
                        "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "napi"), "-fPIC", "-shared",
                        "-o", node, os.path.join(ROOT, "napi", "kmeans_napi.c")], check=True)
        assert "napi_register_module_v1" in _nm(node, "--defined-only")
        wanted = {s for s in _nm(node, "--undefined-only") if s.startswith("km_")}
    header = open(os.path.join(ROOT, "include", "kmeans.h")).read()
    declared = set(re.findall(r"\b(km_[a-z0-9_]+)\s*\(", header))
    assert wanted and wanted <= declared, wanted - declared
    # the four entry points of the façade and what they need
    assert {"km_create", "km_upload_points", "km_set_centroids", "km_init_forgy", "km_assign", "km_update", "km_run",
            "km_get_labels", "km_get_centroids", "km_destroy", "km_last_error"} <= wanted


def test_facade_names_match_python_mirror():
    mjs = open(os.path.join(ROOT, "napi", "kmeans.mjs")).read()
    exported = set(re.findall(r"export (?:function|const) (\w+)", mjs))
    assert exported == {"init", "assign", "update", "step", "run"}
    api = open(os.path.join(ROOT, "k-means-demo_b200", "api.py")).read()
    for name in exported:
        assert re.search(r"^def %s\(" % name, api, re.M), name
