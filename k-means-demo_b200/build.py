"""Build libkmeans.so in-tree for sm_100a.

    python k-means-demo_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU; the .so travels to the GPU box with the snapshot.
cudart is linked statically and NCCL / the driver are only ever dlopen'ed, so the library
loads (and exports its symbols) on a machine with no GPU and no driver.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
OUT = os.path.join(HERE, "libkmeans.so")

NVCC = os.environ.get("NVCC", "nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
          "-Xcompiler", "-fvisibility=hidden", "-Xcompiler", "-Wall"]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def headers_mtime():
    m = os.path.getmtime(os.path.join(HERE, "..", "include", "kmeans.h"))
    for f in os.listdir(CSRC):
        if f.endswith((".h", ".cuh")):
            m = max(m, os.path.getmtime(os.path.join(CSRC, f)))
    return m


def _compile(src, verbose):
    obj = os.path.join(OBJ, src[:-3] + ".o")
    cmd = [NVCC, *ARCH, *CFLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj, r.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hm = headers_mtime()
    todo, objs = [], []
    for s in sources():
        obj = os.path.join(OBJ, s[:-3] + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(hm, os.path.getmtime(os.path.join(CSRC, s))):
            todo.append(s)
    if todo:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            for obj, log in ex.map(lambda s: _compile(s, verbose), todo):
                if verbose and log:
                    print(log)
    if todo or not os.path.exists(OUT):
        cmd = [NVCC, *ARCH, "-shared", "-o", OUT, *objs, "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
