#!/usr/bin/env python
"""bench.py — Lloyd point-iterations/sec on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c5|c2|c3|c4] [--shapes c2,c3,c4|none] [--n-total N]

A "step" is one Lloyd iteration (assignment fused with the partial sums, then the centroid
update) over the whole job.  The default workload is the configuration BASELINE.json's metric
is quoted on across 1..8 GPUs: configs[4], N = 1e9, d = 2, k = 64 — STRONG scaling: the same
1e9 points are split over the `--gpus N` ranks (8 GB of points + 4 GB of labels fit one B200),
so labels, centroids and the iteration count must be the same bits at every N (the
`parity_check` object says so).  The other named GPU shapes (configs[1], [2], [3]) are measured
in the same invocation and reported under `shapes`.  One JSON line on stdout (rank 0); every
field is described in DESIGN.md §7.

`--impl reference` times the CPU restatement of the path (oracle/, all host threads) on the
same workload: the reference itself has no Lloyd code and no JS engine exists on this image
(SURVEY.md §0 F1-F3), so the oracle port *is* the reference arm here.
"""
import argparse
import hashlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "lloyd_point_iterations_per_sec"
UNIT = "point-iterations/s"

# n_total: points in the whole job; `split`: how the job meets the GPUs present
#   "strong"   the same n_total split over the ranks
#   "per_gpu"  every rank holds n_total points (configs[1] names one B200)
#   "c4"       configs[3] names 8 GPUs: n_total = 1e8 when 8 ranks are present, else the
#              1.25e7-point shard of it on every rank
WORKLOADS = {
    # configs[0] names a demo workload the reference does not have: the stand-in is
    # the reference's own card count x 1000 and its maximum k (app.mjs:127, 188, 204-216)
    "c1": dict(n_total=12_000, d=2, k=3, k_true=3, seed=1, init_seed=1, split="strong",
               name="configs[0] stand-in: N=12000 d=2 k=3 blobs (not present in the reference)"),
    "c2": dict(n_total=1_000_000, d=2, k=8, k_true=8, seed=2, init_seed=2, split="per_gpu",
               name="configs[1]: N=1e6 d=2 k=8 blobs per GPU"),
    "c3": dict(n_total=10_000_000, d=32, k=256, k_true=256, seed=3, init_seed=3, split="strong",
               name="configs[2]: N=1e7 d=32 k=256 blobs, split over the GPUs present"),
    "c4": dict(n_total=100_000_000, d=128, k=1024, k_true=1024, seed=4, init_seed=4, split="c4",
               name="configs[3]: N=1e8 d=128 k=1024 blobs over 8 GPUs (fewer GPUs: the 1.25e7-point shard per GPU)"),
    "c5": dict(n_total=1_000_000_000, d=2, k=64, k_true=64, seed=5, init_seed=5, split="strong",
               name="configs[4]: N=1e9 d=2 k=64 blobs, split over the GPUs present"),
}
L2_FLUSH_BYTES = 256 << 20  # > 126 MB L2
PARITY_POINTS = {"c1": 12_000, "c2": 1_000_000, "c3": 200_000, "c4": 50_000, "c5": 1_000_000}


def job_size(wl, world):
    """(n_total of the job as run, scaling word)."""
    if wl["split"] == "strong":
        return wl["n_total"], "strong"
    if wl["split"] == "per_gpu":
        return wl["n_total"] * world, "weak"
    if world >= 8:
        return wl["n_total"], "strong"
    return (wl["n_total"] // 8) * world, "weak"


def shard_range(n_total, rank, world):
    per = -(-n_total // world)
    return min(n_total, rank * per), min(n_total, (rank + 1) * per)


def base_config(wl, n_total, world, scaling):
    """The `config` object — the same keys and values on both arms."""
    return {"workload": wl["name"], "n_total": n_total, "d": wl["d"], "k": wl["k"], "k_true": wl["k_true"],
            "seed": wl["seed"], "init": f"seeded Forgy (seed {wl['init_seed']})", "storage": "float32",
            "scaling": scaling, "parallelism": f"point-sharded x{world}"}


def algorithmic_bytes_per_point(d):
    """fused assign+update, float32 storage: read the point once (4d), read
    the old label (4), write the new label (4)."""
    return 4 * d + 8


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def measured_peak_gbs():
    p = measured_peaks()
    if "hbm_gbs" in p:
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic_bytes(workload, n_local):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the assignment kernel: the
    per-point figure of the committed `ncu --set full` capture (profiles/ncu_traffic.json)
    times the points this launch processes, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            per_point = json.load(f).get(workload, {}).get("dram_bytes_per_point")
        return None if per_point is None else per_point * n_local
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons, sampled every 100 ms while the GPU is loaded."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t_load0, t_load1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.samples:
            if not (t_load0 <= t <= t_load1 + 0.1):
                continue
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples_under_load": len(sm),
                "reasons": sorted(reasons)}


def env_rank():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def host_threads():
    """Host threads this process may use — from the affinity mask, NOT from OMP_NUM_THREADS
    (torchrun exports OMP_NUM_THREADS=1, which must not shrink the CPU arm)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


# ------------------------------------------------------------------------------- CPU arms
def cpu_steps(wl, n_sample, steps, warmup, nthreads, budget_s=None):
    """Lloyd iterations of the oracle port on `nthreads` host threads; returns
    (pt-it/s, steps done, seconds).  Same seeded generator and Forgy init as the GPU arm."""
    import oracle
    x = oracle.gen_blobs(wl["seed"], n_sample, wl["d"], wl["k_true"])
    c = oracle.forgy_init(x, wl["k"], wl["init_seed"])
    labels = None

    def one():
        nonlocal labels, c
        labels, changed, _ = oracle.assign(x, c, labels, nthreads=nthreads)
        if changed:
            c, _ = oracle.update(x, labels, c)

    for _ in range(warmup):
        one()
    done, t0 = 0, time.perf_counter()
    while True:
        one()
        done += 1
        el = time.perf_counter() - t0
        if budget_s is None:
            if done >= steps:
                break
        elif el >= budget_s and done >= 3:
            break
    el = time.perf_counter() - t0
    return n_sample * done / el, done, el


def run_reference(args, wl):
    rank, _, world = env_rank()
    if rank != 0:
        return 0
    n_total, scaling = job_size(wl, args.gpus)
    if args.n_total:
        n_total = args.n_total
    nthreads = host_threads()
    # bound the sample so W+K steps end within minutes on the host cores
    per_point_flops = 3.0 * wl["d"] * wl["k"]
    n_sample = int(min(n_total, max(50_000, 4e10 / per_point_flops / max(1, args.steps + args.warmup))))
    val, done, el = cpu_steps(wl, n_sample, args.steps, args.warmup, nthreads)
    n1 = max(20_000, n_sample // max(1, nthreads))
    v1, done1, el1 = cpu_steps(wl, n1, 0, 1, 1, budget_s=3.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / done, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": base_config(wl, n_total, args.gpus, scaling),
        "details": {"reference_arm": "CPU restatement of the path (oracle/kmeans_oracle.c, OpenMP over points, thread count "
                                     "set explicitly from the affinity mask); the reference has no Lloyd code and no JS "
                                     "engine exists on this image",
                    "sample_points": n_sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": nthreads, "kind": "port",
                         "sample": f"{n_sample} points x {done} Lloyd iterations, same generator/init",
                         "single_thread": {"value": v1, "sample": f"{n1} points x {done1} Lloyd iterations in {el1:.1f} s, 1 thread"}},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------- GPU arm
class Job:
    """torch / torch.distributed plumbing shared by every workload of one invocation."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        import kmeans_b200 as km
        self.torch, self.dist, self.km, self.args = torch, dist, km, args
        self.rank, self.local_rank, self.world = env_rank()
        if self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}: launch with torchrun --nproc-per-node {args.gpus}")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        # a real (non-default) stream shared by torch's events/flush and the library's kernels:
        # the legacy default stream's handle is 0, which the C ABI reads as "make your own"
        self.tstream = torch.cuda.Stream(device=self.dev)
        torch.cuda.set_stream(self.tstream)
        self.stream = self.tstream.cuda_stream
        assert self.stream != 0
        self.flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def new_state(self, **kw):
        km, torch = self.km, self.torch
        st = km.KMeansState(device=self.local_rank, stream=self.stream, profile_kernels=False, **kw)
        if self.world > 1:
            uid = torch.zeros(128, dtype=torch.uint8, device=self.dev)
            if self.rank == 0:
                uid.copy_(torch.frombuffer(bytearray(km.KMeansState.comm_unique_id()), dtype=torch.uint8))
            self.dist.broadcast(uid, 0)
            st.comm_init(bytes(uid.cpu().numpy().tobytes()), self.rank, self.world)
        return st

    def l2_flush(self):
        if not self.args.no_flush:
            self.flush.zero_()

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def roofline_of(key, wl, n_local, kern_avg_ms, path_name):
    d, k = wl["d"], wl["k"]
    if path_name == "tensor":
        # GEMM-form path: algorithmic flops = 2*k*d per point-iteration; the timed
        # bracket covers the whole assignment (norms + tcgen05 coarse stage + exact stage)
        peaks = measured_peaks()
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        peak_src = ("measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if peaks else
                    "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained)")
        flops = 2.0 * k * d * n_local
        achieved = flops / (kern_avg_ms * 1e-3) / 1e12
        return {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "traffic": ncu_traffic_bytes(key, n_local),
                "kernel": "km_tensor_coarse_kernel (+ the norm, order and exact-stage kernels of the same assignment)",
                "kernel_ms": kern_avg_ms, "algorithmic_flops_per_launch": flops, "peak_source": peak_src,
                "hbm_frac_of_measured": algorithmic_bytes_per_point(d) * n_local / (kern_avg_ms * 1e-3) / 1e9 / measured_peak_gbs()[0]}
    alg_bytes = algorithmic_bytes_per_point(d) * n_local
    achieved = alg_bytes / (kern_avg_ms * 1e-3) / 1e9
    peak, peak_src = measured_peak_gbs()
    return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": ncu_traffic_bytes(key, n_local),
            "kernel": "km_assign_small_kernel" if path_name == "filter_f32" else "km_assign_direct_kernel",
            "kernel_ms": kern_avg_ms, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
            "frac_of_nominal_8TBs": achieved / 8000.0}


def measure(job, key, wl, steps, warmup, burn_s, keep_state=False):
    """Generate the workload on the device, time `steps` Lloyd iterations (W+1..W+K of a real run
    from the seeded init, L2 flushed between them), then the same iterations once more with each
    assignment kernel bracketed by its own events, the first pass, and the full-update mode."""
    torch, km, args = job.torch, job.km, job.args
    world, rank = job.world, job.rank
    n_total, scaling = job_size(wl, world)
    if args.n_total and key == args.workload:
        n_total = args.n_total
    lo, hi = shard_range(n_total, rank, world)
    n, d, k = hi - lo, wl["d"], wl["k"]
    st = job.new_state()
    # inputs resident in HBM before anything is timed: device-side seeded generator (A8)
    st.generate_blobs(n, d, wl["k_true"], wl["seed"], sigma=1.0, i0=lo)
    st.init_forgy(k, wl["init_seed"])
    c0 = st.centroids()

    t_load0 = time.time()
    t_end = time.perf_counter() + burn_s
    while True:
        # every rank must issue the same number of steps (each step is a collective): rank 0's
        # clock decides, the others follow
        go = time.perf_counter() < t_end
        if world > 1:
            flag = torch.tensor([1 if go else 0], dtype=torch.int32, device=job.dev)
            job.dist.broadcast(flag, 0)
            go = bool(flag.item())
        if not go:
            break
        for _ in range(10):
            job.l2_flush()
            st.step_async()
        torch.cuda.synchronize()

    def timed_pass(profile):
        st.set_profile(profile)
        st.reset_labels()
        st.set_centroids(c0)
        # first pass (labels all -1: every point is accumulated), timed on its own
        fa, fb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        job.l2_flush()
        fa.record()
        st.step_async()
        fb.record()
        for _ in range(max(0, warmup - 1)):
            job.l2_flush()
            st.step_async()
        st.stats_fetch()
        st.kernel_time(reset=True)
        l0 = st.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        job.barrier()
        for a, b in evs:
            job.l2_flush()        # L2 flush between timed iterations (outside the event pair)
            a.record()
            st.step_async()
            b.record()
        job.barrier()
        return [a.elapsed_time(b) for a, b in evs], st.launch_count() - l0, fa.elapsed_time(fb)

    # pass 1: the timed region proper (no extra events inside a step)
    step_ms, launches, first_ms = timed_pass(False)
    t_load1 = time.time()
    total_ms = job.max_over_ranks(sum(step_ms))
    first_ms = job.max_over_ranks(first_ms)
    last = st.stats_fetch()
    # pass 2: the same iterations again with each assignment kernel bracketed by its own events
    timed_pass(True)
    kern_ms, kern_n = st.kernel_time(reset=True)
    st.set_profile(False)
    kern_avg_ms = kern_ms / max(1, kern_n)
    path_name = {1: "direct_f64", 2: "filter_f32", 3: "tensor"}.get(last["path_used"])

    # full-update mode (every point re-accumulated every iteration; same bits out): a few steps
    full_ms = None
    if not args.no_full_mode:
        st.set_update_mode(km._lib.UPDATE_FULL)
        st.set_centroids(c0)
        for _ in range(warmup):
            job.l2_flush()
            st.step_async()
        fsteps = max(3, min(steps, 5))
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(fsteps)]
        job.barrier()
        for a, b in evs:
            job.l2_flush()
            a.record()
            st.step_async()
            b.record()
        job.barrier()
        full_ms = job.max_over_ranks(sum(a.elapsed_time(b) for a, b in evs)) / fsteps
        st.set_update_mode(km._lib.UPDATE_INCREMENTAL)

    # parity: labels of a prefix of rank 0's shard under the centroids in use == the oracle's
    m = min(n, PARITY_POINTS.get(key, 100_000))
    st.reset_labels()
    st.set_centroids(c0)
    for _ in range(3):
        st.step_async()
    st.stats_fetch()
    c_used = st.centroids()
    st.assign()
    lab = st.labels()
    parity = None
    if rank == 0:
        import oracle
        xs = st.get_points(0, m)
        olab, _, _ = oracle.assign(xs, c_used, nthreads=host_threads())
        parity = {"points": int(m), "mismatches": int(np.count_nonzero(lab[:m] != olab)), "after_iterations": 3,
                  "generator_rows_equal": bool(np.array_equal(xs[:1000], oracle.gen_blobs(wl["seed"], min(m, 1000), d, wl["k_true"]))),
                  "centroid_sha": hashlib.sha256(np.ascontiguousarray(c_used).tobytes()).hexdigest()[:16]}
    del lab

    # the whole-job call on the resident points: km_run(R) from the same start (device time of the call)
    st.set_centroids(c0)
    st.run(2)              # warm-up: the first launch of a kernel pays its module load
    st.set_centroids(c0)
    l0 = st.launch_count()
    rs = st.run(20)
    run_launches = st.launch_count() - l0
    run_ms = job.max_over_ranks(rs["device_ms"])
    run_info = {"iterations": rs["iterations"], "ms_per_iteration": run_ms / max(1, rs["iterations"]), "launches": run_launches,
                "how": "one cooperative launch, points resident in shared memory" if run_launches <= 2 and rs["iterations"] > 2
                       else "one launch sequence per iteration"}

    res = {
        "workload": wl["name"], "n_total": n_total, "n_per_gpu": n, "d": d, "k": k, "scaling": scaling,
        "value": n_total * steps / (total_ms * 1e-3), "ms_per_step": total_ms / steps, "steps": steps,
        "first_pass_ms": first_ms, "full_mode_ms_per_step": full_ms, "path": path_name,
        "roofline": roofline_of(key, wl, n, kern_avg_ms, path_name), "gpu_launches": launches,
        "changed_at_last_step": last["changed"], "parity_check": parity,
        "step_ms_min_med_max": [min(step_ms), statistics.median(step_ms), max(step_ms)],
        "step_ms": [round(v, 4) for v in step_ms],
        "km_run": run_info,
        "collective": "none (one rank)" if world == 1 else
                      ("one-shot kernel over peer memory" if st.comm_info()[2] == 2 else "ncclAllReduce(int64)"),
    }
    extra = dict(st=st, c0=c0, n=n, lo=lo, t_load=(t_load0, t_load1), total_ms=total_ms, launches=launches)
    if not keep_state:
        st.close()
        extra["st"] = None
    return res, extra


def run_ours(args, wl):
    job = Job(args)
    torch, km = job.torch, job.km
    rank, world = job.rank, job.world
    sampler = ClockSampler(job.local_rank) if rank == 0 else None

    main, ex = measure(job, args.workload, wl, args.steps, args.warmup, args.burn, keep_state=True)
    clocks = sampler.stop(*ex["t_load"]) if sampler else None
    st, c0, n = ex["st"], ex["c0"], ex["n"]
    d, k, n_total = wl["d"], wl["k"], main["n_total"]

    # ---- e2e: the public call sequence with HOST buffers, copies inside the timed region
    xbuf, xaddr = km.host_alloc(max(16, n * d * 4))
    lbuf, laddr = km.host_alloc(max(16, n * 4))
    x_host = xbuf[:n * d * 4].view(np.float32).reshape(n, d)
    labels_host = lbuf[:n * 4].view(np.int32)
    st.get_points(out=x_host)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    c_host = np.empty((k, d), dtype=np.float64)

    byte_labels = k <= 255                     # km_get_labels_u8: a quarter of the label bytes
    l8buf, _ = km.host_alloc(max(16, n))
    labels8_host = l8buf[:n].view(np.uint8)

    def e2e_step(narrow):
        st.upload_points_raw(xaddr, n, d)      # H2D: the step's points, from pinned memory
        st.set_centroids(c0)                   # H2D: the step's centroids
        s = st.step()                          # one Lloyd iteration
        if narrow:
            st.labels(out=labels8_host)        # D2H: labels as bytes
        else:
            st.labels(out=labels_host)         # D2H: labels as int32
        st.centroids(out=c_host)               # D2H: updated centroids
        return s

    def e2e_time(narrow):
        for _ in range(2):
            e2e_step(narrow)
        job.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step(narrow)
        job.barrier()
        return n_total * e2e_steps / job.max_over_ranks(time.perf_counter() - t0)

    e2e_i32 = e2e_time(False)
    e2e_value = e2e_time(True) if byte_labels else e2e_i32
    u8_ok = bool(np.array_equal(labels8_host, labels_host.astype(np.uint8))) if byte_labels else None
    # the whole-job call a user makes (upload once, iterate, download once) — and the parity record:
    # labels / centroids / iteration count of this run must be the same bits at every GPU count
    job.barrier()
    t0 = time.perf_counter()
    st.upload_points_raw(xaddr, n, d)
    st.set_centroids(c0)
    rs = st.run(args.run_iters)
    if byte_labels:
        st.labels(out=labels8_host)
    else:
        st.labels(out=labels_host)
    st.centroids(out=c_host)
    job.barrier()
    run_s = job.max_over_ranks(time.perf_counter() - t0)
    st.labels(out=labels_host)                 # (untimed) the int32 labels the parity record below is taken from
    if byte_labels and u8_ok:
        u8_ok = bool(np.array_equal(labels8_host, labels_host.astype(np.uint8)))
    e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": (n * d * 4 + k * d * 8) * world,
           "d2h_bytes_per_step": (n * (1 if byte_labels else 4) + k * d * 8 + 64) * world, "steps": e2e_steps,
           "call": "km_upload_points(pinned)+km_set_centroids+km_step+" + ("km_get_labels_u8" if byte_labels else "km_get_labels") +
                   "+km_get_centroids per step",
           "value_with_int32_labels": e2e_i32, "byte_labels_equal_int32_labels": u8_ok,
           "run_value": n_total * rs["iterations"] / run_s, "run_iterations": rs["iterations"], "run_seconds": run_s,
           "run_call": "km_upload_points+km_set_centroids+km_run(max_iter)+" + ("km_get_labels_u8" if byte_labels else "km_get_labels") +
                       "+km_get_centroids, once"}
    parity = None
    if rank == 0:
        import oracle
        m = min(n, PARITY_POINTS.get(args.workload, 100_000))
        olab, _, _ = oracle.assign(x_host[:m], c_host, nthreads=host_threads())
        parity = {"points": int(m), "mismatches": int(np.count_nonzero(labels_host[:m] != olab)),
                  "iterations": int(rs["iterations"]), "inertia": rs["inertia"],
                  "centroid_sha": hashlib.sha256(c_host.tobytes()).hexdigest()[:16],
                  "labels_sha_prefix": hashlib.sha256(labels_host[:m].tobytes()).hexdigest()[:16],
                  "what": f"km_run({args.run_iters}) from the seeded Forgy init: labels of the first {m} rows against the oracle's "
                          "argmin under the final centroids; centroid_sha / iterations must not depend on the GPU count"}
    km.host_free(xaddr)
    km.host_free(laddr)
    st.close()

    cpu = None
    if rank == 0 and world == 1 and args.cpu_seconds > 0:
        per_point_flops = 3.0 * d * k
        n_sample = int(min(n, max(50_000, 2e9 / per_point_flops)))
        nthreads = host_threads()
        v, done, el = cpu_steps(wl, n_sample, 0, 1, nthreads, budget_s=args.cpu_seconds)
        cpu = {"value": v, "unit": UNIT, "cores": nthreads, "kind": "port",
               "sample": f"{n_sample} points x {done} Lloyd iterations in {el:.1f} s, oracle/kmeans_oracle.c with OpenMP"}
        # one thread: what a single-threaded JS host (the reference's world) would have to work with
        n1 = max(50_000, n_sample // max(1, nthreads))
        v1, done1, el1 = cpu_steps(wl, n1, 0, 1, 1, budget_s=min(3.0, args.cpu_seconds))
        cpu["single_thread"] = {"value": v1, "sample": f"{n1} points x {done1} Lloyd iterations in {el1:.1f} s, 1 thread"}

    # ---- the other named GPU shapes, same invocation
    shapes = {}
    for key in [s for s in args.shapes.split(",") if s and s != "none" and s != args.workload]:
        try:
            r, _ = measure(job, key, WORKLOADS[key], args.shape_steps, 3, 0.3)
            shapes[key] = r
        except Exception as e:   # a shape that fails must not take the main line with it
            shapes[key] = {"error": f"{type(e).__name__}: {e}"}
            if world > 1:
                raise

    if rank == 0:
        config = base_config(wl, n_total, world, main["scaling"])
        line = {
            "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": main["scaling"], "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config,
            "details": {"n_per_gpu": n, "arithmetic": "binary32 filter + binary64 re-score, int64 order-free sums, inertia from the sums",
                        "path": main["path"], "update_mode": "incremental", "collective": main["collective"],
                        "l2": "NOT flushed (diagnostic run, invalid as a bench value)" if args.no_flush else "flushed between timed steps (256 MiB write)",
                        "timed_iterations": f"Lloyd iterations {args.warmup + 1}..{args.warmup + args.steps} from the seeded init",
                        "changed_at_last_step": main["changed_at_last_step"]},
            "roofline": main["roofline"], "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": main["gpu_launches"], "clocks": clocks,
            "first_pass_ms": main["first_pass_ms"], "full_mode_ms_per_step": main["full_mode_ms_per_step"],
            "step_ms_min_med_max": main["step_ms_min_med_max"], "step_ms": main["step_ms"],
            "parity_check": parity, "parity_check_after_3_iterations": main["parity_check"],
            "shapes": shapes,
        }
        emit(line)
    job.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c5")
    ap.add_argument("--shapes", default="c2,c3,c4", help="other named shapes measured in the same invocation ('none' to skip)")
    ap.add_argument("--shape-steps", type=int, default=10)
    ap.add_argument("--n-total", type=int, default=0, help="override the main workload's N (development runs)")
    ap.add_argument("--burn", type=float, default=1.5, help="seconds of untimed load before the timed steps (clock sampling)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--run-iters", type=int, default=50)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-full-mode", action="store_true", help="skip the full-update-mode timing")
    ap.add_argument("--no-flush", action="store_true",
                    help="diagnostic only: skip the L2 flush between timed steps (such a line is not a valid bench value)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    wl = WORKLOADS[args.workload]
    # stdout carries exactly ONE line, the JSON: anything a library writes to file descriptor 1
    # meanwhile (NCCL's version banner, for one) is sent to stderr instead
    sys.stdout.flush()
    global _JSON_FD
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    return run_reference(args, wl) if args.impl == "reference" else run_ours(args, wl)


_JSON_FD = None


def emit(line):
    """Write the bench line to the process's original stdout."""
    text = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(text.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, text)


if __name__ == "__main__":
    sys.exit(main())
