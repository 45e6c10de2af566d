#!/usr/bin/env python
"""bench.py — Lloyd point-iterations/sec on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4|c5]

A "step" is one Lloyd iteration (assignment fused with the partial sums, then the
centroid update) over this rank's shard of seeded synthetic blobs.  Default workload is the
one BASELINE.json's metric ("... at 1/2/4/8 B200 ...") is quoted on across 1..8 GPUs:
configs[4] (N=1e9, d=2, k=64 over 8 B200) in its per-GPU form — every rank holds one shard of
1.25e8 points (weak scaling; --gpus 8 is exactly configs[4], and one shard fits one GPU) and a
step carries one NCCL all-reduce of the packed order-free sums.  `--workload c2` runs
configs[1] (N=1e6, d=2, k=8 per GPU), `c3`/`c4` the tensor-path shapes.  One JSON line on
stdout (rank 0); see DESIGN.md §7 for every field.

`--impl reference` times the CPU restatement of the path (oracle/, all host threads) on the
same workload: the reference itself has no Lloyd code and no JS engine exists on this image
(SURVEY.md §0 F1-F3), so the oracle port *is* the reference arm here.
"""
import argparse
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

# per-GPU shard shapes (weak scaling).  c5/c4 are the per-GPU shards of configs[4]/[3].
WORKLOADS = {
    # configs[0] names a demo workload the reference does not have: the stand-in is
    # the reference's own card count x 1000 and its maximum k (app.mjs:127, 188, 204-216)
    "c1": dict(n=12_000, d=2, k=3, k_true=3, seed=1, init_seed=1,
               name="configs[0] stand-in: N=12000 d=2 k=3 blobs (not present in the reference)"),
    "c2": dict(n=1_000_000, d=2, k=8, k_true=8, seed=2, init_seed=2,
               name="configs[1]: N=1e6 d=2 k=8 blobs per GPU"),
    "c3": dict(n=10_000_000, d=32, k=256, k_true=256, seed=3, init_seed=3,
               name="configs[2]: N=1e7 d=32 k=256 blobs per GPU"),
    "c4": dict(n=12_500_000, d=128, k=1024, k_true=1024, seed=4, init_seed=4,
               name="configs[3] shard: N=1.25e7 d=128 k=1024 blobs per GPU (1e8 on 8)"),
    "c5": dict(n=125_000_000, d=2, k=64, k_true=64, seed=5, init_seed=5,
               name="configs[4] shard: N=1.25e8 d=2 k=64 blobs per GPU (1e9 on 8)"),
}
L2_FLUSH_BYTES = 256 << 20  # > 126 MB L2


def algorithmic_bytes_per_point(d):
    """fused assign+update, float32 storage: read the point once (4d), read
    the old label (4), write the new label (4)."""
    return 4 * d + 8


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic_bytes(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the assignment kernel, from
    the committed `ncu --set full` capture (profiles/ncu_traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(workload, {}).get("dram_bytes_per_launch")
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


# ------------------------------------------------------------------------------- CPU arms
def cpu_steps(wl, n_sample, steps, warmup, budget_s=None, single_thread=False):
    """Lloyd iterations of the oracle port on all host threads (or one: the stand-in for a
    single-threaded JS host); returns (pt-it/s, steps done, threads).  Same seeded generator and
    Forgy init as the GPU arm."""
    import oracle
    x = oracle.gen_blobs(wl["seed"], n_sample, wl["d"], wl["k_true"])
    c = oracle.forgy_init(x, wl["k"], wl["init_seed"])
    labels = None
    nthreads = 1 if single_thread else oracle.num_threads()

    def one():
        nonlocal labels, c
        labels, changed, _ = oracle.assign(x, c, labels, nthreads=1 if single_thread else 0)
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
    return n_sample * done / el, done, nthreads, el


def run_reference(args, wl):
    rank, _, world = env_rank()
    if rank != 0:
        return 0
    # bound the sample so W+K steps end within minutes on the host cores
    per_point_flops = 3.0 * wl["d"] * wl["k"]
    n_sample = int(min(wl["n"], max(50_000, 4e10 / per_point_flops / max(1, args.steps + args.warmup))))
    val, done, nthreads, el = cpu_steps(wl, n_sample, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["name"], "n_per_gpu": wl["n"], "d": wl["d"], "k": wl["k"],
                   "reference_arm": "CPU restatement of the path (oracle/kmeans_oracle.c, OpenMP over points); "
                                    "the reference has no Lloyd code and no JS engine exists on this image"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": nthreads, "kind": "port",
                         "sample": f"{n_sample} points x {done} Lloyd iterations, same generator/init"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------- GPU arm
def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    import kmeans_b200 as km

    rank, local_rank, world = env_rank()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n, d, k = wl["n"], wl["d"], wl["k"]
    n_total = n * world
    # a real (non-default) stream shared by torch's events/flush and the library's kernels:
    # the legacy default stream's handle is 0, which the C ABI reads as "make your own"
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    st = km.KMeansState(device=local_rank, stream=stream, profile_kernels=False)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(km.KMeansState.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        st.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
    # inputs resident in HBM before anything is timed: device-side seeded generator (A8)
    st.generate_blobs(n, d, wl["k_true"], wl["seed"], sigma=1.0, i0=rank * n)
    st.init_forgy(k, wl["init_seed"])
    c0 = st.centroids()
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)

    sampler = ClockSampler(local_rank) if rank == 0 else None
    # keep the GPU loaded long enough for nvidia-smi to see the clocks this workload runs at
    t_load0 = time.time()
    t_end = time.perf_counter() + args.burn
    while True:
        # every rank must issue the same number of steps (each step is a collective): rank 0's
        # clock decides, the others follow
        go = time.perf_counter() < t_end
        if world > 1:
            flag = torch.tensor([1 if go else 0], dtype=torch.int32, device=dev)
            dist.broadcast(flag, 0)
            go = bool(flag.item())
        if not go:
            break
        for _ in range(20):
            if not args.no_flush:
                flush.zero_()
            st.step_async()
        torch.cuda.synchronize()
    def timed_pass(profile):
        """Lloyd iterations W+1..W+K from the seeded init; per-step CUDA events, L2 flushed between steps."""
        st.set_profile(profile)
        st.reset_labels()
        st.set_centroids(c0)
        for _ in range(args.warmup):
            if not args.no_flush:
                flush.zero_()
            st.step_async()
        st.stats_fetch()
        st.kernel_time(reset=True)
        l0 = st.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        barrier()
        for a, b in evs:
            if not args.no_flush:
                flush.zero_()        # L2 flush between timed iterations (outside the event pair)
            a.record()
            st.step_async()
            b.record()
        barrier()
        return [a.elapsed_time(b) for a, b in evs], st.launch_count() - l0

    # pass 1: the timed region proper (no extra events inside a step)
    step_ms, launches = timed_pass(False)
    t_load1 = time.time()
    total_ms = max_over_ranks(sum(step_ms))
    last = st.stats_fetch()
    # pass 2: the same iterations again with each assignment kernel bracketed by its own events
    _, _ = timed_pass(True)
    kern_ms, kern_n = st.kernel_time(reset=True)
    st.set_profile(False)
    clocks = sampler.stop(t_load0, t_load1) if sampler else None

    value = n_total * args.steps / (total_ms * 1e-3)
    kern_avg_ms = kern_ms / max(1, kern_n)
    path_name = {1: "direct_f64", 2: "filter_f32", 3: "tensor"}.get(last["path_used"])
    if path_name == "tensor":
        # GEMM-form path: algorithmic flops = 2*k*d per point-iteration; the
        # timed bracket covers the whole assignment (norms + tcgen05 coarse stage + exact stage)
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        peak_src = ("measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if peaks else
                    "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained)")
        flops = 2.0 * k * d * n
        achieved = flops / (kern_avg_ms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                    "traffic": ncu_traffic_bytes(args.workload),
                    "kernel": "km_tensor_coarse_bf16_kernel (+ km_tensor_cnorm_kernel, km_tensor_stage2_kernel)",
                    "kernel_ms": kern_avg_ms, "algorithmic_flops_per_launch": flops, "peak_source": peak_src,
                    "hbm_frac_of_measured": algorithmic_bytes_per_point(d) * n / (kern_avg_ms * 1e-3) / 1e9 / measured_peak_gbs()[0]}
    else:
        alg_bytes = algorithmic_bytes_per_point(d) * n
        achieved = alg_bytes / (kern_avg_ms * 1e-3) / 1e9
        peak, peak_src = measured_peak_gbs()
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": ncu_traffic_bytes(args.workload),
                    "kernel": "km_assign_small_kernel" if path_name == "filter_f32" else "km_assign_direct_kernel",
                    "kernel_ms": kern_avg_ms, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                    "frac_of_nominal_8TBs": achieved / 8000.0}

    # ---- e2e: the public call sequence with HOST buffers, copies inside the timed region
    xbuf, xaddr = km.host_alloc(n * d * 4)
    lbuf, laddr = km.host_alloc(n * 4)
    x_host = xbuf.view(np.float32).reshape(n, d)
    labels_host = lbuf.view(np.int32)
    x_host[:] = st.get_points()
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    c_host = np.empty((k, d), dtype=np.float64)

    def e2e_step():
        st.upload_points_raw(xaddr, n, d)      # H2D: the step's points, from pinned memory
        st.set_centroids(c0)                   # H2D: the step's centroids
        s = st.step()                          # one Lloyd iteration
        st.labels(out=labels_host)             # D2H: labels
        st.centroids(out=c_host)               # D2H: updated centroids
        return s

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = n_total * e2e_steps / e2e_s
    # context: the whole-job call a user makes (upload once, iterate, download once)
    barrier()
    t0 = time.perf_counter()
    st.upload_points_raw(xaddr, n, d)
    st.set_centroids(c0)
    rs = st.run(args.run_iters)
    st.labels(out=labels_host)
    st.centroids(out=c_host)
    barrier()
    run_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * d * 4 + k * d * 8,
           "d2h_bytes_per_step": n * 4 + k * d * 8 + 64, "steps": e2e_steps,
           "call": "km_upload_points(pinned)+km_set_centroids+km_step+km_get_labels+km_get_centroids per step",
           "run_value": n_total * rs["iterations"] / run_s, "run_iterations": rs["iterations"],
           "run_call": "km_upload_points+km_set_centroids+km_run(max_iter)+km_get_labels+km_get_centroids, once"}
    km.host_free(xaddr)
    km.host_free(laddr)

    cpu = None
    if rank == 0 and world == 1 and args.cpu_seconds > 0:
        per_point_flops = 3.0 * d * k
        n_sample = int(min(n, max(50_000, 2e9 / per_point_flops)))
        v, done, nthreads, el = cpu_steps(wl, n_sample, 0, 1, budget_s=args.cpu_seconds)
        cpu = {"value": v, "unit": UNIT, "cores": nthreads, "kind": "port",
               "sample": f"{n_sample} points x {done} Lloyd iterations in {el:.1f} s, oracle/kmeans_oracle.c with OpenMP"}
        # one thread: what a single-threaded JS host (the reference's world) would have to work with
        n1 = max(50_000, n_sample // max(1, nthreads))
        v1, done1, _, el1 = cpu_steps(wl, n1, 0, 1, budget_s=min(3.0, args.cpu_seconds), single_thread=True)
        cpu["single_thread"] = {"value": v1, "sample": f"{n1} points x {done1} Lloyd iterations in {el1:.1f} s, 1 thread"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl["name"], "n_per_gpu": n, "n_total": n_total, "d": d, "k": k, "k_true": wl["k_true"],
                       "seed": wl["seed"], "init": f"seeded Forgy (seed {wl['init_seed']})", "storage": "float32",
                       "arithmetic": "binary32 filter + binary64 re-score/inertia, int64 order-free sums",
                       "path": path_name,
                       "update_mode": "incremental", "l2": "NOT flushed (diagnostic run, invalid as a bench value)" if args.no_flush else "flushed between timed steps (256 MiB write)",
                       "timed_iterations": f"Lloyd iterations {args.warmup + 1}..{args.warmup + args.steps} from the seeded init",
                       "changed_at_last_step": last["changed"], "parallelism": f"point-sharded x{world}"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "step_ms_min_med_max": [min(step_ms), statistics.median(step_ms), max(step_ms)],
        }
        emit(line)
    st.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c5")
    ap.add_argument("--burn", type=float, default=1.5, help="seconds of untimed load before the timed steps (clock sampling)")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--run-iters", type=int, default=50)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
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
