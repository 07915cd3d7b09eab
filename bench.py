#!/usr/bin/env python
"""Benchmark of the B200 K-Means Lloyd path (BASELINE.json metric: Lloyd point-iterations/s at
N=100M, D=128, K=1024).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c4]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one Lloyd iteration over the whole (sharded) batch: assignment + per-cluster
accumulation on every rank, one NCCL all-reduce of the K*D+K+2 partial doubles (N > 1), centroid
update.  Points are generated on the device and stay resident (the reference keeps them in
`tf.Variable`s, scripts/distribuitedClustering.py:217); N total rows are split over the ranks by
the reference's array_split rule (:184), so the total work is fixed ("strong" scaling).  The data set is
defined by global row blocks, so every GPU count clusters the SAME rows from the same initial centres
(`config.checks.centroid_checksum` / `inertia` agree across --gpus 1, 2, 4, 8).

After the timed region the final iteration is re-checked at full size against a float64 mirror of the oracle
on the GPU (`config.checks.parity`), see oracle/gpu_mirror.py.

One JSON line is printed by rank 0; see DESIGN.md section "Measurement" for every key.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "lloyd_point_iterations_per_sec"
UNIT = "point-iterations/s"
CONFIGS = {
    # name: (N total, D, K, kind)
    "c1": (1_000_000, 2, 8, "blobs"),
    "c2": (3840 * 2160, 3, 16, "image"),
    "c3": (10_000_000, 32, 256, "blobs"),
    "c4": (100_000_000, 128, 1024, "blobs"),
    "c5": (1_000_000_000, 64, 4096, "blobs"),
}
SEED = 123128          # the seed the reference logs (scripts/new_experiment.py:56)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "tflops": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])),
                "tflops_burst": float(p["bf16_tflops"]), "source": "measured"}
    # /opt/skills/guides/B200_PROFILING.md fallback figures
    return {"hbm_gbs": 6650.0, "tflops": 1400.0, "tflops_burst": 1590.0, "source": "fallback"}


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.1):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        self._stop = threading.Event()
        self._period = period_s
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:        # no NVML here: report that instead of inventing clocks
            self.error = repr(e)
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self._nv
        names = {
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "hw_power_brake_slowdown": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(self._period)

    def start(self):
        if self.ok:
            self._t.start()

    def stop(self):
        self._stop.set()
        if self.ok:
            self._t.join(timeout=2)
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ----------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle port of the reference's CPU path, on the host cores
# ----------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def blas_threads():
    """Threads the loaded BLAS / OpenMP pools will really use (threadpoolctl), as a dict per library."""
    try:
        from threadpoolctl import threadpool_info
        return {f"{i.get('user_api')}:{i.get('internal_api')}": i.get("num_threads") for i in threadpool_info()}
    except Exception as e:
        return {"error": repr(e)}


def cpu_sample(cfg_name, sample_rows):
    pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
    N, D, K, kind = CONFIGS[cfg_name]
    if kind == "image":
        side = int(np.sqrt(sample_rows / (16 / 9)))
        X = pkg.datagen.image_numpy(side, side * 16 // 9, SEED)
        C = pkg.datagen.kmeanspp_init(X, K, SEED).astype(np.float64)
    else:
        X, _ = pkg.datagen.blobs_numpy(sample_rows, D, K, SEED)
        C = X[:K].astype(np.float64)
    return X, C


def time_cpu_arms(cfg_name, steps, warmup):
    """The CPU implementations of the same Lloyd iteration, each on a bounded sample of the workload with every
    host thread the affinity mask allows (BLAS / OpenMP pools pinned to that count with threadpoolctl):
      port     the float64 NumPy restatement of reference :221-263 (oracle/), GEMM form where K*D > 4096
      sklearn  sklearn.cluster.KMeans(init=C0, n_init=1, algorithm='lloyd', tol=0): the OpenMP multi-core
               implementation of the same math (the reference's authors compared with a CPU library too,
               notebooks/Testing Images.ipynb:347)
      literal  (config c1 only) the reference-shaped [n, K, D] float64 broadcast of :221-230
    Returns {name: {"value", "rows", "seconds_per_iteration", ...}}."""
    import oracle
    from threadpoolctl import threadpool_limits

    N, D, K, kind = CONFIGS[cfg_name]
    cores = host_cores()
    out = {}
    with threadpool_limits(limits=cores):
        # --- port
        rows = int(min(N, max(4096, 8e10 / (2.0 * K * D))))
        X, C = cpu_sample(cfg_name, rows)
        form = "gemm" if K * D > 4096 else "diff"
        state = {"C": C}

        def step():
            it = oracle.lloyd_iteration(X, state["C"], n_shards=1, form=form)
            state["C"] = it["centroids"]

        for _ in range(max(1, warmup)):
            step()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        dt = (time.perf_counter() - t0) / steps
        out["port"] = {"value": X.shape[0] / dt, "rows": int(X.shape[0]), "seconds_per_iteration": dt,
                       "what": f"float64 {form}-form NumPy restatement (oracle/)"}
        # --- sklearn
        try:
            from sklearn.cluster import KMeans
            rows_s = int(min(N, max(4096, 4e11 / (2.0 * K * D))))
            Xs, Cs = cpu_sample(cfg_name, rows_s)
            iters = max(2, steps)
            km = KMeans(n_clusters=K, init=Cs.astype(np.float32), n_init=1, algorithm="lloyd", tol=0.0, max_iter=iters)
            t0 = time.perf_counter()
            km.fit(Xs)
            dts = (time.perf_counter() - t0) / max(1, km.n_iter_)
            out["sklearn"] = {"value": Xs.shape[0] / dts, "rows": int(Xs.shape[0]), "seconds_per_iteration": dts,
                              "what": f"sklearn KMeans lloyd float32, {km.n_iter_} iterations incl. its set-up"}
        except Exception as e:                       # report, never hide
            out["sklearn"] = {"error": repr(e)}
        # --- literal reference-shaped run (C1 only: [n, K, D] float64 fits)
        if cfg_name == "c1":
            Xl, Cl = cpu_sample(cfg_name, min(N, 250_000))
            t0 = time.perf_counter()
            oracle.literal_iteration(Xl, Cl, n_shards=1)
            dtl = time.perf_counter() - t0
            out["literal"] = {"value": Xl.shape[0] / dtl, "rows": int(Xl.shape[0]), "seconds_per_iteration": dtl,
                              "what": "reference-shaped [n,K,D] float64 broadcast (:221-230), one iteration"}
    return out


def cpu_baseline_record(cfg_name, arms):
    N, D, K, kind = CONFIGS[cfg_name]
    best = max((k for k in arms if "value" in arms[k] and k != "literal"), key=lambda k: arms[k]["value"])
    a = arms[best]
    return {
        "value": a["value"], "unit": UNIT, "cores": host_cores(), "kind": "port",
        "sample": f"{a['rows']} rows of the {cfg_name} workload (D={D}, K={K}); fastest CPU arm = {best}: {a['what']}",
        "blas_threads": blas_threads(),
        "arms": {k: ({"value": v["value"], "rows": v["rows"], "what": v["what"]} if "value" in v else v) for k, v in arms.items()},
    }


def run_reference_arm(args, cfg_name):
    """`--impl reference`: TensorFlow is not installable here, so the reference's CPU path is timed as its CPU
    restatements (see time_cpu_arms) with all host threads; `value` is the fastest of them."""
    N, D, K, kind = CONFIGS[cfg_name]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the BLAS / OpenMP pools of this process would be
    # created with one thread.  Run the measurement in a child with that variable removed.
    if os.environ.get("OMP_NUM_THREADS") and not os.environ.get("KMC_BENCH_CHILD"):
        import subprocess
        env = {k: v for k, v in os.environ.items() if k not in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS")}
        env["KMC_BENCH_CHILD"] = "1"
        for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID"):
            env.pop(k, None)
        r = subprocess.run([sys.executable, os.path.abspath(__file__)] + sys.argv[1:], env=env, capture_output=True, text=True)
        sys.stdout.write(r.stdout)
        sys.stderr.write(r.stderr)
        return
    arms = time_cpu_arms(cfg_name, max(1, args.steps), args.warmup)
    rec = cpu_baseline_record(cfg_name, arms)
    value = rec["value"]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * N / value,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64/f32 (see cpu_baseline.arms)",
        "data": "synthetic",
        "config": {"workload": workload_name(cfg_name), "N": N, "D": D, "K": K,
                   "note": "TensorFlow is not installable here; the reference's CPU path is timed as CPU implementations of "
                           "the same Lloyd iteration on a bounded sample; ms_per_step is that rate scaled to N rows"},
        "cpu_baseline": rec,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def ncu_traffic_bytes(kernel, n_local, D, K):
    """dram__bytes_read + dram__bytes_write of the dominant kernel per launch, from this round's committed
    `ncu --set full` capture (profiles/r2_traffic.json; a profile cannot be taken inside a timed run), scaled to
    the rows this rank's launch covers.  (bytes, source) -- (None, why) if no capture of this kernel and shape
    is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            t = json.load(f)[kernel]
        if t["D"] == D and t["K"] == K:
            return t["dram_bytes"] * (n_local / t["rows"]), t.get("source", "profiles/r2_traffic.json")
        return None, "profiles/r2_traffic.json holds another shape"
    except Exception:
        return None, "no capture committed for this kernel"


def workload_name(cfg_name):
    N, D, K, kind = CONFIGS[cfg_name]
    what = "synthetic 4K RGB image" if kind == "image" else "synthetic Gaussian blobs"
    return f"{cfg_name}: {what} N={N} D={D} K={K} fp32, one Lloyd iteration per step"


# ----------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--rows", type=int, default=0, help="override total rows (debugging only; marks the line invalid)")
    ap.add_argument("--path", default="auto", choices=["auto", "small", "simt", "tc"])
    ap.add_argument("--param", action="append", default=[], help="tuning: ID=VALUE passed to kmc_set_param")
    ap.add_argument("--data", default="blobs", choices=["blobs", "blobs4", "uniform"],
                    help="blobs: K blobs, sigma 1 (BASELINE.json); blobs4: K/4 blobs, so ~4 centroids share a blob and few "
                         "rows are certified by one MMA term; uniform: no cluster structure (speed envelope, not the headline)")
    ap.add_argument("--sigma", type=float, default=1.0, help="blob standard deviation (1.0 = BASELINE.json)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-single-process", action="store_true", help="skip the one-process / all-GPUs e2e leg (N > 1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    if args.impl == "reference":
        run_reference_arm(args, args.config)
        return

    import torch
    import torch.distributed as dist

    pkg = importlib.import_module("tensorflow-distributed-clustering_b200")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs CUDA devices (B200); the product has no CPU fallback")
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    N, D, K, kind = CONFIGS[args.config]
    if args.rows:
        N = args.rows
    r0, r1 = pkg.shard_ranges(N, world)[rank]
    n_local = r1 - r0

    # ---- synthetic data, generated on the device, resident for the whole run; rows [r0, r1) of ONE data set ----
    if kind == "image":
        X = pkg.datagen.image_torch(2160, 3840, SEED, dev)[r0:r1].contiguous()
        C0 = pkg.datagen.kmeanspp_init(X[:: max(1, n_local // 65536)].cpu().numpy(), K, SEED) if rank == 0 else None
    else:
        k_true = max(2, K // 4) if args.data == "blobs4" else K
        X, _ = pkg.datagen.blobs_torch(n_local, D, k_true, SEED, dev, sigma=args.sigma, row0=r0,
                                       kind="uniform" if args.data == "uniform" else "blobs")
        C0 = X[:K].cpu().numpy() if rank == 0 else None          # global rows 0..K-1 (reference convention, :325)
    c0_t = torch.from_numpy(C0).to(dev) if rank == 0 else torch.empty((K, D), dtype=torch.float32, device=dev)
    if world > 1:
        dist.broadcast(c0_t, 0)
    C0 = c0_t.cpu().numpy()

    shard = pkg.LloydShard(n_local, D, K, dev, path=args.path)
    shard.bind_points(X)
    shard.set_centroids(C0)
    shard.enable_profile(True)
    for kv in args.param:
        pid, val = kv.split("=")
        shard.set_param(int(pid), float(val))
    torch.cuda.synchronize(dev)

    def one_step():
        shard.accumulate()
        if world > 1:
            dist.all_reduce(shard.partials)
        shard.update()

    def run_steps(k):
        # one GPU: the product's own loop (kmc_iterate: no host round trip between iterations); several GPUs: accumulate / NCCL all-reduce / update per iteration
        if world == 1:
            if k > 0:
                shard.iterate(k)
        else:
            for _ in range(k):
                one_step()

    run_steps(args.warmup)
    torch.cuda.synchronize(dev)
    shard.phase_ms()                                   # reset the phase ring after warm-up
    launches0 = shard.stats()["launches"]

    sampler = ClockSampler(local_rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run_steps(args.steps)
    ev1.record()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    phases = shard.phase_ms()
    stats = shard.stats()
    launches = stats["launches"] - launches0 + (args.steps if world > 1 else 0)

    ms_per_step = ms_total / max(args.steps, 1)
    value = N * args.steps / (ms_total * 1e-3)

    # ---- checks at the full size: counts, and the SAME result whatever the number of GPUs ----
    part = shard.partials.clone()                      # all-reduced: global sums / counts / inertia of the last step
    counts_sum = float(part[K * D:K * D + K].sum().item())
    cent = shard.centroids_dev.double()
    weights = torch.arange(1, K * D + 1, dtype=torch.float64, device=dev).view(K, D) / float(K * D)
    checks = {"counts_sum_equals_N": bool(counts_sum == N),
              "inertia": float(part[K * D + K].item()),
              "centroid_checksum": float(cent.nan_to_num().sum().item()),
              "centroid_checksum_weighted": float((cent.nan_to_num() * weights).sum().item()),
              "changed_rows_last_step": stats["changed_rows"], "full_pass_last_step": stats["full_pass"]}

    # ---- parity of one more full iteration at the full size against the float64 mirror of the oracle (GPU) ----
    if not args.no_parity:
        checks["parity"] = parity_check(shard, X, K, D, dev, world)

    # ---- roofline of the dominant kernel (the assignment kernel) ----
    peaks = measured_peaks()
    t_assign = phases["assign"] * 1e-3
    fused_small = stats["path"] == "small" and world == 1
    if fused_small:
        # kmc_iterate on the small-D path is two back-to-back launches per iteration (the streaming kernel: assignment +
        # accumulation; a one-CTA fold + update kernel overlapped with it by programmatic dependent launch), with
        # no events inside the library: the pair's duration is the step's
        t_assign = ms_per_step * 1e-3
        phases = dict(phases, assign=ms_per_step)
    flops = 2.0 * n_local * K * D
    bytes_alg = 4.0 * n_local * D
    t_tensor, t_hbm = flops / (peaks["tflops"] * 1e12), bytes_alg / (peaks["hbm_gbs"] * 1e9)
    if t_tensor >= t_hbm:
        roofline = {"bound": "tensor", "achieved": flops / t_assign / 1e12 if t_assign > 0 else None,
                    "peak": peaks["tflops"], "unit": "TFLOP/s"}
    else:
        roofline = {"bound": "hbm", "achieved": bytes_alg / t_assign / 1e9 if t_assign > 0 else None,
                    "peak": peaks["hbm_gbs"], "unit": "GB/s"}
    roofline["frac"] = (roofline["achieved"] / roofline["peak"]) if roofline["achieved"] else None
    roofline["kernel"] = {"tc": "tc_assign_kernel", "small": "small_kernel", "simt": "simt_assign_kernel"}[stats["path"]]
    roofline["traffic"], roofline["traffic_source"] = ncu_traffic_bytes(roofline["kernel"], n_local, D, K)
    roofline["peak_source"] = peaks["source"] + (" bf16 dense, sustained (kernel timed inside a long step)"
                                                 if roofline["bound"] == "tensor" else " copy bandwidth")
    roofline["kernel_ms"] = phases["assign"]
    roofline["phase_ms"] = {k: round(v, 4) for k, v in phases.items() if k != "calls"}
    # whole-iteration roofline of BASELINE.json: slower of 2NKD at peak and 4ND at HBM bandwidth
    roofline["step_frac"] = max(t_tensor, t_hbm) / (ms_per_step * 1e-3)

    result = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.config), "N": N, "D": D, "K": K, "rows_per_gpu": n_local,
                   "kernel_path": stats["path"], "seed": SEED, "parallelism": f"rows sharded over {world} GPU(s)",
                   "l2": "inputs (%.1f GB per GPU) exceed the 126 MB L2; no flush needed" % (n_local * D * 4 / 1e9)
                   if n_local * D * 4 > 4 * 126e6 else "inputs fit in L2: numbers are L2-resident",
                   "refined_rows_last_step": stats["ambiguous"], "walk_overflow_rows_last_step": stats["walk_overflow"],
                   "checks": checks,
                   "iterations_timed": f"Lloyd iterations {args.warmup + 1}..{args.warmup + args.steps} of one run from C0 = X[0:K] "
                                       "(the e2e leg times the whole run from iteration 1)",
                   "accumulation": "incremental: running fp64 sums updated with the rows whose label changed; a device-side "
                                   "gate takes the full pass when more than a quarter of the rows moved (every N*K distance is "
                                   "evaluated either way)" if stats["path"] != "small" else "fused into the assignment kernel",
                   "tensor_terms": "adaptive: 1 fp16 MMA term (fp16 accumulators for K > 256, else fp32) + exact refine of the uncertified rows; device-side fall-back to fp32 accumulators / the 3-term split where the refine would cost more"
                   if stats["path"] == "tc" else None},
        "roofline": roofline, "gpu_launches": int(launches), "clocks": clocks,
    }
    if args.data != "blobs" or args.sigma != 1.0:
        result["config"]["data_variant"] = {"data": args.data, "sigma": args.sigma,
                                            "note": "speed-envelope run on harder data, not the BASELINE.json workload"}
    if args.rows:
        result["config"]["INVALID"] = "--rows override: not the BASELINE.json configuration"

    # ---- end to end through the public entry point, host buffers, copies inside the timed region ----
    if not args.no_e2e:
        host, note = pinned_host_copy(X, n_local, D, K, dev)
        shard.close()
        del shard, X, part, cent, weights
        torch.cuda.empty_cache()
        result["e2e"] = run_e2e(pkg, host, note, C0, D, K, args.steps, dev, world, N)
        if world == 1:
            result["e2e_pageable_f64"] = run_e2e_pageable_f64(pkg, host, C0, D, K, args.steps, dev)
        del host
        if world > 1 and not args.no_single_process:
            # the reference's own call shape: ONE process, GPU_names listing every device (:180-263)
            sp = run_e2e_single_process(pkg, args, N, D, K, kind, C0, dev, world, rank)
            if rank == 0:
                result["e2e_single_process"] = sp
    else:
        result["e2e"] = None

    # ---- CPU baseline: CPU implementations of the same iteration on the host cores, bounded sample (rank 0, N = 1) ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        result["cpu_baseline"] = cpu_baseline_record(args.config, time_cpu_arms(args.config, 3, 1))
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()


def parity_check(shard, X, K, D, dev, world):
    """One more Lloyd iteration at the FULL size, checked against the float64 mirror of the oracle
    (oracle/gpu_mirror.py, the checker -- plain torch float64 on the same GPU): every row's label outside
    near-ties (<= 1e-5 relative), counts, per-cluster sums, inertia of this rank's shard; with several ranks the
    figures are the worst over ranks.  Reference :221-248 on the same rows and centroids."""
    import torch
    import torch.distributed as dist
    from oracle import gpu_mirror

    C = shard.centroids_dev.clone()
    shard.accumulate()                                   # local sums / counts / inertia (not all-reduced)
    torch.cuda.synchronize(dev)
    part = shard.partials.clone()
    labels = shard.last_labels()
    stats = shard.stats()
    chunk = max(1 << 14, min(1 << 18, (1 << 28) // max(K, 1)))       # ~2 GB of float64 distances per chunk
    ref = gpu_mirror.lloyd_pass_f64(X, C, chunk_rows=chunk)
    out = gpu_mirror.compare_pass(ref, labels, part[:K * D].view(K, D), part[K * D:K * D + K],
                                  float(part[K * D + K].item()), X=X)
    out["pass_was_incremental"] = not stats["full_pass"]
    del ref
    if world > 1:
        keys = [k for k in out if isinstance(out[k], (int, float)) and not isinstance(out[k], bool)]
        t = torch.tensor([float(out[k]) for k in keys], dtype=torch.float64, device=dev)
        tsum = t.clone()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        for i, k in enumerate(keys):
            summed = k in ("rows_checked", "label_mismatches_outside_near_ties", "label_differences_inside_near_ties",
                           "near_tie_rows", "counts_abs_diff")
            out[k] = (int(tsum[i].item()) if k != "counts_abs_diff" else float(tsum[i].item())) if summed else float(t[i].item())
        # the shard keeps going: this pass must not leave different centroids on different ranks
        dist.all_reduce(shard.partials)
    shard.update()
    # The result must be the exact computation under a labelling the parity rule allows: labels identical to the
    # oracle's outside near-ties, per-cluster sums equal to the float64 sums of the rows under the product's OWN labels
    # (1e-5: the small-D path adds up to ~100 rows per thread in fp32 before it goes to fp64), counts off by at most
    # the near-tie rows, inertia within 1e-4.  `sums_rel_err` (against the oracle's sums) is reported next to it: it
    # contains the rows that may differ, 2e-5 on the BASELINE.json blobs, 1e-4 on uniform data with 6e4 near-ties.
    out["ok"] = bool(out["label_mismatches_outside_near_ties"] == 0 and out["sums_rel_err_own_labels"] <= 1e-5
                     and out["inertia_rel_err"] <= 1e-4 and out["counts_abs_diff"] <= 2 * out["near_tie_rows"])
    return out


def pinned_host_copy(X_dev, N, D, K, dev):
    """The synthetic points copied to pinned host memory (untimed set-up of the e2e leg)."""
    import psutil
    import torch

    need = N * D * 4
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # every rank of the node pins its own copy at the same time: budget half of this rank's share of the host RAM
    avail = psutil.virtual_memory().available / world * 0.5 if world > 1 else psutil.virtual_memory().available
    rows, note = N, None
    if avail < need * 1.3 + (8 << 30) / world:
        rows = int(max(K, min(N, (avail - (8 << 30) / world) / 1.3 // (D * 4))))
        note = f"host RAM budget of this rank {avail / 1e9:.0f} GB < needed: e2e uses the first {rows} rows"
    host = torch.empty((rows, D), dtype=torch.float32, pin_memory=True)
    host.copy_(X_dev[:rows])
    torch.cuda.synchronize(dev)
    return host, note


def run_e2e(pkg, host, note, C0, D, K, steps, dev, world, n_total):
    """The same metric through the reference-facing entry point (`distribuited_k_means`; with one
    process per GPU its SPMD form), points in pinned HOST memory: the timed region holds the pinned
    H2D staging of every row, the one-time split of the points, `steps` Lloyd iterations (with the
    NCCL all-reduce), and the D2H of the centroids and of the labels (the reference's `cluster_idx`).
    Wall clock between barriers, max over ranks.  Bytes are per job and GPU divided by `steps`."""
    import torch
    import torch.distributed as dist

    rows = host.shape[0]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    if world > 1:
        res = pkg.distribuited_k_means_spmd(host, K, C0, steps, device=dev)
        dist.barrier()
    else:
        res = pkg.distribuited_k_means(host, K, [f"/device:GPU:{dev.index}"], C0, steps)
    dt = time.perf_counter() - t0
    total_rows = rows
    if world > 1:
        t = torch.tensor([dt, float(rows)], dtype=torch.float64, device=dev)
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dt, total_rows = float(tmax[0].item()), int(t[1].item())
    out = {
        "value": total_rows * res["n_iter"] / dt, "unit": UNIT,
        "h2d_bytes_per_step": int(rows * D * 4 / max(steps, 1)),
        "d2h_bytes_per_step": int((K * D * 4 + rows * 8) / max(steps, 1)),   # centroids fp32 + labels int64
        "rows": rows, "iterations": res["n_iter"], "seconds": dt,
        "initialization_time": res["initialization_time"], "computation_time": res["computation_time"],
        "inertia": res["inertia"],
    }
    if note:
        out["note"] = note
    return out


def run_e2e_single_process(pkg, args, N, D, K, kind, C0, dev, world, rank):
    """`distribuited_k_means(host_rows, K, ['/device:GPU:0', ...], C0, steps)` from ONE process driving all the
    GPUs of the node -- the reference's call shape -- while the other ranks of this torchrun job idle: rank 0
    regenerates the whole data set into pinned host memory (bounded by the host RAM), then times the call (shards
    staged side by side by one host thread per GPU, first iteration under the copies, NCCL all-reduce from one
    process, labels back)."""
    import psutil
    import torch
    import torch.distributed as dist

    out = None
    torch.cuda.empty_cache()
    torch.cuda.synchronize(dev)
    # the idle ranks wait on the host (gloo): an NCCL barrier would park a spinning kernel on every GPU that
    # rank 0 is about to use from its own process
    cpu_group = dist.new_group(backend="gloo")
    dist.barrier(group=cpu_group)
    if rank == 0:
        try:
            avail = psutil.virtual_memory().available
            rows = int(min(N, max(K * world, (avail * 0.4) // (D * 4))))
            host = torch.empty((rows, D), dtype=torch.float32, pin_memory=True)
            step = min(rows, 1 << 24)
            for r0 in range(0, rows, step):
                r1 = min(rows, r0 + step)
                if kind == "image":
                    blk = pkg.datagen.image_torch(2160, 3840, SEED, dev)[r0:r1]
                else:
                    k_true = max(2, K // 4) if args.data == "blobs4" else K
                    blk, _ = pkg.datagen.blobs_torch(r1 - r0, D, k_true, SEED, dev, sigma=args.sigma, row0=r0,
                                                     kind="uniform" if args.data == "uniform" else "blobs")
                host[r0:r1].copy_(blk)
                del blk
            torch.cuda.synchronize(dev)
            torch.cuda.empty_cache()
            names = [f"/device:GPU:{i}" for i in range(world)]
            # untimed warm-up on a few rows: this process has used one GPU so far; the CUDA contexts of the others and
            # the single-process NCCL communicator are created on first use (about a second), once per process
            warm = min(rows, max(K * world, 1 << 18))
            pkg.distribuited_k_means(host[:warm], K, names, C0, 2)
            t0 = time.perf_counter()
            res = pkg.distribuited_k_means(host, K, names, C0, args.steps)
            dt = time.perf_counter() - t0
            out = {"value": rows * res["n_iter"] / dt, "unit": UNIT, "rows": rows, "gpus": world, "iterations": res["n_iter"],
                   "seconds": dt, "initialization_time": res["initialization_time"],
                   "computation_time": res["computation_time"], "inertia": res["inertia"],
                   "h2d_bytes_per_step": int(rows * D * 4 / max(args.steps, 1)),
                   "d2h_bytes_per_step": int((K * D * 4 + rows * 8) / max(args.steps, 1)),
                   "first_iteration_streamed": res.get("first_iteration_streamed")}
            del host, res
        except Exception as e:                      # a leg of its own: report, do not lose the line
            out = {"error": repr(e)}
    dist.barrier(group=cpu_group)
    return out


def run_e2e_pageable_f64(pkg, host_f32, C0, D, K, steps, dev):
    """The drop-in's real input (reference :303, notebooks/Testing Images.ipynb:329): an ordinary PAGEABLE float64
    NumPy array handed to `distribuited_k_means`.  The array is a float64 copy of a bounded number of rows of the
    same workload (a float64 copy of all 100M rows of config 4 would need 102 GB of host memory); the timed
    region is the whole call: host threads convert to fp32 into pinned chunks, H2D, plane split,
    `steps` iterations, centroids and int64 labels back."""
    import psutil
    import torch

    avail = psutil.virtual_memory().available
    rows = int(min(host_f32.shape[0], 16 << 20, max(K, (avail - (8 << 30)) // (3 * D * 8))))
    X64 = host_f32[:rows].numpy().astype(np.float64)          # pageable
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    res = pkg.distribuited_k_means(X64, K, [f"/device:GPU:{dev.index}"], C0.astype(np.float64), steps)
    dt = time.perf_counter() - t0
    return {"value": rows * res["n_iter"] / dt, "unit": UNIT, "rows": rows, "iterations": res["n_iter"], "seconds": dt,
            "h2d_bytes_per_step": int(rows * D * 4 / max(steps, 1)),     # the host threads convert to fp32 on their way into the pinned chunks
            "host_bytes_read_per_step": int(rows * D * 8 / max(steps, 1)),
            "d2h_bytes_per_step": int((K * D * 4 + rows * 8) / max(steps, 1)),
            "initialization_time": res["initialization_time"], "computation_time": res["computation_time"],
            "input": "pageable float64 NumPy array (the reference's callers' dtype)"}


if __name__ == "__main__":
    main()
