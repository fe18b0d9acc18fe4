#!/usr/bin/env python
"""bench.py — radar path samples/s on the car-body configuration (BASELINE.json configs[1]).

One "step" = one bf_receive of `--samples` radar path samples per GPU on the committed scene
(max_depth 4 = 3 bounces, 1 Tx / 1 Rx; car-body mesh = procedural stand-in, the named .ply is missing
from the reference tree), followed at N > 1 by one NCCL reduce of the per-rank range profiles.
Samples are sharded across ranks by sample index (rank r takes [r*n, (r+1)*n) of every step): weak scaling.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--samples S] [--impl native|reference]

`--impl reference` times the reference's CPU path for the same metric on the host cores: the oracle's
scalar sample loop calling the reference's own Embree 3.8.0 (oracle/_ref, built from /root/reference/ext/embree)
when present, else the oracle's BVH.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "radar path samples/s"
UNIT = "samples/s"
MAX_DEPTH = 4
SPECTRUM_WIDTH = 1


def algorithmic_bytes(n_tris, channels):
    """B_path = 80*ceil(log8(N/4)); closest = B_path + 48, shadow = B_path."""
    import math
    b_path = 80 * math.ceil(math.log(max(n_tris, 8) / 4.0, 8))
    return b_path, b_path + 48, b_path, 8 * channels


class ClockSampler:
    """SM clock / throttle reasons sampled through NVML every 10 ms while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.stop_flag, self.thread, self.err = [], set(), False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()
        except Exception as e:   # noqa: BLE001
            self.err = f"nvml unavailable: {e}"

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception as e:   # noqa: BLE001
                self.err = str(e)
                return
            time.sleep(0.01)

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=1)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [self.err or "no samples"], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": float(self.max_sm), "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_reference_run(scene_blob, n_samples, steps, warmup, threads):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import OracleScene, EMBREE
    use_embree = os.path.exists(EMBREE)
    orc = OracleScene(scene_blob, accel=OracleScene.EMBREE_ACCEL if use_embree else OracleScene.BVH)
    for i in range(warmup):
        orc.receive(max(n_samples // 8, 1024), seed=100 + i, threads=threads)
    t0 = time.perf_counter()
    q = 0
    for i in range(steps):
        orc.receive(n_samples, seed=i, first=0, threads=threads)
        q += orc.last_queries
    dt = time.perf_counter() - t0
    kind = "reference" if use_embree else "port"
    what = ("oracle scalar sample loop + the reference's Embree 3.8.0 traversal (oracle/_ref/libembree3.so)" if use_embree
            else "oracle scalar sample loop + oracle BVH (Embree build absent)")
    return n_samples * steps / dt, dt, kind, what, q / max(steps * n_samples, 1)


class DevArray:
    """__cuda_array_interface__ view of the library's device profile (for torch / NCCL)."""
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--samples", type=int, default=1 << 28, help="radar path samples per step per GPU")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-samples", type=int, default=0, help="samples per step of the CPU arm (0 = calibrate to ~10 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    steps, warmup = args.steps, max(args.warmup, 0)

    from beifong_b200 import configs
    scene = configs.c2_car_body(spectrum_width=SPECTRUM_WIDTH, max_depth=MAX_DEPTH)
    workload = {"workload": "BASELINE configs[1]: car-body mesh radar range profile, 1 Tx / 1 Rx, 3 bounces (max_depth 4), "
                            "94 GHz FMCW wignertransmitter/wignerreceiver, 1024 range bins; Car-body.ply is missing from the "
                            "reference tree -> procedural stand-in mesh of 69 696 triangles + 4 rectangles",
                "spectrum_width": SPECTRUM_WIDTH, "max_depth": MAX_DEPTH, "triangles": 69696}
    threads = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return 0
        blob = scene.export()
        n = args.cpu_samples
        if n <= 0:
            rate, _, _, _, _ = cpu_reference_run(blob, 1 << 16, 1, 0, threads)
            n = int(max(1 << 16, min(1 << 31, rate * 20.0 / max(steps + min(warmup, 1) / 8.0, 1))))
        value, dt, kind, what, qps = cpu_reference_run(blob, n, steps, min(warmup, 1), threads)
        out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
               "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
               "data": "synthetic", "config": dict(workload, samples_per_step=n, parallelism="cpu threads"),
               "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": f"{steps} x {n} samples; {what}"},
               "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "queries_per_sample": qps}
        print(json.dumps(out))
        return 0

    # Everything below may print to stdout from native code (NCCL prints its version line there): keep fd 1 for the one JSON
    # line and send the rest to stderr.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    scene.commit(local_rank)
    n = args.samples
    ptr, nfloats, stream_ptr = scene.profile_device()
    stream = torch.cuda.ExternalStream(stream_ptr, device=torch.device("cuda", local_rank))
    prof_t = torch.as_tensor(DevArray(ptr, nfloats), device=torch.device("cuda", local_rank))

    def step(i, e2e=False):
        if e2e:
            scene.reupload()                                   # host -> device: flattened scene from pinned memory
        scene.receive(n, seed=i, first_sample=rank * n, accumulate=False)
        if world > 1:
            with torch.cuda.stream(stream):
                dist.reduce(prof_t, dst=0)
        if e2e:
            return scene.profile()                             # device -> host: the range profile

    def sync_all():
        scene.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for i in range(warmup):
        step(1000 + i)
    sync_all()
    st0 = scene.stats()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for i in range(steps):
        step(i)
    ev1.record(stream)
    sync_all()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clk = clocks.stop() if clocks else None
    st1 = scene.stats()
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    value = world * n * steps / (ms * 1e-3)

    # ---- end to end through the C ABI with host buffers -------------------------------------------------
    geom_bytes = scene.geometry_bytes()
    step(2000, e2e=True)
    sync_all()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i, e2e=True)
    sync_all()
    e2e_s = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = world * n * steps / float(e2e_s.item())

    out = None
    if rank == 0:
        # ---- roofline of the dominant kernel (closest-hit traversal), timed live with CUDA events ----------
        shape = scene.profile_shape()
        b_path, b_closest, b_shadow, b_bins = algorithmic_bytes(workload["triangles"], shape[2])
        scene.set_trace_timing(True)
        km, q_closest, q_shadow, defer_ms = {}, 0, 0, [0.0, 0]
        reps = 3
        for i in range(reps):
            a = scene.stats()
            scene.receive(n, seed=3000 + i, first_sample=0)
            scene.synchronize()
            b = scene.stats()
            q_closest += b["closest_queries"] - a["closest_queries"]
            q_shadow += b["shadow_queries"] - a["shadow_queries"]
            for k, (m, c) in scene.last_kernel_ms().items():
                km.setdefault(k, [0.0, 0])
                km[k][0] += m; km[k][1] += c
            # a k_trace<closest> launch right behind a k_primary launch is the second half of the primary query (the rays that
            # enter the mesh bounds, traversed densely: k_primary<DEFER>)
            ll = scene.last_launch_list()
            for j in range(1, len(ll)):
                if ll[j][0] == "trace_closest" and ll[j - 1][0] == "primary":
                    defer_ms[0] += ll[j][1]; defer_ms[1] += 1
        scene.set_trace_timing(False)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        # Dominant kernel: k_primary (receiver ray generation + first closest-hit query + compaction, fused); it issues one
        # closest-hit query per sample.  Later closest-hit launches (k_trace<closest>) are reported beside it.
        pr_ms, pr_launch = km.get("primary", (0.0, 0))
        cl_ms, cl_launch = km["trace_closest"]
        if pr_launch and defer_ms[1]:
            # the primary query is the launch pair: its time is the sum, the later-depth launches exclude the second half
            dom_name = "k_primary<DEFER> + k_trace<closest> over the rays that enter the mesh bounds (raygen + closest-hit query + compaction)"
            dom_ms, dom_launch, dom_queries = pr_ms + defer_ms[0], pr_launch, reps * n
            cl_ms, cl_launch = cl_ms - defer_ms[0], cl_launch - defer_ms[1]
        elif pr_launch:
            dom_name, dom_ms, dom_launch, dom_queries = "k_primary (raygen + closest-hit query + compaction)", pr_ms, pr_launch, reps * n
        else:
            dom_name, dom_ms, dom_launch, dom_queries = "k_trace<closest>", cl_ms, cl_launch, q_closest
        achieved = dom_queries * b_closest / (dom_ms * 1e-3) / 1e9
        total_ms = sum(v[0] for v in km.values())
        traffic = None
        try:   # DRAM bytes per launch of the same kernel from the committed ncu --set full capture, scaled to this launch size
            tr = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
            traffic = tr["dram_bytes_per_query"] * dom_queries / max(dom_launch, 1)
        except Exception:
            pass
        roofline = {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_query": b_closest, "queries_per_launch": dom_queries / max(dom_launch, 1),
                    "avg_launch_ms": dom_ms / max(dom_launch, 1), "kernel_share_of_step": dom_ms / max(total_ms, 1e-9),
                    "later_closest_launches": {"queries": q_closest - (reps * n if pr_launch else 0), "ms": cl_ms, "launches": cl_launch,
                                               "achieved": (q_closest - (reps * n if pr_launch else 0)) * b_closest / max(cl_ms * 1e-3, 1e-12) / 1e9 if pr_launch else None},
                    "kernel_ms_per_step": {k: v[0] / reps for k, v in km.items()},
                    "primary_pair_ms_per_step": {"k_primary": pr_ms / reps, "k_trace_deferred": defer_ms[0] / reps} if defer_ms[1] else None,
                    "queries_per_sample": {"closest": q_closest / (reps * n), "shadow": q_shadow / (reps * n)},
                    "algorithmic_bytes_per_sample": (q_closest * b_closest + q_shadow * b_shadow) / (reps * n) + b_bins,
                    "whole_step_frac": ((q_closest * b_closest + q_shadow * b_shadow) / (reps * n) + b_bins) * (n * steps / (ms * 1e-3)) / 1e9 / peak}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            blob = scene.export()
            rate, _, _, _, _ = cpu_reference_run(blob, 1 << 16, 1, 0, threads)
            cn = args.cpu_samples or int(max(1 << 16, min(1 << 31, rate * 12.0)))
            v, dt, kind, what, _ = cpu_reference_run(blob, cn, 1, 0, threads)
            cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": f"1 x {cn} samples of the same workload ({dt:.1f} s); {what}"}
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": dict(workload, samples_per_step_per_gpu=n, parallelism=f"sample-index sharding x{world}, scene replicated, 1 NCCL reduce/step",
                              l2="wave state (132 B per surviving sample, 2^25 samples per wave, 8 waves in flight) is streamed and exceeds L2 by "
                                 "two orders of magnitude; a new seed every step; the 8.6 MB scene stays L2-resident by design"),
               "clocks": clk, "gpu_launches": int(launches),
               "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(geom_bytes), "d2h_bytes_per_step": int(nfloats * 4)},
               "roofline": roofline, "cpu_baseline": cpu}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    scene.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
