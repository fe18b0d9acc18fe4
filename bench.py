#!/usr/bin/env python
"""bench.py — radar path samples/s on the configurations BASELINE.json names (default: configs[1], the car-body scene).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config c1|c2|c3|c4|c5] [--samples S] [--scaling weak|strong]
                  [--mesh procedural|FILE.ply] [--impl native|reference] [--no-cpu-baseline]

One "step":
  c1 / c2 / c3   one bf_receive of S radar path samples per GPU on the committed scene (+ ONE reduce of the per-rank profiles at N > 1,
                 bf_multi_reduce = ncclReduce inside the library).  Samples are sharded by sample index: weak scaling (S per GPU,
                 the default) or strong scaling (S in total, --scaling strong).
  c4             the 256 frames of the moving-target sequence, S samples each (default 1000, animated_trans_rad.py:39), as one
                 bf_receive_batch per GPU; frames are dealt round-robin to the GPUs, no reduce (a frame stays on one GPU).
  c5             the 64 elements of the receive array, S samples each (default 2^20), one bf_receive_batch per GPU; elements are dealt
                 round-robin to the GPUs, no reduce.
`value` is device-timed (CUDA events on the scene's stream, scene resident in HBM, max over ranks); `e2e` is the same work through
the C ABI with host buffers inside the timed region (geometry re-upload from pinned memory / per-item poses in, profiles out).

`--impl reference` times the reference's CPU path for the same metric and config on the host cores: the oracle's scalar sample loop
calling the reference's own Embree 3.8.0 (oracle/_ref/libembree3.so, built from /root/reference/ext/embree) when present, else the
oracle's BVH.  Its scene comes from a blob file written at build time (oracle/_ref/scenes/), so that arm never maps the product library.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "radar path samples/s"
UNIT = "samples/s"
SCENE_DIR = os.path.join(ROOT, "oracle", "_ref", "scenes")
DEFAULT_SAMPLES = {"c1": 1 << 26, "c2": 1 << 28, "c3": 1 << 26, "c4": 1000, "c5": 1 << 20}
N_ITEMS = {"c4": 256, "c5": 64}


def algorithmic_bytes(n_tris, channels):
    """B_path = 80*ceil(log8(N/4)); closest = B_path + 48, shadow = B_path; no BVH (C1 / C4): 3 x 64 B matrices."""
    import math
    if n_tris <= 0:
        return 192, 192, 192, 8 * channels
    b_path = 80 * math.ceil(math.log(max(n_tris, 8) / 4.0, 8))
    return b_path, b_path + 48, b_path, 8 * channels


L2_NOTE = ("inputs are generated on the device from (seed, sample index); a new seed every step; per-step state written and read "
           "(spilled lanes + k_path parking areas, >= 0.3 GB per 2^28 samples) exceeds L2; the scene itself is meant to stay L2-resident")


def workload_of(name, mesh="procedural", mesh_tris=None):
    """The workload description of BASELINE config `name` (identical in both arms; no scene is built)."""
    if name == "c1":
        w = {"workload": "BASELINE configs[0]: python_scripts/trans_rad.xml verbatim (3 rectangles, spot emitter, fluxmeter, "
                         "time(pathtime) 50 bins x 0.5 ns, max_depth -1, rr_depth 5)", "name": "c1", "spectrum_width": 1, "max_depth": -1, "triangles": 0}
    elif name == "c2":
        if mesh != "procedural":
            what = f"mesh file {os.path.basename(mesh)} scaled to 4.5 m in the car pose of Render.py:305-316"
            tris = int(mesh_tris) if mesh_tris is not None else -1
        else:
            what, tris = "Car-body.ply is missing from the reference tree -> procedural stand-in mesh of 69 696 triangles", 69696
        w = {"workload": "BASELINE configs[1]: car-body mesh radar range profile, 1 Tx / 1 Rx, 3 bounces (max_depth 4), 94 GHz FMCW "
                         "wignertransmitter/wignerreceiver, 1024 range bins; " + what + " + 4 rectangles",
             "name": "c2", "spectrum_width": 1, "max_depth": 4, "triangles": tris}
    elif name == "c3":
        w = {"workload": "BASELINE configs[2]: multi-object mesh scene (Motorbike + Bus are missing from the reference tree -> 14 procedural "
                         "bodies, 975 744 triangles), 1 Tx / 1 Rx, max_depth 4, 1024 x 1024 ADC frame (time x beat frequency, Receive.ipynb cell 7)",
             "name": "c3", "spectrum_width": 1, "max_depth": 4, "triangles": 975744}
    elif name == "c4":
        w = {"workload": "BASELINE configs[3]: python_scripts/animated_trans_rad.py moving-target sequence, 256 frames (Tx + Rx rotate about z), "
                         "3 rectangles, area emitter, perspective sensor, range(pathlength) dr 0.2 x 50 bins, max_depth -1",
             "name": "c4", "spectrum_width": 1, "max_depth": -1, "triangles": 0, "frames": 256}
    elif name == "c5":
        w = {"workload": "BASELINE configs[4]: synthetic urban scene, 100 x 100 tessellated blocks = 1.0e7 triangles, 1 wignertransmitter, "
                         "64 receive elements on a line (lambda / 2 pitch), each its own 1024-bin range profile, max_depth 4",
             "name": "c5", "spectrum_width": 1, "max_depth": 4, "triangles": 10000002, "elements": 64}
    else:
        raise ValueError(name)
    w["l2"] = L2_NOTE
    return w


def mesh_triangles(mesh):
    if mesh == "procedural":
        return None
    from beifong_b200 import meshes
    return int(len(meshes.load_ply(mesh)[1]))


def build_config(name, mesh="procedural"):
    """(un-committed scene, workload description, ids) of BASELINE config `name`.  Host only: no GPU needed."""
    from beifong_b200 import configs, meshes
    ids, tris = {}, None
    if name == "c1":
        s = configs.c1_trans_rad(spectrum_width=1)
    elif name == "c2":
        if mesh != "procedural":
            p, f = meshes.load_ply(mesh)[:2]
            lo, hi = p.min(axis=0), p.max(axis=0)
            p = (p - 0.5 * (lo + hi)) * (4.5 / float((hi - lo).max()))          # longest side 4.5 m, then the car pose
            p = meshes.transform_points((configs.car_pose() * meshes.Transform.rotate([1, 0, 0], 90.0)).matrix32(), p.astype(np.float32))
            s = configs.radar_scene(p, f.astype(np.uint32), spectrum_width=1, max_depth=4)
            tris = len(f)
        else:
            s = configs.c2_car_body(spectrum_width=1, max_depth=4)
    elif name == "c3":
        s = configs.c3_multi_object(1)
        s.set_adc(t_bins=1024, f_bins=1024, t_bandwidth=200e-6, f_bandwidth=1024 * 0.03 * 2.0 / configs.C0 * (6e9 / 200e-6))
    elif name == "c4":
        s, ids["tx"] = configs.c4_animated(frame=0)
    elif name == "c5":
        s, ids["rx"] = configs.c5_urban()
    else:
        raise ValueError(name)
    return s, workload_of(name, mesh, tris), ids


def batch_items(name, ids, which):
    """bf_receive_batch items of config c4 (frames) / c5 (receive elements) for the item indices `which`."""
    from beifong_b200 import configs, meshes
    items = []
    for k in which:
        if name == "c4":
            pose = configs.c4_pose(k)
            items.append(dict(seed=0, rects=[(ids["tx"], (pose * meshes.Transform.scale([0.5, 0.5, 1.0])).matrix32())],
                              receiver=(None, configs.c4_receiver(pose))))
        else:
            items.append(dict(seed=k, rects=[(ids["rx"], configs.c5_element_pose(k))]))
    return items


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


# ---- the CPU arm (oracle = test infrastructure; this is one of the two places bench.py may execute it) ---------------------------
def scene_blob(name, mesh, scene=None):
    """Flat scene blob of config `name` for the oracle: the file written by __graft_entry__.build() when present, else an export
    through the library (host only).  Returns (blob, source description)."""
    path = os.path.join(SCENE_DIR, f"{name}.blob")
    if mesh == "procedural" and os.path.exists(path):
        return np.fromfile(path, dtype=np.uint8), f"blob file oracle/_ref/scenes/{name}.blob"
    if scene is None:
        scene, _, _ = build_config(name, mesh)
    return scene.export(), "bf_scene_export of the host-side scene (blob file absent)"


def cpu_run(blob, n_samples, steps, warmup, threads, accel="auto"):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import OracleScene, EMBREE
    use_embree = accel == "auto" and os.path.exists(EMBREE)
    a = {"auto": OracleScene.EMBREE_ACCEL if use_embree else OracleScene.BVH, "bvh": OracleScene.BVH, "brute": OracleScene.BRUTE}[accel]
    orc = OracleScene(blob, accel=a)
    for i in range(warmup):
        orc.receive(max(n_samples // 8, 1024), seed=100 + i, threads=threads)
    t0 = time.perf_counter()
    q = 0
    for i in range(steps):
        orc.receive(n_samples, seed=i, first=0, threads=threads)
        q += orc.last_queries
    dt = time.perf_counter() - t0
    orc.close()
    kind = "reference" if use_embree else "port"
    what = {"auto": "oracle scalar sample loop + the reference's Embree 3.8.0 traversal (oracle/_ref/libembree3.so)" if use_embree
            else "oracle scalar sample loop + oracle BVH (Embree build absent)",
            "bvh": "oracle scalar sample loop + oracle BVH", "brute": "oracle scalar sample loop + brute-force intersector (stands in for the kd-tree)"}[accel]
    return n_samples * steps / dt, dt, kind, what, q / max(steps * n_samples, 1)


def reference_arm(args, workload, threads):
    blob, src = scene_blob(args.config, args.mesh)
    steps, warmup = args.steps, max(args.warmup, 0)
    n = args.cpu_samples
    if n <= 0:
        rate, _, _, _, _ = cpu_run(blob, 1 << 19, 1, 0, threads)
        n = int(max(1 << 15, min(1 << 31, rate * 20.0 / max(steps + min(warmup, 1) / 8.0, 1))))
    value, dt, kind, what, qps = cpu_run(blob, n, steps, min(warmup, 1), threads)
    note = ""
    if args.config in N_ITEMS:
        note = f"; one item (frame / element 0) of the {N_ITEMS[args.config]}, every item costs the same"
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload,
            "run": {"samples_per_step": n, "parallelism": f"{threads} cpu threads, static split of the index range", "scene_source": src},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": f"{steps} x {n} samples; {what}{note}"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "queries_per_sample": qps}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--samples", type=int, default=0, help="samples per step per GPU (c1-c3; in total with --scaling strong) / per item (c4, c5)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--mesh", default="procedural", help="c2: 'procedural' or a .ply file to use as the target mesh")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-samples", type=int, default=0, help="samples per step of the CPU arm (0 = calibrate to ~10 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the small-call and CPU-row measurements")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    steps, warmup = args.steps, max(args.warmup, 0)
    threads = os.cpu_count() or 1
    name = args.config
    S = args.samples or DEFAULT_SAMPLES[name]

    if args.impl == "reference":
        if rank != 0:
            return 0
        workload = workload_of(name, args.mesh, mesh_triangles(args.mesh))
        print(json.dumps(reference_arm(args, workload, threads)))
        return 0

    scene, workload, ids = build_config(name, args.mesh)
    # Everything below may print to stdout from native code (NCCL prints its version line there): keep fd 1 for the one JSON
    # line and send the rest to stderr.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    from beifong_b200 import multi
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    scene.commit(local_rank)
    ptr, nfloats, stream_ptr = scene.profile_device()
    stream = torch.cuda.ExternalStream(stream_ptr, device=torch.device("cuda", local_rank))
    itemised = name in N_ITEMS
    needs_reduce = world > 1 and not itemised
    if needs_reduce:
        multi.init_native_reduce(scene, rank, world)      # the library's own NCCL communicator; torch only ships the 128-byte id

    if itemised:
        mine = multi.shard_items(N_ITEMS[name], rank, world)
        items = batch_items(name, ids, mine)
        packed = scene.pack_batch(items)      # the bf_batch_item array the C ABI takes (host buffers), marshalled once
        n_local, n_global = S * len(mine), S * N_ITEMS[name]
        first = 0
    elif args.scaling == "strong":
        first, n_local = multi.shard_samples(S, rank, world)
        n_global = S
    else:
        first, n_local, n_global = rank * S, S, S * world

    def step(i, e2e=False):
        if itemised:
            return scene.receive_batch(packed, S, read=e2e)         # e2e: per-item poses in, item profiles out
        if e2e:
            scene.reupload()                                        # host -> device: flattened scene from pinned memory
        scene.receive(n_local, seed=i, first_sample=first, accumulate=False)
        if needs_reduce:
            scene.multi_reduce(0)
        if e2e:
            return scene.profile()                                  # device -> host: the profile

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
    value = n_global * steps / (ms * 1e-3)

    # ---- end to end through the C ABI with host buffers -------------------------------------------------
    step(2000, e2e=True)
    sync_all()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i, e2e=True)
    sync_all()
    e2e_s = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = n_global * steps / float(e2e_s.item())
    if itemised:
        h2d = sum(64 * len(it.get("rects") or []) + (300 if it.get("receiver") is not None else 0) for it in items)
        d2h = int(nfloats * 4 * len(items))
    else:
        h2d, d2h = int(scene.geometry_bytes()), int(nfloats * 4)

    # ---- N > 1: is the reduced profile the profile of the whole index range? --------------------------------
    reduce_check = None
    if needs_reduce:
        m = min(n_local, 1 << 24)
        f0 = rank * m
        scene.receive(m, seed=77, first_sample=f0, accumulate=False)
        scene.multi_reduce(0)
        sync_all()
        if rank == 0:
            reduced = scene.profile().astype(np.float64)
            scene.receive(m * world, seed=77, first_sample=0, accumulate=False)
            whole = scene.profile().astype(np.float64)
            fin = np.isfinite(whole[..., 0]) & np.isfinite(reduced[..., 0])
            reduce_check = {"rel_l2_Y": float(np.linalg.norm(reduced[..., 0][fin] - whole[..., 0][fin]) / max(np.linalg.norm(whole[..., 0][fin]), 1e-300)),
                            # alpha / weight channels are fp32 counts: exact below 2^24 per bin, rounded above (C2 puts half of all samples
                            # -- the receiver rays that hit nothing -- into one bin), in the reference's fp32 ADC storage just the same
                            "counts_rel_max": float(np.max(np.abs(reduced[..., 1:3] - whole[..., 1:3]) / np.maximum(whole[..., 1:3], 1.0))),
                            "counts_equal": bool(np.array_equal(reduced[..., 1:3], whole[..., 1:3])), "samples": int(m * world),
                            "what": "rank 0's reduced profile of the sharded run vs one GPU running the whole index range", "bytes_reduced": int(nfloats * 4)}
        sync_all()

    out = None
    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
               "higher_is_better": True, "scaling": args.scaling if not itemised else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": workload,
               "run": {"samples_per_step": int(n_global), "samples_per_step_this_gpu": int(n_local),
                       "parallelism": (f"{N_ITEMS[name]} items dealt round-robin to {world} GPU(s), no reduce" if itemised else
                                       f"sample-index sharding x{world}, scene replicated" + (f", 1 ncclReduce of {int(nfloats * 4)} B per step (bf_multi_reduce)" if world > 1 else "")),
                       "e2e_excludes": "bf_scene_add_mesh + SAH build (reported as build_ms); the CPU arm excludes its BVH build too",
                       "build_ms": st1["build_ms"], "upload_ms": st1["upload_ms"]},
               "clocks": clk, "gpu_launches": int(launches),
               "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)}}
        if itemised:
            out["items_per_s"] = N_ITEMS[name] * steps / (ms * 1e-3)
            out["e2e"]["items_per_s"] = N_ITEMS[name] * steps / float(e2e_s.item())
        if reduce_check:
            out["reduce_check"] = reduce_check
        out["roofline"] = roofline_block(scene, workload, n_local if not itemised else S, first, value / world, itemised, items if itemised else None)
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            blob, src = scene_blob(name, args.mesh, scene)
            rate, _, _, _, _ = cpu_run(blob, 1 << 19, 1, 0, threads)
            cn = args.cpu_samples or int(max(1 << 15, min(1 << 31, rate * 15.0)))
            v, dt, kind, what, _ = cpu_run(blob, cn, 1, 0, threads)
            cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": f"1 x {cn} samples of the same workload ({dt:.1f} s); {what}"}
            if not args.no_extras and name == "c2":
                out["cpu_rows"] = cpu_rows(blob, cpu, threads)
        out["cpu_baseline"] = cpu
        if world == 1 and not args.no_extras and not itemised:
            out["small_calls"] = small_calls(scene)
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    scene.close()
    return 0


def roofline_block(scene, workload, n, first, per_gpu_value, itemised, items):
    """Roofline of the dominant kernel, timed live with CUDA events on the launching stream (bf_set_trace_timing)."""
    shape = scene.profile_shape()
    b_path, b_closest, b_shadow, b_bins = algorithmic_bytes(workload["triangles"], shape[2])
    scene.set_trace_timing(True)
    km, tot = {}, {"closest_queries": 0, "shadow_queries": 0, "path_closest_queries": 0, "path_shadow_queries": 0, "samples": 0}
    reps = 3
    for i in range(reps):
        a = scene.stats()
        if itemised:
            # items k_mega can run (no meshes) are ONE launch for the whole batch: time that launch; otherwise one item's waves
            scene.receive_batch(items if workload["triangles"] == 0 else items[:1], n, read=False)
        else:
            scene.receive(n, seed=3000 + i, first_sample=first)
        scene.synchronize()
        b = scene.stats()
        for k in tot:
            tot[k] += b[k] - a[k]
        for k, (m, c) in scene.last_kernel_ms().items():
            km.setdefault(k, [0.0, 0])
            km[k][0] += m; km[k][1] += c
    scene.set_trace_timing(False)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:   # noqa: BLE001
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    total_ms = sum(v[0] for v in km.values())
    dom = max(km, key=lambda k: km[k][0])
    dom_ms, dom_launch = km[dom]
    samples = max(tot["samples"], 1)
    if dom == "primary" and tot["path_closest_queries"]:
        kernel = ("k_path: receiver ray generation, rectangle / mesh-bounds tests, one shade step, the spawned rays' pre-tests and the range-bin "
                  "scatter for every path that stays clear of the meshes; spills the rest into the wavefront")
        alg_per_launch = (tot["path_closest_queries"] * b_closest + tot["path_shadow_queries"] * b_shadow + samples * b_bins) / max(dom_launch, 1)
        unit_note = "queries decided inside k_path x (B_path [+ 48]) + 8 C per sample"
    elif dom == "primary" and workload["triangles"] == 0:
        kernel = ("k_mega (k_mega_batch when the step is a bf_receive_batch): one thread per sample, receiver ray to range bin in one launch; "
                  "the scene has no meshes, so SURVEY 8(d)'s bytes per query are an upper bound that no kernel here moves")
        alg_per_launch = samples * b_closest / max(dom_launch, 1)
        unit_note = "one closest-hit query per sample x (B_path + 48)"
    elif dom == "primary":
        kernel = "k_primary: receiver ray generation + first closest-hit query + compaction"
        alg_per_launch = samples * b_closest / max(dom_launch, 1)
        unit_note = "one closest-hit query per sample x (B_path + 48)"
    else:
        q = {"trace_closest": tot["closest_queries"] - tot["path_closest_queries"], "trace_shadow": tot["shadow_queries"] - tot["path_shadow_queries"]}.get(dom, samples)
        per = {"trace_closest": b_closest, "trace_shadow": b_shadow}.get(dom, b_bins)
        kernel = {"trace_closest": "k_trace<closest>: 8-wide BVH traversal + Moller-Trumbore", "trace_shadow": "k_trace<any-hit>", "shade": "k_shade", "finalize": "k_finalize"}.get(dom, dom)
        alg_per_launch = q * per / max(dom_launch, 1)
        unit_note = "queries (or samples) of the launch x SURVEY 8(d) bytes per unit"
    achieved = alg_per_launch / (dom_ms / max(dom_launch, 1) * 1e-3) / 1e9
    alg_per_sample = (tot["closest_queries"] * b_closest + tot["shadow_queries"] * b_shadow) / samples + b_bins
    traffic, issue = None, None
    try:   # DRAM bytes and warp instructions per sample from the committed ncu --set full capture of the same kernel / step
        tr = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        if tr.get("config") == workload["name"]:
            if dom == "primary":
                traffic = tr["k_path_dram_bytes_per_sample"] * samples / max(dom_launch, 1)
            sm_clock = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
            ceiling = 148 * 4 * sm_clock / tr["step_warp_inst_per_sample"]
            issue = {"warp_inst_per_sample": tr["step_warp_inst_per_sample"], "k_path_warp_inst_per_sample": tr.get("k_path_warp_inst_per_sample"),
                     "ceiling_samples_per_s": ceiling, "frac": per_gpu_value / ceiling, "dram_bytes_per_sample_step": tr.get("step_dram_bytes_per_sample"),
                     "dram_frac_step": tr.get("step_dram_bytes_per_sample", 0) * per_gpu_value / 1e9 / peak,
                     "traversal_thread_inst_per_inst": tr.get("traversal_thread_inst_per_inst"), "source": tr.get("source")}
    except Exception:   # noqa: BLE001
        pass
    return {"bound": "hbm", "limiter": "issue slots (ncu: the scene is L2-resident, DRAM moves a few % of peak; see `issue`)" if workload["triangles"] < 5_000_000 else "hbm / L2 (the scene exceeds L2)",
            "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg_per_launch, "algorithmic_unit": unit_note, "avg_launch_ms": dom_ms / max(dom_launch, 1),
            "kernel_share_of_step": dom_ms / max(total_ms, 1e-9), "issue": issue,
            "kernel_ms_per_step": {k: v[0] / reps for k, v in km.items()}, "launches_per_step": {k: v[1] / reps for k, v in km.items()},
            "queries_per_sample": {"closest": tot["closest_queries"] / samples, "shadow": tot["shadow_queries"] / samples,
                                   "decided_in_k_path": (tot["path_closest_queries"] + tot["path_shadow_queries"]) / max(tot["closest_queries"] + tot["shadow_queries"], 1)},
            "algorithmic_bytes_per_sample": alg_per_sample, "whole_step_frac": alg_per_sample * per_gpu_value / 1e9 / peak}


def small_calls(scene):
    """Latency of small receive() calls (the reference's drivers use 1 000 - 65 536 samples per call): microseconds per call, pipelined
    (calls back to back, one synchronisation at the end) and with the profile read back after every call."""
    rows = []
    for n in (4096, 65536):
        for _ in range(20):
            scene.receive(n, seed=1)
        scene.synchronize()
        reps = 200
        t0 = time.perf_counter()
        for i in range(reps):
            scene.receive(n, seed=i)
        scene.synchronize()
        t_pipe = (time.perf_counter() - t0) / reps
        t0 = time.perf_counter()
        for i in range(reps):
            scene.receive(n, seed=i)
            scene.profile()
        t_read = (time.perf_counter() - t0) / reps
        rows.append({"samples_per_call": n, "us_per_call_pipelined": t_pipe * 1e6, "us_per_call_with_profile_read": t_read * 1e6,
                     "samples_per_s_pipelined": n / t_pipe, "samples_per_s_with_profile_read": n / t_read})
    return rows


def cpu_rows(blob, cpu_n, threads):
    """BASELINE.md 3's CPU rows: CPU-1 (one thread: what the reference's receive() does today, integrator.cpp:659-663), CPU-N (all
    cores) on this config, and CPU-BF (brute-force intersector, the kd-tree stand-in) on C1."""
    v1, dt1, kind, what, _ = cpu_run(blob, 1 << 18, 1, 0, 1)
    rows = {"CPU-1": {"value": v1, "unit": UNIT, "cores": 1, "kind": kind, "sample": f"1 x {1 << 18} samples ({dt1:.1f} s); {what}"},
            "CPU-N": cpu_n}
    try:
        b1, _ = scene_blob("c1", "procedural")
        vb, dtb, _, whatb, _ = cpu_run(b1, 1 << 20, 1, 0, threads, accel="brute")
        rows["CPU-BF"] = {"value": vb, "unit": UNIT, "cores": threads, "kind": "port", "config": "c1", "sample": f"1 x {1 << 20} samples of C1 ({dtb:.1f} s); {whatb}"}
    except Exception as e:   # noqa: BLE001
        rows["CPU-BF"] = {"unavailable": str(e)}
    return rows


if __name__ == "__main__":
    sys.exit(main())
