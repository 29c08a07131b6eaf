#!/usr/bin/env python
"""bench.py — LBVH build Mtri/s + traversal Mrays/s (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload soup1m|flake10m]

A step is one pass of the hot path over one batch: build_bvh over the scene's
triangles, then trace_rays over the batch of primary rays (and, for flake10m,
4 any-hit AO rays per pixel).  At N=1 the workload is BASELINE.json configs[1]
(1M-triangle soup, 1920x1080 primary rays).

  value          traversal Mrays/s, rays of all ranks / max-over-ranks device time
                 of the trace kernels, inputs resident in HBM, L2 flushed first
  build          the other half of the metric: Mtri/s of build_bvh, same rules
  e2e            the same two figures through the C ABI with HOST buffers
                 (pinned), H2D/D2H copies inside the timed region
  roofline       the dominant kernel (traversal) against measured HBM bandwidth,
                 plus one entry per build stage in `stages`
  cpu_baseline   the CPU oracle on this box's cores, bounded sample (N=1 only)

With --impl reference the CPU oracle is the thing timed (the reference itself
holds no code to run: SURVEY.md §0), on rank 0 only.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "LBVH build Mtri/s + traversal Mrays/s"
UNIT = "Mrays/s"

# algorithmic bytes per triangle of each build stage
STAGE_BYTES = {
    30: {"bounds": 48, "morton": 56, "sort": 68, "tree": 20 + 196, "total": 388},
    63: {"bounds": 48, "morton": 60, "sort": 200, "tree": 24 + 196, "total": 528},
}
# "tree" = A4 topology + A5 leaves + refit: one fused bottom-up pass (csrc/refit.cu), so the
# library's "topology" timer is empty and its time is reported with "refit".


def measured_peak():
    """HBM GB/s: MEASURED_PEAKS.json (driver-written) or the recipe's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def physical_gpu_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# ---------------------------------------------------------------------------
# CPU oracle leg (cpu_baseline of our arm; the whole of --impl reference)
# ---------------------------------------------------------------------------
def oracle_pass(wl, sample_rays: int, threads: int = 0):
    """One bounded pass of the workload on the host cores with the CPU oracle:
    full build of the scene, trace of a strided sample of the primary rays.
    Returns (build_s, trace_s, n_traced, nodes, tests, cores)."""
    import numpy as np
    import oracle as O  # test infrastructure: allowed here (cpu_baseline / reference arm)

    if threads:
        O.set_threads(threads)
    cores = O.max_threads()
    kind, p0, p1, seed = wl.scene
    tris = O.gen_scene(kind, p0, p1, seed)
    t0 = time.perf_counter()
    bvh = O.OracleBvh(tris, wl.key_bits)
    build_s = time.perf_counter() - t0
    total = wl.primary_rays
    stride = max(1, total // max(sample_rays, 1))
    rows = []
    # every stride-th ray of the image, generated by index like the GPU does
    cam = wl.camera()
    rays = O.gen_primary_rays(cam, wl.width, wl.height)[::stride]
    rays = np.ascontiguousarray(rays)
    t0 = time.perf_counter()
    hits, (nodes, tests) = bvh.trace(rays, counters=True)
    trace_s = time.perf_counter() - t0
    del rows
    bvh.close()
    return build_s, trace_s, int(rays.shape[0]), nodes, tests, cores


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the host cores.
    The mounted reference has no code, so this is the oracle port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from gpu_raytracer_b200.configs import WORKLOADS
    wl = WORKLOADS[args.workload]
    sample = 1 << 19
    for _ in range(min(args.warmup, 1)):
        oracle_pass(wl, sample)
    bs = ts = 0.0
    nr = 0
    steps = max(1, min(args.steps, 20))
    for _ in range(steps):
        b, t, n, nodes, tests, cores = oracle_pass(wl, sample)
        bs, ts, nr = bs + b, ts + t, nr + n
    mrays = nr / ts / 1e6
    mtri = wl.n_tris * steps / bs / 1e6
    sample_desc = (f"full CPU build of the {wl.n_tris}-triangle scene + every "
                   f"{max(1, wl.primary_rays // sample)}th of the {wl.primary_rays} primary rays "
                   f"({nr // steps} rays) per step")
    line = {
        "impl": "reference", "metric": METRIC, "value": mrays, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": min(args.warmup, 1), "ms_per_step": (bs + ts) / steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": wl.name, "n_tris": wl.n_tris, "rays_per_step": nr // steps,
                   "note": "CPU oracle port (reference has no code), OpenMP on all host threads"},
        "build": {"value": mtri, "unit": "Mtri/s", "ms": bs / steps * 1e3},
        "cpu_baseline": {"value": mrays, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample_desc, "build_mtri_s": mtri},
        "e2e": {"value": mrays, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "build_mtri_s": mtri},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import gpu_raytracer_b200 as g
    from gpu_raytracer_b200.configs import WORKLOADS
    from gpu_raytracer_b200 import sharding

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: gpu_raytracer_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(vals):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    wl = WORKLOADS[args.workload]
    K, W = args.steps, args.warmup
    kind, p0, p1, seed = wl.scene
    tris = g.gen_scene(kind, p0, p1, seed)                      # device resident, [n, 12]
    n_tris = int(tris.shape[0])
    # weak scaling: the frame grows to height x world rows of the same frustum and is dealt
    # to the ranks row by row (camera.interlaced_field), so every rank traces width x height
    # rays spread over the whole frustum: equal work per rank.  Contiguous
    # bands gave the ranks that see the middle of the sphereflake 1.4x the work of rank 0.
    cam = g.camera.interlaced_field(wl.camera(wl.width, wl.height), rank, world)
    n_rays_rank = wl.width * wl.height
    count = n_rays_rank
    rays = g.gen_primary_rays(cam, wl.width, wl.height)
    hits = torch.empty((count, 4), dtype=torch.float32, device="cuda")
    ao_rays = ao_hits = None
    if wl.ao_spp:
        ao_rays = torch.empty((count * wl.ao_spp, 8), dtype=torch.float32, device="cuda")
        ao_hits = torch.empty((count * wl.ao_spp, 4), dtype=torch.float32, device="cuda")
    rays_per_step_rank = count * (1 + wl.ao_spp)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB of L2

    def ev():
        return torch.cuda.Event(enable_timing=True)

    # ---- BVH strategy at N > 1: redundant build vs build on rank 0 + broadcast ----
    strategy = {"chosen": "redundant"}
    if world > 1:
        def timed_build(kind):
            barrier()
            t0 = time.perf_counter()
            if kind == "redundant":
                b = g.Bvh.build(tris, key_bits=wl.key_bits)
            else:
                b = g.Bvh.build(tris, key_bits=wl.key_bits) if rank == 0 else None
                b = sharding.broadcast_bvh(b, 0)
            barrier()
            dt = time.perf_counter() - t0
            b.close()
            return dt
        best = {}
        for kind in ("redundant", "broadcast"):
            for _ in range(2):
                timed_build(kind)          # warm the pools and NCCL
            best[kind] = min(timed_build(kind) for _ in range(3))
        t_red, t_bc = max_over_ranks([best["redundant"], best["broadcast"]])
        strategy = {"redundant_ms": t_red * 1e3, "broadcast_ms": t_bc * 1e3,
                    "chosen": "redundant" if t_red <= t_bc else "broadcast"}

    radius = None

    def step(timed):
        """build_bvh, then trace_rays; returns (build_ms, trace_ms) event pairs."""
        nonlocal radius
        e = [ev() for _ in range(6)]
        flush.zero_()
        e[0].record()
        if strategy["chosen"] == "broadcast":
            bvh = g.Bvh.build(tris, key_bits=wl.key_bits) if rank == 0 else None
            bvh = sharding.broadcast_bvh(bvh, 0)
        else:
            bvh = g.Bvh.build(tris, key_bits=wl.key_bits)
        e[1].record()
        flush.zero_()
        e[2].record()
        bvh.trace_async(rays, hits)
        e[3].record()
        if wl.ao_spp:
            if radius is None:
                box = bvh.bounds
                radius = wl.ao_radius_frac * float(np.linalg.norm(box[3:] - box[:3]))
            g.gen_ao_rays(tris, rays, hits, wl.ao_spp, wl.ao_seed, radius, out=ao_rays)
            flush.zero_()
            e[4].record()
            bvh.trace_async(ao_rays, ao_hits, g.RT_TRACE_ANY)
            e[5].record()
        bvh.close()
        return e

    for _ in range(W):
        step(False)
    barrier()
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    launches0 = g.kernel_launches()
    t0 = time.perf_counter()
    events = [step(True) for _ in range(K)]
    barrier()
    wall = time.perf_counter() - t0
    launches = g.kernel_launches() - launches0
    clocks = sampler.finish()
    build_ms = sum(e[0].elapsed_time(e[1]) for e in events)
    trace_ms = sum(e[2].elapsed_time(e[3]) for e in events)
    ao_ms = sum(e[4].elapsed_time(e[5]) for e in events) if wl.ao_spp else 0.0

    # ---- algorithmic bytes of the traversal: counted by the instrumented kernel ----
    with g.Bvh.build(tris, key_bits=wl.key_bits, timings=True) as bvh:
        stage_ms = bvh.build_times()
        _, nodes, tests = bvh.trace_stats(rays)
        hit_frac = float((g.hit_prims(hits) >= 0).float().mean())
        if wl.ao_spp:
            _, an, at = bvh.trace_stats(ao_rays, g.RT_TRACE_ANY)
            nodes_ao, tests_ao = an, at
    trace_bytes = count * 48 + 64 * nodes + 48 * tests

    # ---- e2e: host (pinned) buffers through the C ABI, copies inside the timing ----
    if args.staged:
        g.set_param("trace.streamed", 0)
    h_tris = tris.cpu().pin_memory()
    h_rays = rays.cpu().pin_memory()
    h_hits = torch.empty((count, 4), dtype=torch.float32).pin_memory()
    Ke = max(3, min(K, 10))
    eb = et = 0.0
    for it in range(2 + Ke):
        barrier()
        t0 = time.perf_counter()
        bvh = g.Bvh.build(h_tris, key_bits=wl.key_bits)
        t1 = time.perf_counter()
        bvh.trace(h_rays, out=h_hits)
        t2 = time.perf_counter()
        bvh.close()
        if it >= 2:
            eb, et = eb + (t1 - t0), et + (t2 - t1)
    e2e_ok = bool(torch.equal(h_hits.view(torch.int32), hits.cpu().view(torch.int32)))  # bit compare: miss prim = -1 is a NaN pattern

    build_ms, trace_ms, ao_ms, wall, eb, et = max_over_ranks([build_ms, trace_ms, ao_ms, wall, eb, et])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    total_trace_ms = trace_ms + ao_ms
    value = rays_per_step_rank * world * K / total_trace_ms / 1e3
    build_mtri = n_tris * K / build_ms / 1e3
    trace_gbs = trace_bytes / (trace_ms / K) / 1e6
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(wl.name, {}).get("trace_kernel_dram_bytes")
    except Exception:
        pass
    sb = STAGE_BYTES[wl.key_bits]
    stage_ms = dict(stage_ms)
    stage_ms["tree"] = stage_ms.get("topology", 0.0) + stage_ms.get("refit", 0.0)
    stages = {k: {"ms": stage_ms[k], "alg_bytes_per_tri": sb[k],
                  "achieved_gbs": sb[k] * n_tris / stage_ms[k] / 1e6,
                  "frac": sb[k] * n_tris / stage_ms[k] / 1e6 / peak}
              for k in ("bounds", "morton", "sort", "tree", "total") if stage_ms[k] > 0}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": (build_ms + total_trace_ms) / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": wl.name, "n_tris": n_tris, "key_bits": wl.key_bits,
            "rays_per_step_per_gpu": rays_per_step_rank, "image": [wl.width, wl.height * world],
            "ao_spp": wl.ao_spp, "hit_fraction": hit_frac,
            "sharding": "rows of the frame dealt to the ranks one by one (interlaced fields), no collective",
            "l2": "256 MiB memset before build and before each trace of every step (outside the events)",
            "bvh_strategy": strategy,
            "value_is": "rays / device time of the trace kernels; `build` holds the Mtri/s half of the metric",
        },
        "build": {"value": build_mtri, "unit": "Mtri/s", "ms": build_ms / K,
                  "target_ms_10m": 10.0},
        "trace": {"ms": total_trace_ms / K, "primary_ms": trace_ms / K, "ao_ms": ao_ms / K},
        "wall_ms_per_step": wall / K * 1e3,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "e2e": {"value": count * world * Ke / et / 1e6, "unit": UNIT,
                "h2d_bytes_per_step": int(n_tris * 48 + count * 32),
                "d2h_bytes_per_step": int(count * 16 + 64),
                "build_mtri_s": n_tris * Ke / eb / 1e6, "build_ms": eb / Ke * 1e3,
                "trace_ms": et / Ke * 1e3, "steps": Ke, "hits_equal_device_path": e2e_ok,
                "note": "pinned host tris/rays/hits through build_bvh / trace_rays; primary rays only"},
        "roofline": {"bound": "hbm", "kernel": "trace_kernel<closest>", "achieved": trace_gbs,
                     "peak": peak, "unit": "GB/s", "frac": trace_gbs / peak, "traffic": traffic,
                     "peak_source": peak_src,
                     "alg_bytes_per_launch": int(trace_bytes),
                     "per_ray": {"nodes_fetched": nodes / count, "tris_tested": tests / count,
                                 "bytes": trace_bytes / count},
                     "note": "latency/divergence-bound kernel; HBM is the stated denominator (SURVEY.md §8d)",
                     "stages": stages},
    }
    if world == 1 and not args.no_cpu:
        b_s, t_s, n, cn, ct, cores = oracle_pass(wl, 1 << 20)
        line["cpu_baseline"] = {
            "value": n / t_s / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"full CPU build ({wl.n_tris} tris, {b_s:.2f} s) + {n} of the "
                      f"{wl.primary_rays} primary rays ({t_s:.2f} s)",
            "build_mtri_s": wl.n_tris / b_s / 1e6,
            "per_ray": {"nodes_fetched": cn / n, "tris_tested": ct / n}}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="soup1m", choices=["soup1m", "flake10m", "cornell"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--staged", action="store_true",
                    help="e2e leg through staged chunks instead of the streamed single launch "
                         "(for runs under a replaying profiler)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
