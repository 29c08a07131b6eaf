#!/usr/bin/env python
"""bench.py — LBVH build Mtri/s + traversal Mrays/s (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload flake10m|soup1m|cornell|random1b] [--no-cpu] [--no-configs]

A step is one pass of the hot path over one batch: build_bvh over the scene's
triangles, then trace_rays over the batch of primary rays and, for flake10m, 4
any-hit AO rays per pixel.  The default workload is BASELINE.json configs[2]
(`flake10m`: 10M-triangle sphereflake, 3840x2160 primary + 4 AO rays/px), the
configuration both north-star targets are quoted on (10M build <= 10 ms with
the sort fraction; primary+AO Mrays/s).  `soup1m` (configs[1]) stays selectable.

  value          traversal Mrays/s, rays of all ranks / max-over-ranks device time
                 of the trace kernels, inputs resident in HBM, L2 flushed first
  build          the other half of the metric: Mtri/s of build_bvh, same rules,
                 with per-stage ms and roofline fractions under roofline.stages
  e2e            the same figures through the C ABI with HOST (pinned) inputs and
                 outputs, every copy inside the timed region
  roofline       the dominant kernel (the traversal that takes most of the step)
                 against measured HBM bandwidth: algorithmic bytes, DRAM bytes
                 (ncu), L1 sectors (ncu)
  configs        sub-records: config 4 (100M terrain, 63- and 30-bit builds; N=1),
                 config 5 (`random1b`: ONE batch of 10^9 random rays cut N ways —
                 strong scaling — plain and with RT_TRACE_REORDER; at N=1 also the
                 work counters of a 2^24-ray sample, which make its roofline a
                 genuine HBM fraction: the 1.1 GB BVH does not fit in L2)
  cpu_baseline   the CPU oracle on this box's cores, same workload (N=1 only)

`scaling` names the headline `value`: weak (every rank traces a full frame of the
same frustum, rows interlaced).  configs.random1b is the strong-scaling record.

With --impl reference the CPU oracle is the thing timed (the reference itself
holds no code to run: SURVEY.md §0), on rank 0 only, on all host threads.
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

# algorithmic bytes per triangle of each build stage.  "encode" = A1 scene
# bounds + A2 Morton codes, reported as ONE stage: the bounds pass also stores the centroids
# the Morton pass reads, so the split of the bytes between the two kernels is not the split
# of the table (48 + 56/60).  "tree" = A4 topology + A5 leaves + refit, one fused bottom-up
# pass (csrc/refit.cu): the library's "topology" timer is empty and counted with "refit".
STAGE_BYTES = {
    30: {"encode": 48 + 56, "sort": 68, "tree": 20 + 196, "total": 388},
    63: {"encode": 48 + 60, "sort": 200, "tree": 24 + 196, "total": 528},
}
RANDOM_BATCH = 1 << 25


def measured_peak():
    """HBM GB/s: MEASURED_PEAKS.json (driver-written) or the recipe's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_record(name):
    """Per-launch ncu counters of a trace kernel, written by tools/summarize_ncu.py."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(name, {})
    except Exception:
        return {}


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
def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class OracleWorkload:
    """The workload's inputs on the host, made once; `step()` is one timed pass of the CPU
    oracle over ALL of them: full build, every primary ray, every AO ray (or, for random1b,
    a stated slice of the 10^9 rays)."""

    def __init__(self, wl, threads: int, random_slice: int = 1 << 22):
        import numpy as np
        import oracle as O  # test infrastructure: allowed here (cpu_baseline / reference arm)

        self.O, self.np, self.wl = O, np, wl
        # torch.distributed.run exports OMP_NUM_THREADS=1 when nproc > 1: set the count
        # explicitly so the CPU arm always uses every core the process may run on
        O.set_threads(threads)
        self.cores = O.max_threads()
        kind, p0, p1, seed = wl.scene
        self.tris = O.gen_scene(kind, p0, p1, seed)
        self.random_slice = random_slice
        self.rays = None
        if wl.primary_rays:
            self.rays = O.gen_primary_rays(wl.camera(), wl.width, wl.height)

    def step(self):
        """-> dict(build_s, trace_s, rays, nodes, tests)"""
        O, np, wl = self.O, self.np, self.wl
        t0 = time.perf_counter()
        bvh = O.OracleBvh(self.tris, wl.key_bits)
        build_s = time.perf_counter() - t0
        out = {"build_s": build_s, "trace_s": 0.0, "rays": 0, "nodes": 0, "tests": 0, "primary_s": 0.0}
        if wl.primary_rays:
            t0 = time.perf_counter()
            hits, (nodes, tests) = bvh.trace(self.rays, counters=True)
            dt = time.perf_counter() - t0
            out.update(trace_s=dt, primary_s=dt, rays=int(self.rays.shape[0]), nodes=nodes, tests=tests)
            if wl.ao_spp:
                box = bvh.bounds
                radius = wl.ao_radius_frac * float(np.linalg.norm(box[3:] - box[:3]))
                ao = O.gen_ao_rays(self.tris, self.rays, hits, wl.ao_spp, wl.ao_seed, radius)  # untimed, like ours
                t0 = time.perf_counter()
                bvh.trace(ao, O.TRACE_ANY)
                out["trace_s"] += time.perf_counter() - t0
                out["rays"] += int(ao.shape[0])
                del ao
        elif wl.random_rays:
            rays = O.gen_random_rays(bvh.bounds, wl.random_seed, 0, self.random_slice)
            t0 = time.perf_counter()
            _, (nodes, tests) = bvh.trace(rays, counters=True)
            out.update(trace_s=time.perf_counter() - t0, rays=self.random_slice, nodes=nodes, tests=tests)
        bvh.close()
        return out

    def sample_text(self):
        wl = self.wl
        if wl.primary_rays:
            return (f"the whole workload per step: CPU build of the {wl.n_tris}-triangle scene, all "
                    f"{wl.primary_rays} primary rays" +
                    (f" and all {wl.primary_rays * wl.ao_spp} AO rays (any-hit)" if wl.ao_spp else ""))
        return (f"CPU build of the {wl.n_tris}-triangle scene + the first {self.random_slice} of the "
                f"{wl.random_rays} random rays per step (a rate: the rays are i.i.d.)")


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the host cores.
    The mounted reference has no code, so this is the oracle port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from gpu_raytracer_b200.configs import WORKLOADS
    wl = WORKLOADS[args.workload]
    ow = OracleWorkload(wl, host_threads())
    K, W = max(1, args.steps), max(0, args.warmup)
    budget_s = float(os.environ.get("RT_REFERENCE_BUDGET_S", "420"))
    t_begin = time.perf_counter()
    warmed = 0
    while warmed < W:
        r = ow.step()
        warmed += 1
        # a slow box: keep the whole run within the budget by warming up less, never by
        # timing fewer steps than asked
        if (time.perf_counter() - t_begin) + (K + 1) * (r["build_s"] + r["trace_s"]) * 1.3 > budget_s:
            break
    acc = {"build_s": 0.0, "trace_s": 0.0, "rays": 0, "nodes": 0, "tests": 0, "primary_s": 0.0}
    for _ in range(K):
        r = ow.step()
        for k in acc:
            acc[k] += r[k]
    mrays = acc["rays"] / acc["trace_s"] / 1e6
    mtri = wl.n_tris * K / acc["build_s"] / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": mrays, "unit": UNIT, "n_gpus": args.gpus,
        "steps": K, "warmup": warmed, "ms_per_step": (acc["build_s"] + acc["trace_s"]) / K * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": wl.name, "n_tris": wl.n_tris, "key_bits": wl.key_bits,
                   "rays_per_step": acc["rays"] // K, "ao_spp": wl.ao_spp,
                   "note": "CPU oracle port (reference has no code), OpenMP on all host threads; one "
                           "frame per step whatever --gpus says (a rate)"},
        "build": {"value": mtri, "unit": "Mtri/s", "ms": acc["build_s"] / K * 1e3},
        "trace": {"ms": acc["trace_s"] / K * 1e3, "primary_ms": acc["primary_s"] / K * 1e3},
        "cpu_baseline": {"value": mrays, "unit": UNIT, "cores": ow.cores, "kind": "port",
                         "sample": ow.sample_text(), "build_mtri_s": mtri},
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

    def sum_over_ranks(vals):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.tolist()

    def ev():
        return torch.cuda.Event(enable_timing=True)

    peak, peak_src = measured_peak()
    K, W = args.steps, args.warmup
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB of L2

    # -----------------------------------------------------------------------
    # config 5: ONE batch of `total` random rays cut `world` ways (strong scaling)
    # -----------------------------------------------------------------------
    def random_batch(bvh, total, reorder, passes=1, warm=True):
        """This rank's share of rays [0, total) in batches; returns (trace_ms, hits) of the
        timed passes.  Rays are generated on the device per batch, outside the events."""
        wl5 = WORKLOADS["random1b"]
        box = bvh.bounds
        first, count = sharding.shard_range(total, rank, world)
        flags = g.RT_TRACE_REORDER if reorder else 0
        cap = min(RANDOM_BATCH, max(count, 1))
        rays = torch.empty((cap, 8), dtype=torch.float32, device="cuda")
        hits = torch.empty((cap, 4), dtype=torch.float32, device="cuda")
        e0, e1 = ev(), ev()
        if warm and count:
            c = min(cap, count)
            g.gen_random_rays(box, wl5.random_seed, first, c, out=rays)
            bvh.trace_async(rays[:c], hits[:c], flags)
        barrier()
        ms, nhit = 0.0, 0
        for _ in range(passes):
            done = 0
            while done < count:
                c = min(cap, count - done)
                g.gen_random_rays(box, wl5.random_seed, first + done, c, out=rays)
                e0.record()
                bvh.trace_async(rays[:c], hits[:c], flags)
                e1.record()
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1)
                nhit += int((g.hit_prims(hits[:c]) >= 0).sum())
                done += c
        return ms / passes, nhit // passes

    def attach_ncu(leg, key):
        """ncu's view of a config-5 launch (profiles/traffic.json), rescaled to this run's rate"""
        nc = ncu_record(key)
        r = leg["roofline"]
        if nc.get("dram_bytes_per_ray"):
            r["dram_bytes_per_ray_ncu"] = nc["dram_bytes_per_ray"]
            r["dram_frac"] = nc["dram_bytes_per_ray"] * leg["mrays_s"] / 1e3 / peak
        if nc.get("ncu_us") and nc.get("rays"):
            scale = (nc["ncu_us"] / nc["rays"]) * leg["mrays_s"]
            if nc.get("l1_lsu_wavefront_pct"):
                r["l1_frac"] = nc["l1_lsu_wavefront_pct"] / 100.0 * scale
            if nc.get("issue_pct"):
                r["issue_frac"] = nc["issue_pct"] / 100.0 * scale

    def config5_record(bvh, total, passes):
        rec = {"rays": total, "n_tris": bvh.n_tris, "scaling": "strong",
               "sharding": "contiguous ranges of ONE batch (sharding.shard_range), rays generated per "
                           "shard on device from (seed, index), no collective",
               "batch_rays": RANDOM_BATCH, "passes": passes}
        for name, reorder in (("plain", False), ("reorder", True)):
            ms, nhit = random_batch(bvh, total, reorder, passes)
            (ms_max,) = max_over_ranks([ms])
            (hit_sum,) = sum_over_ranks([float(nhit)])
            rec[name] = {"mrays_s": total / ms_max / 1e3, "trace_ms": ms_max,
                         "hit_fraction": hit_sum / total}
        if world == 1:
            # work counters on a 2^24-ray sample: the §8(d) bytes per ray.  The 1.1 GB BVH is
            # far larger than L2, so achieved / peak is a genuine HBM fraction here.
            wl5 = WORKLOADS["random1b"]
            ns = min(1 << 24, total)
            rays = g.gen_random_rays(bvh.bounds, wl5.random_seed, 0, ns)
            _, nodes, tests = bvh.trace_stats(rays)
            bpr = 48 + 64 * nodes / ns + 48 * tests / ns
            rec["per_ray"] = {"nodes_fetched": nodes / ns, "tris_tested": tests / ns, "bytes": bpr,
                              "sample_rays": ns}
            for name in ("plain", "reorder"):
                gbs = bpr * rec[name]["mrays_s"] / 1e3
                rec[name]["roofline"] = {"achieved_gbs": gbs, "frac": gbs / peak}
            for name, key in (("plain", "random1b"), ("reorder", "random1b:reorder")):
                attach_ncu(rec[name], key)
        # the same batch through the 32-byte nodes (RT_BUILD_COMPACT): packed here, outside the timing
        bvh.compact()
        for name, reorder in (("compact", False), ("compact_reorder", True)):
            ms, nhit = random_batch(bvh, total, reorder, passes)
            (ms_max,) = max_over_ranks([ms])
            (hit_sum,) = sum_over_ranks([float(nhit)])
            rec[name] = {"mrays_s": total / ms_max / 1e3, "trace_ms": ms_max, "hit_fraction": hit_sum / total}
        if world == 1:
            _, nodes_c, tests_c = bvh.trace_stats(rays)
            bpr_c = 48 + 32 * nodes_c / ns + 48 * tests_c / ns
            rec["per_ray_compact"] = {"nodes_fetched": nodes_c / ns, "tris_tested": tests_c / ns, "bytes": bpr_c}
            for name, key in (("compact", "random1b:compact"), ("compact_reorder", "random1b:reorder:compact")):
                gbs = bpr_c * rec[name]["mrays_s"] / 1e3
                rec[name]["roofline"] = {"achieved_gbs": gbs, "frac": gbs / peak}
                attach_ncu(rec[name], key)
        return rec

    wl = WORKLOADS[args.workload]
    kind, p0, p1, seed = wl.scene
    tris = g.gen_scene(kind, p0, p1, seed)                      # device resident, [n, 12]
    n_tris = int(tris.shape[0])

    if wl.random_rays:   # --workload random1b: config 5 as the headline (strong scaling)
        for _ in range(max(W, 3)):
            g.Bvh.build(tris, key_bits=wl.key_bits).close()
        barrier()
        sampler = ClockSampler(physical_gpu_index(local_rank))
        sampler.start()
        launches0 = g.kernel_launches()
        e0, e1 = ev(), ev()
        e0.record()
        bvh = g.Bvh.build(tris, key_bits=wl.key_bits)
        e1.record()
        torch.cuda.synchronize()
        (build_ms,) = max_over_ranks([e0.elapsed_time(e1)])
        rec = config5_record(bvh, wl.random_rays, max(1, min(K, 3)))
        launches = g.kernel_launches() - launches0
        clocks = sampler.finish()
        bvh.close()
        if rank == 0:
            best = max(("plain", "reorder", "compact", "compact_reorder"), key=lambda k: rec[k]["mrays_s"])
            line = {"metric": METRIC, "value": rec[best]["mrays_s"], "unit": UNIT, "n_gpus": world,
                    "steps": rec["passes"], "warmup": max(W, 3), "ms_per_step": rec[best]["trace_ms"],
                    "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                    "data": "synthetic",
                    "config": {"workload": wl.name, "n_tris": n_tris, "rays": wl.random_rays,
                               "variant": best, "l2": "working set 1.1 GB > 126 MB of L2"},
                    "build": {"value": n_tris / build_ms / 1e3, "unit": "Mtri/s", "ms": build_ms},
                    "configs": {"random1b": rec}, "gpu_launches": int(launches), "clocks": clocks,
                    "e2e": {"value": rec[best]["mrays_s"], "unit": UNIT, "h2d_bytes_per_step": 96,
                            "d2h_bytes_per_step": 8,
                            "note": "rays are generated on the device from (seed, index): 10^9 rays = 32 GB "
                                    "never cross PCIe (SURVEY.md §8a A8); only the box and the hit count do"}}
            if "roofline" in rec[best]:
                r = rec[best]["roofline"]
                line["roofline"] = {"bound": "hbm", "kernel": "trace_kernel_dyn<closest>", "achieved": r["achieved_gbs"],
                                    "peak": peak, "unit": "GB/s", "frac": r["frac"], "traffic": None,
                                    "peak_source": peak_src, "per_ray": rec["per_ray"]}
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    # weak scaling: the frame grows to height x world rows of the same frustum and is dealt
    # to the ranks row by row (camera.interlaced_field), so every rank traces width x height
    # rays spread over the whole frustum: equal work per rank.  Contiguous
    # bands gave the ranks that see the middle of the sphereflake 1.4x the work of rank 0.
    cam = g.camera.interlaced_field(wl.camera(wl.width, wl.height), rank, world)
    count = wl.width * wl.height
    rays = g.gen_primary_rays(cam, wl.width, wl.height)
    hits = torch.empty((count, 4), dtype=torch.float32, device="cuda")
    ao_rays = ao_hits = None
    if wl.ao_spp:
        ao_rays = torch.empty((count * wl.ao_spp, 8), dtype=torch.float32, device="cuda")
        ao_hits = torch.empty((count * wl.ao_spp, 4), dtype=torch.float32, device="cuda")
    rays_per_step_rank = count * (1 + wl.ao_spp)

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

    def step():
        """build_bvh, then trace_rays; returns the event pairs around each part."""
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
        step()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    launches0 = g.kernel_launches()
    t0 = time.perf_counter()
    events = [step() for _ in range(K)]
    barrier()
    wall = time.perf_counter() - t0
    launches = g.kernel_launches() - launches0
    clocks = sampler.finish()
    build_ms = sum(e[0].elapsed_time(e[1]) for e in events)
    trace_ms = sum(e[2].elapsed_time(e[3]) for e in events)
    ao_ms = sum(e[4].elapsed_time(e[5]) for e in events) if wl.ao_spp else 0.0

    # ---- per-stage build times (median of 5 timed builds) and the traversal's work counters ----
    stage_runs = []
    for _ in range(5):
        with g.Bvh.build(tris, key_bits=wl.key_bits, timings=True) as b:
            stage_runs.append(b.build_times())
    stage_ms = {k: sorted(r[k] for r in stage_runs)[len(stage_runs) // 2] for k in stage_runs[0]}
    nodes_ao = tests_ao = 0
    with g.Bvh.build(tris, key_bits=wl.key_bits) as bvh:
        _, nodes, tests = bvh.trace_stats(rays)
        hit_frac = float((g.hit_prims(hits) >= 0).float().mean())
        if wl.ao_spp:
            _, nodes_ao, tests_ao = bvh.trace_stats(ao_rays, g.RT_TRACE_ANY)
    n_ao = count * wl.ao_spp
    bytes_primary = count * 48 + 64 * nodes + 48 * tests
    bytes_ao = n_ao * 48 + 64 * nodes_ao + 48 * tests_ao

    # ---- e2e: host (pinned) inputs and outputs through the C ABI, copies inside the timing ----
    if args.staged:
        g.set_param("trace.streamed", 0)
    h_tris = tris.cpu().pin_memory()
    h_rays = rays.cpu().pin_memory()
    h_hits = torch.empty((count, 4), dtype=torch.float32).pin_memory()
    h_rgb = torch.empty((count, 3), dtype=torch.float32).pin_memory()
    Ke = max(3, min(K, 10))
    s_down = torch.cuda.Stream()
    ev_hits = torch.cuda.Event()
    eb = et = ec = es = eca = 0.0
    d_rgb = None
    one_call_rgb = None
    for it in range(2 + Ke):
        barrier()
        t0 = time.perf_counter()
        if wl.ao_spp:
            # the AO generator and the shader index the ORIGINAL triangles: the host copy goes up
            # once and build_bvh uses it in place
            # N > 1: every rank holds the same host array; each uploads 1/N of it and one NCCL
            # all-gather over NVLink hands everybody the rest (sharding.upload_scene), so the
            # scene crosses the host's PCIe once instead of N times
            d_tris = sharding.upload_scene(h_tris, rank, world) if world > 1 else h_tris.cuda(non_blocking=True)
            bvh = g.Bvh.build(d_tris, key_bits=wl.key_bits)
        else:
            bvh = g.Bvh.build(sharding.upload_scene(h_tris, rank, world) if world > 1 else h_tris, key_bits=wl.key_bits)
        t1 = time.perf_counter()
        if wl.ao_spp:
            # one call: host rays up, primary hits + the shaded frame down, in 1M-ray bands over
            # three streams; the AO rays are made, traced and consumed on the device
            bvh.render_ao(d_tris, h_rays, wl.ao_spp, wl.ao_seed, radius, hits=h_hits, rgb=h_rgb)
        else:
            bvh.trace(h_rays, out=h_hits)
        t2 = time.perf_counter()
        # the same frame from the 12-float camera: no ray upload (trace_camera)
        bvh.trace_camera(cam, wl.width, wl.height, out=h_hits)
        t3 = time.perf_counter()
        if wl.ao_spp:
            # the same frame through the four separate calls (whole-frame upload, then compute)
            if it == 0:
                one_call_rgb = h_rgb.clone()
            d_rays = h_rays.cuda(non_blocking=True)
            bvh.trace_async(d_rays, hits)
            ev_hits.record()
            with torch.cuda.stream(s_down):                # the hits go down while the AO pass runs
                s_down.wait_event(ev_hits)
                h_hits.copy_(hits, non_blocking=True)
            g.gen_ao_rays(d_tris, d_rays, hits, wl.ao_spp, wl.ao_seed, radius, out=ao_rays)
            bvh.trace_async(ao_rays, ao_hits, g.RT_TRACE_ANY)
            d_rgb = g.shade(d_tris, d_rays, hits, ao_hits, wl.ao_spp)
            h_rgb.copy_(d_rgb, non_blocking=True)
            torch.cuda.synchronize()
        t4 = time.perf_counter()
        if wl.ao_spp:
            # and from the camera alone: primary rays made on the device, hits + frame down
            bvh.render_camera_ao(d_tris, cam, wl.width, wl.height, wl.ao_spp, wl.ao_seed, radius, hits=h_hits, rgb=h_rgb)
        t5 = time.perf_counter()
        bvh.close()
        if it >= 2:
            eb, et, ec, es = eb + (t1 - t0), et + (t2 - t1), ec + (t3 - t2), es + (t4 - t3)
            eca += t5 - t4
    e2e_ok = bool(torch.equal(h_hits.view(torch.int32), hits.cpu().view(torch.int32)))  # bit compare: miss prim = -1 is a NaN pattern
    if one_call_rgb is not None:                                # rt_render_ao's frame == the separate calls' frame
        e2e_ok = e2e_ok and bool(torch.equal(one_call_rgb.view(torch.int32), h_rgb.view(torch.int32)))
    del h_tris, h_rays

    build_ms, trace_ms, ao_ms, wall, eb, et, ec, es, eca = max_over_ranks([build_ms, trace_ms, ao_ms, wall, eb, et, ec, es, eca])

    # ---- sub-records: config 5 at every N (strong scaling), config 4 at N = 1 ----
    configs = {}
    if not args.no_configs:
        if wl.scene == WORKLOADS["random1b"].scene:
            scene5 = tris
        else:
            scene5 = g.gen_scene(*WORKLOADS["random1b"].scene)
        with g.Bvh.build(scene5) as bvh5:
            configs["random1b"] = config5_record(bvh5, WORKLOADS["random1b"].random_rays, 1)
        del scene5
        if world == 1:
            del ao_rays, ao_hits
            ao_rays = ao_hits = None
            torch.cuda.empty_cache()
            wl4 = WORKLOADS["terrain100m"]
            t4 = g.gen_scene(*wl4.scene)
            rec4 = {"n_tris": int(t4.shape[0])}
            for kb in (63, 30):
                runs = []
                for i in range(2 + 5):
                    with g.Bvh.build(t4, key_bits=kb, timings=True) as b:
                        if i >= 2:
                            runs.append(b.build_times())
                med = {k: sorted(r[k] for r in runs)[len(runs) // 2] for k in runs[0]}
                n4, sb = rec4["n_tris"], STAGE_BYTES[kb]
                med["encode"] = med["bounds"] + med["morton"]
                med["tree"] = med["topology"] + med["refit"]
                rec4[f"k{kb}"] = {
                    "build_ms": med["total"], "mtri_s": n4 / med["total"] / 1e3,
                    "frac": sb["total"] * n4 / med["total"] / 1e6 / peak,
                    "stages": {k: {"ms": med[k], "frac": sb[k] * n4 / med[k] / 1e6 / peak}
                               for k in ("encode", "sort", "tree")}}
            configs["terrain100m"] = rec4
            del t4
            g.trim()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_trace_ms = trace_ms + ao_ms
    value = rays_per_step_rank * world * K / total_trace_ms / 1e3
    build_mtri = n_tris * K / build_ms / 1e3
    sb = STAGE_BYTES[wl.key_bits]
    stage_ms = dict(stage_ms)
    stage_ms["encode"] = stage_ms.get("bounds", 0.0) + stage_ms.get("morton", 0.0)
    stage_ms["tree"] = stage_ms.get("topology", 0.0) + stage_ms.get("refit", 0.0)
    stages = {k: {"ms": stage_ms[k], "alg_bytes_per_tri": sb[k],
                  "achieved_gbs": sb[k] * n_tris / stage_ms[k] / 1e6,
                  "frac": sb[k] * n_tris / stage_ms[k] / 1e6 / peak}
              for k in ("encode", "sort", "tree", "total") if stage_ms[k] > 0}

    # the dominant kernel = the traversal launch that takes most of the step
    def kernel_roofline(name, kernel, alg_bytes, ms, n, nd, nt):
        gbs = alg_bytes / ms / 1e6
        nc = ncu_record(name)
        r = {"kernel": kernel, "ms": ms, "achieved": gbs, "frac": gbs / peak,
             "alg_bytes_per_launch": int(alg_bytes),
             "per_ray": {"nodes_fetched": nd / n, "tris_tested": nt / n, "bytes": alg_bytes / n},
             "traffic": nc.get("dram_bytes"),
             # (i) what the launch really moved over HBM, as a fraction of peak
             "dram_frac": (nc["dram_bytes"] / ms / 1e6 / peak) if nc.get("dram_bytes") else None,
             # (ii) L1 roofline: the unit ncu shows busiest is the LSU data pipe (one wavefront per
             # clock and SM): its share of peak under ncu, rescaled to this run's duration
             "l1_sectors": nc.get("l1_sectors"),
             "l1_frac": (nc["l1_lsu_wavefront_pct"] / 100.0 * (nc["ncu_us"] / 1e3) / ms)
                        if nc.get("l1_lsu_wavefront_pct") and nc.get("ncu_us") else None,
             # (iii) the other limit the kernel sits at: issue slots in use (ncu, same rescaling)
             "issue_frac": (nc["issue_pct"] / 100.0 * (nc["ncu_us"] / 1e3) / ms)
                           if nc.get("issue_pct") and nc.get("ncu_us") else None}
        return r
    kr = {"primary": kernel_roofline(wl.name + ":primary", "trace_kernel_dyn<closest>", bytes_primary,
                                     trace_ms / K, count, nodes, tests)}
    if wl.ao_spp:
        kr["ao"] = kernel_roofline(wl.name + ":ao", "trace_kernel_dyn<any>", bytes_ao, ao_ms / K, n_ao,
                                   nodes_ao, tests_ao)
    dom = max(kr, key=lambda k: kr[k]["ms"])
    roof = {"bound": "hbm", "kernel": kr[dom]["kernel"], "achieved": kr[dom]["achieved"], "peak": peak,
            "unit": "GB/s", "frac": kr[dom]["frac"], "traffic": kr[dom]["traffic"],
            "peak_source": peak_src, "dominant": dom, "kernels": kr,
            "note": "achieved = algorithmic bytes (48 + 64 x nodes fetched + 48 x triangles tested per ray, "
                    "counted by trace_rays_stats) / CUDA-event time.  The BVH of a coherent frame is served "
                    "from L1/L2, so `frac` can exceed 1 and says nothing about distance to a bound; "
                    "`dram_frac` (ncu dram bytes / time / peak), `l1_frac` (ncu: LSU data-pipe wavefronts, share of the pipe's peak) "
                    "and `issue_frac` (ncu: issue slots in use) are the fractions of real limits, and configs.random1b.*.roofline is the genuine "
                    "HBM fraction of §8(d) (working set 1.1 GB >> L2)",
            "stages": stages}
    n_img = count
    e2e_rays = count * (1 + wl.ao_spp)
    # which of the two public paths a host with this many ranks should use
    if wl.ao_spp and world >= 4:
        e2e_best, e2e_path = es, "separate_calls"
    else:
        e2e_best, e2e_path = et, ("rt_render_ao" if wl.ao_spp else "trace_rays")
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
            "value_is": "rays (primary + AO) / device time of the trace kernels; `build` holds the Mtri/s "
                        "half of the metric; configs.random1b is the strong-scaling record",
        },
        "build": {"value": build_mtri, "unit": "Mtri/s", "ms": build_ms / K,
                  "target_ms_10m": 10.0, "sort_frac": stages.get("sort", {}).get("frac"),
                  "sort_target_frac": 0.70},
        "trace": {"ms": total_trace_ms / K, "primary_ms": trace_ms / K, "ao_ms": ao_ms / K,
                  "primary_mrays_s": count * world * K / trace_ms / 1e3,
                  "ao_mrays_s": (n_ao * world * K / ao_ms / 1e3) if wl.ao_spp else None},
        "wall_ms_per_step": wall / K * 1e3,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "e2e": {"value": e2e_rays * world * Ke / e2e_best / 1e6, "unit": UNIT,
                "h2d_bytes_per_step": int(n_tris * 48 + count * 32),
                "d2h_bytes_per_step": int(count * 16 + (n_img * 12 if wl.ao_spp else 0) + 64),
                "build_mtri_s": n_tris * Ke / eb / 1e6, "build_ms": eb / Ke * 1e3,
                "build_note": ("host triangles -> 1/N per rank over PCIe + NCCL all-gather -> build_bvh on every rank"
                               if world > 1 else "host triangles -> build_bvh"),
                "trace_ms": e2e_best / Ke * 1e3, "steps": Ke, "hits_equal_device_path": e2e_ok,
                "path": e2e_path,
                "one_call": ({"mrays_s": e2e_rays * world * Ke / et / 1e6, "ms": et / Ke * 1e3,
                              "note": "rt_render_ao: rays up, hits + frame down in 1M-ray bands over four streams "
                                      "(both PCIe directions at once)"} if wl.ao_spp else None),
                "separate_calls": ({"mrays_s": e2e_rays * world * Ke / es / 1e6, "ms": es / Ke * 1e3,
                                    "note": "trace_rays + rt_gen_ao_rays + trace_rays(any) + rt_shade, whole-frame copies "
                                            "(one PCIe direction at a time, the hits go down under the AO pass)"}
                                   if wl.ao_spp else None),
                "camera": {"mrays_s": count * world * Ke / ec / 1e6, "ms": ec / Ke * 1e3,
                           "h2d_bytes": 48, "d2h_bytes": int(count * 16),
                           "note": "trace_camera: primary rays generated on the device, hits to pinned host",
                           "frame": ({"mrays_s": e2e_rays * world * Ke / eca / 1e6, "ms": eca / Ke * 1e3,
                                      "d2h_bytes": int(count * 28),
                                      "note": "rt_render_camera_ao: the whole frame (primary + AO + shading) from the "
                                              "12-float camera, hits + frame to pinned host"} if wl.ao_spp else None)},
                "note": ("pinned host triangles -> build_bvh; pinned host primary rays up, primary hits + shaded "
                         "frame down, AO rays are device intermediates.  Both public paths are timed every step and "
                         "`value` is the one INTEGRATION.md recommends for this many ranks per host: rt_render_ao "
                         "where a rank has the host's PCIe to itself (N <= 2), the separate calls from four ranks "
                         "on, where the hosts measured move less in both directions at once than in one"
                         if wl.ao_spp else
                         "pinned host tris/rays/hits through build_bvh / trace_rays")},
        "roofline": roof,
        "configs": configs,
    }
    if world == 1 and not args.no_cpu:
        ow = OracleWorkload(wl, host_threads())
        r = ow.step()
        line["cpu_baseline"] = {
            "value": r["rays"] / r["trace_s"] / 1e6, "unit": UNIT, "cores": ow.cores, "kind": "port",
            "sample": ow.sample_text() + f" ({r['build_s']:.2f} s build, {r['trace_s']:.2f} s trace)",
            "build_mtri_s": wl.n_tris / r["build_s"] / 1e6,
            "per_ray_primary": {"nodes_fetched": r["nodes"] / count, "tris_tested": r["tests"] / count}}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="flake10m", choices=["flake10m", "soup1m", "cornell", "random1b"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the config 4 / config 5 sub-records")
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
