"""Sort stage alone (A3): rt_sort_pairs against cub::DeviceRadixSort::SortPairs on the same
sizes and key widths ("Compare against cub::DeviceRadixSort on the same
box as a third-party yardstick").  One JSON line per (impl, n, key width).

    python tools/bench_sort.py [--no-cub]

Keys are uniform random (a hash of the position), payload = position; the L2 is flushed before
every timed run; median of 9 after 3 warm-ups.  `alg_gbs` charges SURVEY.md §8(d)'s bytes of
the stage (68 B per 30-bit pair, 200 B per 63-bit pair) to the measured time, for both
implementations, so the two columns are comparable whatever passes each performs.
"""
import json
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gpu_raytracer_b200 as g  # noqa: E402

SIZES = (1_000_000, 9_979_112, 100_026_368)


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def keys_for(n, bits):
    i = torch.arange(n, dtype=torch.int64, device="cuda")
    # splitmix64 finaliser in int64 arithmetic (wraps like uint64)
    z = i + (-7046029254386353131)
    z = (z ^ ((z >> 30) & ((1 << 34) - 1))) * (-4658895280553007687)
    z = (z ^ ((z >> 27) & ((1 << 37) - 1))) * (-7723592293110705685)
    z = z ^ ((z >> 31) & ((1 << 33) - 1))
    k = (z >> (64 - bits)) & ((1 << bits) - 1)
    return k.to(torch.int32) if bits == 30 else k


def ours(n, bits, variant=None):
    if variant is not None:
        for name, v in variant.items():
            g.set_param(name, v)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ms = []
    for it in range(12):
        k = keys_for(n, bits)
        v = torch.arange(n, dtype=torch.int32, device="cuda")
        flush.fill_(it)
        t = g.sort_pairs(k, v, bits)
        if it >= 3:
            ms.append(t)
    ok = bool((k[1:] >= k[:-1]).all())
    ms.sort()
    med = ms[len(ms) // 2]
    alg = (68.0 if bits == 30 else 200.0) * n
    rec = {"impl": "rt_sort_pairs", "n": n, "key_bits": bits, "ms": med, "min_ms": ms[0],
           "gpairs_s": n / med / 1e6, "alg_gbs": alg / med / 1e6, "frac_of_peak": alg / med / 1e6 / peak(),
           "sorted": ok}
    if variant:
        rec["params"] = variant
    print(json.dumps(rec), flush=True)


def main():
    for n in SIZES:
        for bits in (30, 63):
            ours(n, bits)
    if "--no-cub" not in sys.argv:
        exe = os.path.join(ROOT, "tools", "ubench", "cub_sort")
        if os.path.exists(exe):
            out = subprocess.run([exe], capture_output=True, text=True)
            for line in out.stdout.splitlines():
                try:
                    rec = json.loads(line)
                    rec["frac_of_peak"] = rec["alg_gbs"] / peak()
                    print(json.dumps(rec), flush=True)
                except ValueError:
                    print(line, flush=True)
        else:
            print(json.dumps({"impl": "cub", "error": "tools/ubench/cub_sort not built (nvcc it first)"}))


if __name__ == "__main__":
    main()
