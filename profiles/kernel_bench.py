#!/usr/bin/env python
"""Batched-window kernel micro-benchmark (SURVEY.md section 8d "Kernel micro-benchmarks").

A frozen configs[1] window (7 keyframes, ~2000 points, ~12 k residuals) is replicated x B (B = 1, 8, 64: the
residual population of B independent sequences, configs[4]) and pushed through the window kernels K7 / K8 / K9 / K10
as ONE launch each.  Per-kernel CUDA-event times (ldso_gpu_profile_*) against the algorithmic bytes of DESIGN.md
section 4 give the HBM roofline fraction at a size where the launch is bandwidth- rather than latency-bound.
L2 is flushed (256 MiB write) before every repetition.

usage: python profiles/kernel_bench.py [--batches 1,8,64] [--reps 5] [--json out.json]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ldso_b200 import capi, synth  # noqa: E402

CALIB = dict(w=640, h=480, fx=384.0, fy=384.0, cx=319.5, cy=239.5)
# algorithmic bytes per unit
BYTES = {"k7": 900.0, "k8": 288.0 + 32.0, "k9": 72.0, "k10": 60.0}
SLOTS = {"k7": (2, "residuals"), "k8": (3, "residuals"), "k9": (4, "residuals"), "k10": (8, "points")}


replicate = synth.replicate_window


def measure(gpu, batches, reps, peak):
    """rows of {batch, kernel, avg_us, achieved_gbs, frac, ...}: CUDA-event times of K7/K8/K9/K10 on B-fold windows"""
    import torch
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    W0 = synth.make_window(pts_per_frame=286, **CALIB)
    rows = []
    for B in batches:
        W = replicate(W0, B) if B > 1 else W0
        np_, nr = len(W["u"]), len(W["target"])
        cfg = gpu.default_config(n_frame_slots=8, max_points=np_ + 64, max_residuals=nr + 64, **CALIB)
        ctx = gpu.create(cfg)
        for i, img in enumerate(W0["imgs"]):
            ctx.make_images(i, img)
        ctx.window_load(W)
        for rep in range(reps + 2):
            if rep == 2:
                ctx.sync()
                ctx.profile_enable(True)
            flush_buf.fill_(1)
            torch.cuda.synchronize()
            ctx.window_linearize(False)
            ctx.window_apply_res()
            acc = ctx.window_accumulate()
            x = ctx.window_solve(0, 1e-5, acc, W["HM"] * 1e6, W["bM"] * 0, W["delta"] * 0)
            ctx.window_resubstitute(x, 1e-5)
        for key, (slot, unit) in SLOTS.items():
            ms, n, units = ctx.profile_get(slot)
            if n == 0:
                continue
            per = units / n
            us = 1e3 * ms / n
            gbs = per * BYTES[key] / (us * 1e-6) / 1e9
            rows.append(dict(batch=B, kernel=key, launches=int(n), units_per_launch=per, unit=unit, avg_us=us,
                             bytes_per_unit=BYTES[key], achieved_gbs=gbs, peak_gbs=peak, frac=gbs / peak,
                             points=np_, residuals=nr))
        ctx.close()
    del flush_buf
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="1,8,64")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    gpu = capi.load()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("hbm_gbs", 6650.0)
    rows = measure(gpu, [int(b) for b in args.batches.split(",")], args.reps, peak)
    print("%5s %5s %12s %10s %12s %8s" % ("batch", "kern", "units/launch", "avg us", "GB/s (alg.)", "frac"))
    for r in rows:
        print("%5d %5s %12.0f %10.2f %12.1f %8.3f" % (r["batch"], r["kernel"], r["units_per_launch"], r["avg_us"],
                                                      r["achieved_gbs"], r["frac"]))
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
