#!/usr/bin/env python3
"""Print the headline figures of bench / shard-probe JSON files under gpurun_out/ (local convenience)."""
import json, sys, os
d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
for f in ("bench.json", "bench_n2.json", "bench_n4.json", "bench_n8.json"):
    p = os.path.join(d, f)
    if not os.path.exists(p): continue
    try:
        b = json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    r = b["roofline"]
    print(f, "value %.0f e2e %.0f frac %.3f k1 %.2fus k2 %.2fus | batch1: frac %.3f k1 %.2fus k2 %.1fus" % (
        b["value"], b["e2e"]["value"], r["frac"], 1e3 * r["pack_alloc_kernel_ms_per_frame"], 1e3 * r["fuse_kernel_ms_per_frame"],
        r["single_frame_pass"]["frac"], 1e3 * r["single_frame_pass"]["pack_alloc_kernel_ms_per_frame"], 1e3 * r["single_frame_pass"]["avg_launch_ms"]))
p = os.path.join(d, "shard_probe.json")
if os.path.exists(p):
    for row in json.load(open(p))["rows"]:
        print("probe world %d batch %d: %.2f us/frame (%.0f f/s) k1 %.2f k2 %.2f" % (row["world"], row["batch"], row["us_per_frame"], row["rank_frames_per_s"], row["k1_us_per_frame"], row["k2_us_per_frame"]))
