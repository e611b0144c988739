#!/usr/bin/env python3
"""Render a few frames of one configuration through the C ABI and print the
library's own statistics as JSON -- the short command that is run plain and then
under ncu (profiles/).  Run on the GPU box.

    python tools/profile_frame.py scene5 480 270 --frames 3 -- -g path -n 4
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.chdir(ROOT)

import _pkg  # noqa: E402

_pkg.load()
from c_raytracer_b200 import _abi, device, host  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene")
    ap.add_argument("width", type=int)
    ap.add_argument("height", type=int)
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--count", action="store_true", help="also run one instrumented frame first")
    ap.add_argument("--shards", type=int, default=1, help="render each of N shards in turn on this GPU (what N ranks would do)")
    argv = sys.argv[1:]
    flags = []
    if "--" in argv:
        k = argv.index("--")
        argv, flags = argv[:k], argv[k + 1:]
    a = ap.parse_args(argv)
    a.flags = flags
    sc = host.Scene(os.path.join(ROOT, "scenes", a.scene + ".json"))
    dev = device.DeviceScene(sc, 0)
    out = {"scene": a.scene, "size": [a.width, a.height], "flags": a.flags,
           "bvh": {"nodes": dev.info.num_nodes, "prims": dev.info.num_bounded, "depth": dev.info.max_depth,
                   "build_ms": dev.info.build_ms, "upload_ms": dev.info.upload_ms, "bytes": dev.info.device_bytes}}
    if a.count:
        o = sc.options(a.width, a.height, *a.flags)
        o.flags |= _abi.CRT_RENDER_COUNT_STATS
        st = _abi.Stats()
        dev.render(o, st)
        d = st.as_dict()
        rays = d["rays_closest"] + d["rays_shadow"]
        d["flops"] = _abi.algorithmic_flops(st)
        d["flops_per_ray"] = d["flops"] / max(rays, 1)
        d["box_per_ray"] = d["box_tests"] / max(rays, 1)
        out["counted"] = d
    o = sc.options(a.width, a.height, *a.flags)
    ms = []
    for _ in range(a.frames):
        st = _abi.Stats()
        dev.render(o, st)
        ms.append(st.kernel_ms)
    out["kernel_ms"] = ms
    if a.shards > 1:
        per = []
        for i in range(a.shards):
            o.shard_index, o.shard_count = i, a.shards
            st = _abi.Stats()
            dev.render(o, st)
            per.append(round(st.kernel_ms, 2))
        out["shard_kernel_ms"] = per
        out["shards_max_ms"], out["shards_mean_ms"] = max(per), sum(per) / len(per)
    if a.count:
        rays = out["counted"]["rays_closest"] + out["counted"]["rays_shadow"]
        best = min(ms)
        out["mrays_s"] = rays / best / 1e3
        out["tflops"] = out["counted"]["flops"] / best / 1e9
    print(json.dumps(out))


if __name__ == "__main__":
    main()
