#!/usr/bin/env python3
"""Fills BASELINE.md section 5: every configuration of BASELINE.json on one GPU
(device-timed frames, counted rays / box tests / FLOPs, roofline fraction) with
the compiled reference (oracle/_ref, test infrastructure) timed and compared
beside it on the host cores.  Run on the GPU box; prints JSON lines and a
markdown table.

    python tools/results_table.py [--only scene5]
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.chdir(ROOT)

import _pkg  # noqa: E402

_pkg.load()
from c_raytracer_b200 import _abi, device, host, parity  # noqa: E402
from oracle import oracle  # noqa: E402  (checker and CPU baseline only)

# (scene, width, height, flags, (cpu_w, cpu_h) of the reference run)
CONFIGS = [
    ("scene1", 1920, 1080, [], (960, 540)),
    ("scene3", 1920, 1080, [], (960, 540)),
    ("scene4", 1920, 1080, [], (480, 270)),
    ("scene5", 1920, 1080, ["-g", "path", "-n", "16"], (192, 108)),
    ("scene6", 3840, 2160, ["-g", "path", "-n", "64"], (480, 270)),
]


def stock_seconds(scene, w, h, flags, exe):
    with tempfile.TemporaryDirectory() as d:
        out = subprocess.run([exe, scene, os.path.join(d, "o.tif"), str(w), str(h), "-m", "max", "-p", "real", "-f"] + flags,
                             cwd=ROOT, check=True, capture_output=True, text=True).stdout
    cores = int(re.search(r"Using (\d+) threads", out).group(1))
    t0 = float(re.search(r"\[(\d+\.\d+)\].*Commencing raytracing", out).group(1))
    t1 = float(re.search(r"\[(\d+\.\d+)\].*Saving image", out).group(1))
    return t1 - t0, cores


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--stock", action="store_true", help="also time the reference as shipped (global rand(): 6-60x slower)")
    a = ap.parse_args()
    peak = device.measure_fp32_peak(0)
    rows = []
    for scene, w, h, flags, (cw, ch) in CONFIGS:
        if a.only and a.only != scene:
            continue
        path = os.path.join("scenes", scene + ".json")
        sc = host.Scene(os.path.join(ROOT, path))
        dev = device.DeviceScene(sc, 0)
        o = sc.options(w, h, *flags)
        o.flags |= _abi.CRT_RENDER_COUNT_STATS
        st = _abi.Stats()
        dev.render(o, st)
        rays = st.rays_closest + st.rays_shadow
        flops = _abi.algorithmic_flops(st)
        o = sc.options(w, h, *flags)
        ms = []
        for _ in range(a.frames):
            s2 = _abi.Stats()
            dev.render(o, s2)
            ms.append(s2.kernel_ms)
        best = min(ms)
        # end to end through crt_render(): D2H of raster + z inside the call (Stats.total_ms)
        #   pageable = fresh numpy arrays per frame, the way the reference's image_init() allocates;
        #   touched  = the same arrays again (pages already faulted in);
        #   pinned   = those arrays after crt_host_register()
        e2e = {}
        s3 = _abi.Stats()
        dev.render(o, s3)
        e2e["pageable_fresh_ms"] = s3.total_ms
        raster = np.empty((h, w, 3), dtype=np.float32)
        zbuf = np.empty((h, w), dtype=np.float32)
        dev.render_into(o, raster, zbuf, s3)
        t_touch = []
        for _ in range(a.frames):
            dev.render_into(o, raster, zbuf, s3)
            t_touch.append(s3.total_ms)
        e2e["pageable_touched_ms"] = min(t_touch)
        t = time.perf_counter()
        device.host_register(raster)
        device.host_register(zbuf)
        e2e["register_ms"] = (time.perf_counter() - t) * 1e3
        t_pin = []
        for _ in range(a.frames):
            dev.render_into(o, raster, zbuf, s3)
            t_pin.append(s3.total_ms)
        e2e["pinned_ms"] = min(t_pin)
        e2e["kernel_ms_same_frames"] = s3.kernel_ms
        e2e["d2h_bytes"] = raster.nbytes + zbuf.nbytes
        device.host_unregister(raster)
        device.host_unregister(zbuf)
        # the reference beside it: the RNG-threaded build for parity, the same build and the stock one for time
        t = time.time()
        ref, ref_z, cnt = oracle.ref_render(path, cw, ch, flags, seed=1)
        t_instr = time.time() - t
        got, got_z = dev.render(sc.options(cw, ch, *flags))
        s2s = None
        if "path" in flags: # stochastic tier: the bound is twice the reference's own seed-to-seed distance
            other = oracle.ref_render(path, cw, ch, flags, seed=2)[0]
            s2s = float(np.linalg.norm(other.astype(np.float64) - ref) / np.linalg.norm(ref.astype(np.float64)))
        p = parity.frame_parity(got, ref, s2s)
        q8 = int(np.abs(host.quantize(got).astype(int) - host.quantize(ref).astype(int)).max())
        secs, cores = stock_seconds(path, cw, ch, flags, oracle.ENGINE_REF)
        secs_stock = stock_seconds(path, cw, ch, flags, oracle.ENGINE_REF_STOCK)[0] if a.stock else float("nan")
        cpu_rays = cnt["cast_ray"] + cnt["shadow"]
        row = {
            "config": f"{scene} {w}x{h} {' '.join(flags)}".strip(), "gpu_ms": best, "rays": rays,
            "gpu_mrays_s": rays / best / 1e3, "box_per_ray": st.box_tests / max(rays, 1), "flop_per_ray": flops / max(rays, 1),
            "tflops": flops / best / 1e9, "fp32_peak_tflops": peak, "roofline_frac": flops / best / 1e9 / peak,
            "e2e": e2e, "e2e_mrays_s_pinned": rays / e2e["pinned_ms"] / 1e3,
            "bvh": {"nodes": dev.info.num_nodes, "prims": dev.info.num_bounded, "build_ms": dev.info.build_ms},
            "cpu": {"size": [cw, ch], "cores": cores, "rays": cpu_rays, "seconds_rng_patched": secs, "seconds_stock": secs_stock,
                    "mrays_s_rng_patched": cpu_rays / secs / 1e6, "mrays_s_stock": cpu_rays / secs_stock / 1e6,
                    "instrumented_seconds": t_instr, "box_per_ray": cnt["box_tests"] / max(cpu_rays, 1)},
            "parity": {"size": [cw, ch], "rel_rmse": p["rel_rmse"], "rel_rmse_all": p["rel_rmse_all"], "outliers": p["outliers"],
                       "allowed": p["allowed"], "max_8bit_delta": q8, "z_equal": bool(np.array_equal(got_z, ref_z)), "ok": bool(p["ok"]),
                       "tier": p.get("tier"), "seed_to_seed_rel_rmse": s2s},
        }
        rows.append(row)
        print(json.dumps(row), flush=True)
        del dev
    print()
    print("| Config | GPU ms/frame | GPU Mrays/s | e2e ms: pageable fresh / touched / pinned (+ register once) | box/ray GPU / ref | FLOP/ray | FP32 frac | CPU ref (cores, size) Mrays/s patched / stock | rel-RMSE / rel-RMSE all / 8-bit max / outliers |")
    print("|---|---|---|---|---|---|---|---|---|")
    for r in rows:
        c, p = r["cpu"], r["parity"]
        e = r["e2e"]
        print(f"| {r['config']} | {r['gpu_ms']:.1f} | {r['gpu_mrays_s']:.0f} | {e['pageable_fresh_ms']:.1f} / {e['pageable_touched_ms']:.1f} / {e['pinned_ms']:.1f} (+{e['register_ms']:.1f}) | "
              f"{r['box_per_ray']:.1f} / {c['box_per_ray']:.1f} | {r['flop_per_ray']:.0f} | "
              f"{r['roofline_frac']:.3f} | {c['cores']} cores, {c['size'][0]}x{c['size'][1]}: {c['mrays_s_rng_patched']:.2f} / {c['mrays_s_stock']:.2f} | "
              f"{p['rel_rmse']:.1e} / {p['rel_rmse_all']:.1e} / {p['max_8bit_delta']} / {p['outliers']} of {p['allowed']} allowed |")


if __name__ == "__main__":
    main()
