#!/usr/bin/env python3
"""Times the post-processing kernels (crt_postprocess_device) on a frame rendered
on the device, with the reference's postprocess executable (oracle/_ref, test
infrastructure) timed beside it on the host cores.  Run on the GPU box.

    python tools/profile_post.py [scene3 1920 1080] [--no-cpu]

Prints one JSON line per operation: device ms (CUDA events, frame resident in
HBM, L2 flushed before each run), algorithmic GB/s for the per-pixel maps
(28 B per pixel: 12 read + 4 depth + 12 written), gathered sources per second
for the depth of field, and the CPU reference's seconds.
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.chdir(ROOT)

import _pkg  # noqa: E402

_pkg.load()
import torch  # noqa: E402

from c_raytracer_b200 import _abi, device, host  # noqa: E402
from oracle import oracle  # noqa: E402  (CPU baseline only)

CASES = [
    ("brighten", ["-b", "1.5"]),
    ("mist", ["--mist", "2", "12", "quad", "0.3", "0.3", "0.4"]),
    ("brighten+mist", ["-b", "1.5", "--mist", "2", "12", "quad", "0.3", "0.3", "0.4"]),
    # scene3 depths run from 7.6 to 22.5: circles of confusion up to 3 and up to 15 pixels
    ("dof small", ["--dof", "0.5", "-3.5"]),
    ("dof large", ["--dof", "2.0", "-14"]),
    ("brighten+dof large+mist", ["-b", "1.5", "--dof", "2.0", "-14", "--mist", "2", "12", "quad", "0.3", "0.3", "0.4"]),
]


def cpu_seconds(raw, flags):
    with tempfile.TemporaryDirectory() as d:
        out = subprocess.run([oracle.POSTPROCESS_REF, raw, os.path.join(d, "o.tif"), "-p", "real"] + flags, check=True,
                             capture_output=True, text=True, cwd=ROOT).stdout
    t0 = float(re.search(r"\[(\d+\.\d+)\].*Commencing Postprocessing", out).group(1))
    t1 = float(re.search(r"\[(\d+\.\d+)\].*Saving image", out).group(1))
    return t1 - t0


def main():
    argv = [a for a in sys.argv[1:] if not a.startswith("--")]
    scene, w, h = (argv + ["scene3", "1920", "1080"])[:3] if len(argv) < 3 else argv[:3]
    w, h = int(w), int(h)
    cpu = "--no-cpu" not in sys.argv and os.path.exists(oracle.POSTPROCESS_REF)
    sc = host.Scene(os.path.join(ROOT, "scenes", scene + ".json"))
    dev = device.DeviceScene(sc, 0)
    raster = torch.empty((h, w, 3), dtype=torch.float32, device="cuda:0")
    z = torch.empty((h, w), dtype=torch.float32, device="cuda:0")
    dev.render_device(sc.options(w, h), raster.data_ptr(), z.data_ptr(), None)
    dev.sync()
    frame = raster.clone()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    tmp = tempfile.mkdtemp()
    raw = os.path.join(tmp, "frame.tif")
    if cpu:
        host.save_tiff(raw, frame.cpu().numpy(), z.cpu().numpy(), raw=True)
    zc = z.cpu().numpy()
    for name, flags in CASES:
        o = host.post_options(*flags)
        ms = []
        st = _abi.PostStats()
        for _ in range(5):
            raster.copy_(frame)
            flush.zero_()
            torch.cuda.synchronize()
            device.postprocess_device(o, w, h, raster.data_ptr(), z.data_ptr(), 0, None, st)
            ms.append(st.kernel_ms)
        best = min(ms[1:])
        line = {"op": name, "flags": flags, "frame": f"{scene} {w}x{h}", "kernel_ms": best, "launches": st.launches,
                "dof_max_radius": st.dof_max_radius}
        if not (o.ops & (_abi.CRT_POST_DOF | _abi.CRT_POST_DOF_CAMERA)):
            line["algorithmic_GBps"] = 28.0 * w * h / best / 1e6
        else:
            coc = np.abs(zc * np.float32(st.dof_scale) + np.float32(st.dof_bias)) * np.float32(0.5)
            r = coc.astype(np.int64)
            disc = float((np.pi * r.astype(np.float64) ** 2 + 1).sum())
            line["scattered_pixels"] = disc  # the reference's inner-loop iterations, approximately
            line["Gpixel_adds_per_s"] = disc / best / 1e6
        if cpu:
            line["cpu_ref_seconds"] = cpu_seconds(raw, flags)
            line["speedup_vs_cpu_ref"] = line["cpu_ref_seconds"] * 1e3 / best
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
