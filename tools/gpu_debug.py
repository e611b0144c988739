"""Debug helper (run on the GPU box): render one configuration on the GPU and with the
CPU oracle, print the worst pixels and dump both images to gpurun_out/debug_<tag>.npz.

    python tools/gpu_debug.py scene5 96 54 -- -g path -n 4
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.chdir(ROOT)
import _pkg  # noqa: E402

_pkg.load()
from c_raytracer_b200 import _abi, device, host  # noqa: E402
from oracle import oracle  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("scene")
ap.add_argument("width", type=int)
ap.add_argument("height", type=int)
argv = sys.argv[1:]
flags = []
if "--" in argv:
    k = argv.index("--")
    argv, flags = argv[:k], argv[k + 1:]
a = ap.parse_args(argv)
a.flags = flags

sc = host.Scene(os.path.join(ROOT, "scenes", a.scene + ".json"))
dev = device.DeviceScene(sc, 0)
orc = oracle.OracleScene(sc)
o = sc.options(a.width, a.height, *a.flags)
o.flags |= _abi.CRT_RENDER_COUNT_STATS
st, ost = _abi.Stats(), _abi.Stats()
r, z = dev.render(o, st)
rr, rz = orc.render(o, ost)
rel = np.linalg.norm(r - rr) / np.linalg.norm(rr)
print("rel-RMSE", rel, "rays", st.rays_closest, st.rays_shadow, "oracle", ost.rays_closest, ost.rays_shadow,
      "z mismatches", int((z != rz).sum()))
d = np.abs(r - rr).max(axis=2)
order = np.argsort(d.ravel())[::-1][:12]
for p in order:
    y, x = divmod(int(p), a.width)
    print(f"  px ({x},{y}) gpu {r[y, x]} oracle {rr[y, x]} z {z[y, x]} {rz[y, x]}")
print("pixels with |diff| > 1e-6:", int((d > 1e-6).sum()), "of", d.size)
tag = a.scene + "_" + "_".join(f.strip("-") for f in a.flags)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", f"debug_{tag}.npz"), gpu=r, oracle=rr, zg=z, zo=rz)
