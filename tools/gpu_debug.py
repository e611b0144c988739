"""Debug helper (run on the GPU box): dump GPU and oracle images + ray-batch diffs to gpurun_out/."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import _pkg; _pkg.load()
from c_raytracer_b200 import _abi, device, host
from oracle import oracle

out = {}
for name, W, H, flags in [("scene1", 96, 54, ()), ("scene3", 96, 54, ()), ("scene4", 64, 36, ()), ("scene1", 48, 27, ("-g", "path", "-n", "4"))]:
    sc = host.Scene(os.path.join(ROOT, "scenes", name + ".json"))
    dev = device.DeviceScene(sc, 0)
    orc = oracle.OracleScene(sc)
    o = sc.options(W, H, *flags)
    o.flags |= _abi.CRT_RENDER_COUNT_STATS
    st = _abi.Stats(); ost = _abi.Stats()
    r, z = dev.render(o, st)
    rr, rz = orc.render(o, ost)
    key = name + "_" + "_".join(flags)
    out[key + "_gpu"] = r; out[key + "_ora"] = rr; out[key + "_zg"] = z; out[key + "_zo"] = rz
    rel = np.linalg.norm(r - rr) / np.linalg.norm(rr)
    print(key, "rel", rel, "rays", st.rays_closest, st.rays_shadow, ost.rays_closest, ost.rays_shadow, "zdiff", (z != rz).sum(), "pxdiff>1e-5", (np.abs(r - rr).max(axis=2) > 1e-5).sum())
    # ray batches
    rng = np.random.default_rng(1)
    n = 20000
    objs = sc.objects
    b = objs[objs["type"] != 2]
    cen = b["a"]
    lo, hi = cen.min(0) - 1, cen.max(0) + 1
    org = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    tgt = cen[rng.integers(0, len(cen), n)] + rng.normal(0, 0.3, size=(n, 3))
    d = (tgt - org).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([d, org], 1).astype(np.float32)
    go, gt, gn = dev.trace_closest(rays)
    oo, ot, on = orc.trace_closest(rays)
    print("  closest: obj mismatch", (go != oo).sum(), "t maxrel", np.nanmax(np.abs(gt - ot)[go == oo] / np.maximum(ot[go == oo], 1e-9)) if (go == oo).any() else None, "normal maxabs", np.abs(gn - on)[go == oo].max())
    dist = rng.uniform(0.2, 12, n).astype(np.float32)
    em = np.full(n, -1, np.int32)
    light = np.ones((n, 3), np.float32)
    gb, gl = dev.trace_shadow(rays, dist, em, light)
    ob, ol = orc.trace_shadow(rays, dist, em, light)
    print("  shadow: blocked mismatch", (gb != ob).sum(), "light maxabs", np.abs(gl - ol)[(gb == 0) & (ob == 0)].max() if ((gb == 0) & (ob == 0)).any() else None)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "debug.npz"), **out)
