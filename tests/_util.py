"""Shared helpers of the test-suite."""
import glob
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

from c_raytracer_b200 import host  # noqa: E402


def scene_path(name):
    return os.path.join(ROOT, "scenes", name + ".json")


def golden_images():
    return sorted(os.path.basename(p)[len("image_"):-4] for p in glob.glob(os.path.join(GOLDEN, "image_*.npz")))


def load_golden_image(name):
    g = np.load(os.path.join(GOLDEN, f"image_{name}.npz"))
    flags = [str(f) for f in g["flags"]]
    return {"raster": g["raster"], "z": g["z"], "scene": str(g["scene"]), "width": int(g["width"]),
            "height": int(g["height"]), "flags": flags, "cast_ray": int(g["cast_ray"]), "shadow": int(g["shadow"]),
            "seed_to_seed": float(g["seed_to_seed_rel_rmse"])}


def open_case(case):
    """(Scene, RenderOpts) of a golden image case; -r goes to the loader, the rest to the options."""
    flags = list(case["flags"])
    rescale = None
    if "-r" in flags:
        i = flags.index("-r")
        rescale = flags[i + 1]
        del flags[i:i + 2]
    sc = host.Scene(scene_path(case["scene"]), rescale=rescale)
    return sc, sc.options(case["width"], case["height"], *flags)


def rel_rmse(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b) / max(np.linalg.norm(b.astype(np.float64)), 1e-30))


def quant_diff(a, b):
    """max |difference| after the reference's truncating 8-bit quantiser, and the
    number of pixels that differ by more than one code."""
    qa = host.quantize(a).astype(np.int32)
    qb = host.quantize(b).astype(np.int32)
    d = np.abs(qa - qb).max(axis=-1)
    return int(d.max()), int((d > 1).sum())
