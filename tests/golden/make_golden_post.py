#!/usr/bin/env python3
"""Golden fixtures of the post-processing path, made by RUNNING THE REFERENCE:
the unmodified src/postprocess/postproc.c + image.c (compiled in place into
oracle/_ref/ref_post by oracle/Makefile, IEEE flags) applied to golden frames of
the reference engine (tests/golden/image_*.npz), through the reference's own
command-line flags.  The FP32 raster postprocess() leaves behind is stored.

Runs only where /root/reference exists; the outputs are committed.

    python tests/golden/make_golden_post.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import _pkg  # noqa: E402

_pkg.load()
from c_raytracer_b200 import host  # noqa: E402
from oracle import oracle  # noqa: E402

# (name, input golden image, postprocess flags)
POST_CASES = [
    ("brighten", "scene1_ambient", ("-b", "1.75")),
    ("mist_lin", "scene1_ambient", ("--mist", "3", "9", "lin", "0.5", "0.55", "0.7")),
    ("mist_quad", "scene4_ambient", ("--mist", "2.5", "6", "quad", "0.8", "0.8", "0.9")),
    ("mist_invquad", "scene3_ambient", ("--mist", "1", "14", "inv-quad", "0.1", "0.2", "0.3")),
    ("dof_small", "scene3_ambient", ("--dof", "0.9", "-6")),
    ("dof_sky", "scene1_ambient", ("--dof", "0.35", "-3")),        # z = 0 background pixels in front of everything
    ("dof_wide", "scene4_ambient", ("--dof", "2.5", "-14")),       # discs larger than the frame
    ("dof_zero_radius", "scene2_ambient", ("--dof", "0.01", "0.5")),
    ("dof_odd_size", "scene2_odd_size", ("--dof", "0.6", "-4")),
    ("dof_camera", "scene3_ambient", ("--dof-camera", "9", "0.6", "5.5")),
    ("dof_wins_over_camera", "scene3_ambient", ("--dof-camera", "9", "0.6", "5.5", "--dof", "0.5", "-3")),
    ("all_three", "scene4_ambient", ("-b", "0.8", "--dof", "1.1", "-6", "--mist", "3", "10", "quad", "0.6", "0.65", "0.7")),
    ("brighten_mist", "scene5s_ambient", ("--mist", "4", "4", "lin", "1", "1", "1", "-b", "2.5")),
    ("path_frame", "scene6s_path4", ("--dof", "0.8", "-5", "-b", "1.2")),
]


def main():
    if not os.path.exists(oracle.REF_POST):
        sys.exit("oracle/_ref/ref_post is missing: run `make -C oracle ref` where /root/reference exists")
    with tempfile.TemporaryDirectory() as d:
        for name, image, flags in POST_CASES:
            g = np.load(os.path.join(HERE, f"image_{image}.npz"))
            raw = os.path.join(d, "in.tif")
            host.save_tiff(raw, g["raster"], g["z"], raw=True)
            out = oracle.ref_postprocess(raw, flags)
            np.savez_compressed(os.path.join(HERE, f"post_{name}.npz"), raster=out, image=image, flags=np.array(flags, dtype=str))
            print(f"post_{name}: {image} {' '.join(flags)} -> max {out.max():.4f} changed {float((out != g['raster']).mean()):.2f}")


if __name__ == "__main__":
    main()
