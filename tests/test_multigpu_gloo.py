"""The N > 1 path on CPU: world_size 2 and 3 over gloo.  The host logic of
c-raytracer_b200/multigpu.py (shard options, assembly by reduce) is the code
bench.py runs over NCCL; here the per-rank render callable is the CPU oracle
(test infrastructure), which implements the same frame partition of
include/crt.h, so the assembled frame must equal the unsharded frame bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _gloo_worker
from _util import ROOT, scene_path
from c_raytracer_b200 import host, multigpu
from oracle import oracle


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,scene,w,h,flags", [
    (2, "scene1", 96, 54, ()),
    (3, "scene3", 70, 37, ("-g", "path", "-n", "2")),
])
def test_sharded_frame_equals_whole_frame(tmp_path, world, scene, w, h, flags):
    out = str(tmp_path / "frame.npz")
    mp.spawn(_gloo_worker.worker, args=(world, _free_port(), scene, w, h, flags, out), nprocs=world, join=True)
    got = np.load(out)
    sc = host.Scene(scene_path(scene))
    want, want_z = oracle.OracleScene(sc).render(sc.options(w, h, *flags))
    np.testing.assert_array_equal(got["raster"], want)
    np.testing.assert_array_equal(got["z"], want_z)


def test_partition_is_a_partition():
    for world in (1, 2, 3, 8):
        owner = multigpu.shard_of_pixels(1920, 1080, world)
        counts = torch.bincount(owner.flatten(), minlength=world)
        assert int(counts.sum()) == 1920 * 1080
        # round-robin tiles: shard sizes differ by at most a few tiles of 32 pixels
        assert int(counts.max() - counts.min()) <= 8 * 32
    with pytest.raises(ValueError):
        from c_raytracer_b200 import _abi
        multigpu.shard_options(_abi.RenderOpts(), 2, 2)
