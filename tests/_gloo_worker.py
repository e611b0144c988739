"""Worker of tests/test_multigpu_gloo.py (a separate module so that spawned
processes can import it: it sets up the package loader itself)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import _pkg  # noqa: E402

_pkg.load()

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from _util import scene_path  # noqa: E402
from c_raytracer_b200 import host, multigpu  # noqa: E402
from oracle import oracle  # noqa: E402


def worker(rank, world, port, scene, w, h, flags, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["OMP_NUM_THREADS"] = "2"
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = host.Scene(scene_path(scene))
        orc = oracle.OracleScene(sc)
        opts = sc.options(w, h, *flags)
        raster = torch.zeros((h, w, 3), dtype=torch.float32)
        z = torch.zeros((h, w), dtype=torch.float32)

        def render(o, r, zb):
            a, b = orc.render(o)
            r.copy_(torch.from_numpy(a))
            zb.copy_(torch.from_numpy(b))

        # what this rank rendered covers exactly its pixels of the partition
        multigpu.shard_options(opts, rank, world)
        render(opts, raster, z)
        owner = multigpu.shard_of_pixels(w, h, world)
        mine = owner == rank
        assert float(raster[~mine].abs().sum()) == 0.0 and float(z[~mine].abs().sum()) == 0.0
        multigpu.assemble(raster, z, 0)
        if rank == 0:
            np.savez(out_path, raster=raster.numpy(), z=z.numpy())
    finally:
        dist.destroy_process_group()


