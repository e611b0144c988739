"""Frame sharding across the GPUs of one node -- one process per GPU.

The reference parallelises render() over image rows with OpenMP (reference
src/raytracer/render.c:349-352); pixels are independent and the scene is
read-only.  Here the scene is replicated on every GPU and the frame is cut by
the partition of include/crt.h (8x4-pixel tiles dealt round-robin to the
ranks, `crt_render_opts.shard_index / shard_count`): the interleave is fine
enough (32 pixels) to even out the >10x cost differences
between sky, floor and glass.  Every rank renders its chunks into a zeroed
full-size frame; the only exchange is the assembly of the disjoint shards on
rank 0 -- one reduce(sum) over NVLink (NCCL) per buffer, 16 B per pixel.
x + 0 is exact, and the RNG is keyed on the pixel, so the assembled frame is
bit-identical to the single-GPU frame.

The render callable is injected so that the same host logic runs on the CUDA
path (device.DeviceScene.render_device) and, in the CPU test-suite, over gloo.
"""
import torch
import torch.distributed as dist


def shard_options(opts, rank, world):
    """Marks `opts` (an _abi.RenderOpts) as shard `rank` of `world`."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    opts.shard_index, opts.shard_count = rank, world
    return opts


def shard_of_pixels(width, height, world, tile_w=8, tile_h=4, chunk_tiles=1):
    """(H, W) int tensor: owning shard of every pixel (the partition of include/crt.h)."""
    tiles_x = (width + tile_w - 1) // tile_w
    ys = torch.arange(height).unsqueeze(1)
    xs = torch.arange(width).unsqueeze(0)
    tile = (ys // tile_h) * tiles_x + xs // tile_w
    return (tile // chunk_tiles) % world


def assemble(raster, z_buffer, dst=0, group=None):
    """Sums the zero-elsewhere shard frames onto rank `dst` (in place there).
    Tensors live wherever the process group's backend needs them (CUDA for
    NCCL, CPU for gloo)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return raster, z_buffer
    dist.reduce(raster, dst, op=dist.ReduceOp.SUM, group=group)
    dist.reduce(z_buffer, dst, op=dist.ReduceOp.SUM, group=group)
    return raster, z_buffer


def render_sharded(render, opts, raster, z_buffer, group=None):
    """One frame on all ranks: `render(opts, raster, z_buffer)` fills this rank's
    shard of the zero-initialised frame, then the shards are assembled on rank 0."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    shard_options(opts, rank, world)
    render(opts, raster, z_buffer)
    return assemble(raster, z_buffer, 0, group)
