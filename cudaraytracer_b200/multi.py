"""Host-side logic of one-process-per-GPU rendering.

The image is cut into CRT_TILE x CRT_TILE tiles numbered row-major; rank r of
`world` owns tiles t with t % world == r (include/crt.h). A rank's
crt_render_accum_device writes its own pixels and zeros elsewhere, so the only
exchange step of the path is a sum-reduce of the accumulators to rank 0
(ncclReduce over NVLink; gloo in the CPU tests). Adding zeros is exact, so the
combined image is bit-identical to the single-GPU one.
"""
import numpy as np

TILE = 32


def tile_grid(nx, ny):
    return (nx + TILE - 1) // TILE, (ny + TILE - 1) // TILE


def local_tile_count(nx, ny, rank, world):
    tx, ty = tile_grid(nx, ny)
    tiles = tx * ty
    return (tiles - rank + world - 1) // world if tiles > rank else 0


def owned_mask(nx, ny, rank, world):
    """Boolean (ny, nx) mask of the pixels rank `rank` renders (row 0 = bottom,
    like the reference's colorBuffer)."""
    tx, _ = tile_grid(nx, ny)
    y, x = np.mgrid[0:ny, 0:nx]
    t = (y // TILE) * tx + (x // TILE)
    return (t % world) == rank


def paths_per_rank(nx, ny, spp, world):
    return [int(owned_mask(nx, ny, r, world).sum()) * spp for r in range(world)]


def combine_accum(accum, dst=0, group=None):
    """Sum-reduce the per-rank accumulators onto rank `dst` (in place)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst, op=dist.ReduceOp.SUM, group=group)
    return accum
