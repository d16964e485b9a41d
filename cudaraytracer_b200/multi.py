"""Host-side logic of one-process-per-GPU rendering.

The image is cut into CRT_TILE x CRT_TILE tiles numbered row-major; rank r of
`world` owns tiles t with t % world == r (include/crt.h). A rank's
crt_render_accum_device writes its own pixels and zeros elsewhere, so the only
exchange step of the path is a sum-reduce of the accumulators to rank 0. Adding
zeros is exact, so the combined image is bit-identical to the single-GPU one.

The exchange itself lives behind the C ABI (csrc/comm.cu: crt_comm_init,
crt_reduce_accum_device, crt_scene_broadcast, crt_render_dist; NCCL over
NVLink). This module is a CALLER of it: `init_comm` carries the communicator id
from rank 0 to the other ranks over an existing torch.distributed group (any
transport would do; torchrun has already set one up), the rest is tile
arithmetic the tests use. `combine_accum` over torch.distributed remains for
the CPU (gloo) tests, where there is no NCCL.
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


def init_comm(ctx, group=None):
    """Give `ctx` its rank in a library-owned NCCL communicator spanning the torch.distributed group: rank 0 draws the
    id (crt_comm_unique_id), the group carries the 128 bytes, every rank calls crt_comm_init."""
    import torch.distributed as dist
    from . import comm_unique_id
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    ctx.comm_init(box[0], rank, world)
    return rank, world


def combine_accum(accum, dst=0, group=None):
    """Sum-reduce per-rank accumulators held in torch tensors onto rank `dst` (in place) over torch.distributed:
    the CPU (gloo) stand-in of crt_reduce_accum_device."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst, op=dist.ReduceOp.SUM, group=group)
    return accum
