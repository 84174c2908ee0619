"""Multi-GPU rendering: one process per GPU, the baked volume replicated in every GPU's HBM, the
(tile, wave) job space sharded by WAVES, films combined with ONE collective.

The reference has a single shared-memory strategy: num_workers jthreads pulling (tile, wave) jobs off one
atomic counter (src/main.cpp:63-87, src/tile_provider.cpp:27-67). Jobs are independent given their id
(jid = (wave-1)*ntiles + tile seeds the job's RNG, include/vpt/random.hpp:93-95) and the film is a plain sum
(src/worker.cpp:203-204), so rank r of R renders waves {first + r, first + r + R, ...} over the full frame
with the jids unchanged: the union of the ranks' RNG streams is exactly the single-GPU run, sample counts add
exactly, and radiance sums differ from the serial order only by fp32 add order.

The only exchange step is the film sum: torch.distributed.reduce (NCCL over NVLink / NVSwitch on GPUs; gloo in
the CPU tests) of H*W*4 floats to rank 0.
"""
from __future__ import annotations

from typing import Tuple


def shard_waves(first_wave: int, n_waves: int, rank: int, world: int) -> Tuple[int, int, int]:
    """-> (first, count, stride) for vptc_render_device: waves first + k*stride, k < count.
    Interleaved so every rank gets waves from the whole range (progressive previews stay unbiased)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    if first_wave < 1 or n_waves < 0:
        raise ValueError("waves are 1-based (tile_provider.cpp:30)")
    count = (n_waves - rank + world - 1) // world if n_waves > rank else 0
    return first_wave + rank, count, world


def shard_tile_rows(scene, rank: int, world: int):
    """Tile-stripe partition for low spp / large frames (the reference numbers tiles row-major, src/tile_provider.cpp:95-105): rank r takes
    the ROWS of tiles r, r + world, r + 2*world, ... — interleaved, so that the ranks share the expensive part of the frame evenly (contiguous
    halves of the 4K cloud differed by 70 % in cost). -> (tile_begin, tile_count, tile_row_stride, rows) for vptc_job_range /
    Renderer.run_device_range, `rows` = the film rows this rank renders (a torch-indexable list of slices, one per row of tiles).
    Disjoint pixels across ranks, so the films are GATHERED, not summed."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    W, H, tw, th = scene.width, scene.height, int(scene.tile_size[0]), int(scene.tile_size[1])
    ntx, nty = -(-W // tw), -(-H // th)
    mine = list(range(rank, nty, world))
    return rank * ntx, len(mine) * ntx, world, [slice(r * th, min(H, (r + 1) * th)) for r in mine]


def reduce_film(film, dst: int = 0):
    """Sum the per-rank (sum XYZ, N) films onto rank `dst`. `film` is a torch tensor on this rank's device."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def gather_film_stripes(film, scene, rank: int, world: int, dst: int = 0):
    """Tile-stripe sharding: every rank's bands of film rows (shard_tile_rows) are gathered into rank `dst`'s film with ONE collective
    (NCCL gather over NVLink on GPUs, gloo in the CPU tests). Ranks differ by at most one row of tiles: padded to the largest for the call."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return film

    def row_index(r):
        sl = shard_tile_rows(scene, r, world)[3]
        return torch.cat([torch.arange(s.start, s.stop) for s in sl]) if sl else torch.zeros(0, dtype=torch.long)
    idx = [row_index(r) for r in range(world)]
    rows = max(len(i) for i in idx)
    send = torch.zeros((rows,) + tuple(film.shape[1:]), dtype=film.dtype, device=film.device)
    mine = idx[rank].to(film.device)
    send[: len(mine)] = film.index_select(0, mine)
    recv = [torch.empty_like(send) for _ in range(world)] if rank == dst else None
    dist.gather(send, recv, dst=dst)
    if rank == dst:
        for r in range(world):
            if r != dst and len(idx[r]):
                film.index_copy_(0, idx[r].to(film.device), recv[r][: len(idx[r])])
    return film


def render_sharded(renderer, scene, d_film, first_wave: int, n_waves: int, rank: int, world: int, reduce: bool = True, scratch=None):
    """Render this rank's waves of [first_wave, first_wave + n_waves) and accumulate the COMBINED result into rank 0's `d_film`
    (H, W, 4 float32 on the renderer's device). The reduce runs on a per-call partial film (`scratch`, allocated when not given), never on
    the accumulated one: reducing `d_film` itself would re-add every earlier call's contribution of the other ranks on the second call."""
    import torch
    first, count, stride = shard_waves(first_wave, n_waves, rank, world)
    if world == 1 or not reduce:
        if count:
            renderer.run_device(scene, d_film.data_ptr(), first, count, stride)
        return d_film
    part = scratch if scratch is not None else torch.empty_like(d_film)
    part.zero_()
    if count:
        renderer.run_device(scene, part.data_ptr(), first, count, stride)
    reduce_film(part)
    if rank == 0:
        d_film.add_(part)
    return d_film
