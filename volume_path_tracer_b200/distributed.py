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


def shard_tiles(ntiles: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous tile-stripe partition [begin, end) — the low-spp alternative (disjoint pixels, gather instead of
    reduce). Kept for completeness of §8e; the wave sharding above is what render_sharded uses."""
    base, rem = divmod(ntiles, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def reduce_film(film, dst: int = 0):
    """Sum the per-rank (sum XYZ, N) films onto rank `dst`. `film` is a torch tensor on this rank's device."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def render_sharded(renderer, scene, d_film, first_wave: int, n_waves: int, rank: int, world: int, reduce: bool = True):
    """Render this rank's waves of [first_wave, first_wave + n_waves) into the device film tensor `d_film`
    (H, W, 4 float32 on the renderer's device, accumulated into), then reduce onto rank 0."""
    first, count, stride = shard_waves(first_wave, n_waves, rank, world)
    if count:
        renderer.run_device(scene, d_film.data_ptr(), first, count, stride)
    if reduce:
        reduce_film(d_film)
    return d_film
