"""Multi-GPU host logic on CPU: wave sharding + the one film reduce, world_size 2 over gloo.

Each rank renders ITS waves with the CPU oracle (tests may use the oracle as the stand-in renderer: what is
under test here is the sharding arithmetic and the collective, not the kernels), reduces the float4 film onto
rank 0 through volume_path_tracer_b200.distributed.reduce_film, and rank 0 compares with the single-process
render of all waves: sample counts bit-equal, radiance sums equal up to fp32 add order.
"""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_shard_waves_partition():
    from volume_path_tracer_b200.distributed import shard_tiles, shard_waves
    for first, n, world in [(1, 8, 2), (1, 7, 4), (5, 3, 8), (1, 0, 2), (3, 1, 1), (1, 4096, 8)]:
        seen = []
        for r in range(world):
            f, c, s = shard_waves(first, n, r, world)
            assert s == world
            seen += [f + k * s for k in range(c)]
        assert sorted(seen) == list(range(first, first + n)), (first, n, world)
    for nt, world in [(11000, 8), (7, 4), (3, 8)]:
        spans = [shard_tiles(nt, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == nt
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    with pytest.raises(ValueError):
        shard_waves(0, 4, 0, 2)
    with pytest.raises(ValueError):
        shard_waves(1, 4, 2, 2)


def _worker(rank, world, port, out_path):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    from oracle import oracle_lib
    from volume_path_tracer_b200 import api, scenes
    from volume_path_tracer_b200.distributed import reduce_film, shard_waves
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene, spec = scenes.small_variant("fire", (40, 32), 5, 0.2)
        grids = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
        vol = oracle_lib.Oracle("own").volume(grids)
        film = np.zeros((scene.height, scene.width, 4), np.float32)
        first, count, stride = shard_waves(1, scene.num_waves, rank, world)
        for k in range(count):
            vol.render(scene, film, first_wave=first + k * stride, n_waves=1, threads=2)
        t = torch.from_numpy(film)
        reduce_film(t, dst=0)
        if rank == 0:
            full = np.zeros_like(film)
            vol.render(scene, full, first_wave=1, n_waves=scene.num_waves, threads=2)
            np.savez(out_path, reduced=t.numpy(), full=full)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_wave_sharded_render_equals_single_process(built, tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "films.npz")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    z = np.load(out)
    reduced, full = z["reduced"], z["full"]
    assert np.array_equal(reduced[..., 3], full[..., 3]) and (full[..., 3] == 5).all()
    assert np.allclose(reduced[..., :3], full[..., :3], rtol=2e-6, atol=1e-7)
    assert full[..., :3].sum() > 0
