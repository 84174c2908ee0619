#!/usr/bin/env python
"""bench.py — pixel-samples/s of the volumetric path integrator on N B200s, beside the host-CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config NAME] [--waves-per-step M] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A STEP is one pass of the hot path (vpt::run, src/worker.cpp:92-208) over one batch of synthetic input:
`waves_per_step` waves (samples per pixel) of the full frame of the workload, i.e. ntiles * waves_per_step
(tile, wave) jobs with the reference's job ids; successive steps render successive waves, like the
reference's progressive renderer. Default workload = BASELINE.json configs[3], the largest single-GPU
configuration: the dense 512^3 procedural density grid at 1920x1080 (bricks 772 MB + tables, > 126 MB of L2).
N > 1: weak scaling — every rank renders its own `waves_per_step` waves per step (wave-sharded job space,
volume replicated) and the float4 films are summed onto rank 0 with one NCCL reduce per step.

value : whole-job pixel-samples/s, film resident in HBM (vptc_render_device), timed with CUDA events over
        exactly K steps between barrier + synchronize pairs, max over ranks.
e2e   : the same metric through the reference-facing call with HOST buffers (Renderer.run = vptc_render: scene
        POD in, kernel, float4 film device->pinned host, accumulated into the caller's host film).
e2e_small_calls: the same host-film job handed over 16 waves per call through vptc_render_async (device-side range ring: no drain between calls).
strong (N >= 1): fixed total jobs (BASELINE configs 3 and 5) split over the N ranks, end to end, with the N = 1 time measured in the same run.
parity: N = 1: wave 1 of the frame, GPU vs the CPU reference, equal bit for bit; N > 1: the NCCL-reduced film vs rank 0 rendering the same waves alone.
roofline: counted algorithmic bytes (32*N_rho + 32*N_T + 4*N_maj + 32 per pixel-sample, N_* from
        the kernel's own work counters on one step) / mean kernel duration (CUDA events on the launching stream,
        vptc_last_render_ms) against MEASURED_PEAKS.json hbm_gbs.
cpu_baseline: the reference's own CPU code for the path (its src/*.cpp compiled unchanged into oracle/_ref/libvpt_reference.so; kind
        "reference") — or the oracle's restatement where that library is absent (kind "port") — on all host threads, whole waves of the frame.
--impl reference: times that CPU implementation alone, same metric/config, all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC, UNIT = "pixel_samples_per_sec", "pixel-samples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="dense512", help="fire | fire_lowscattering | wdas_cloud | dense512 | wdas_cloud_4k")
    ap.add_argument("--waves-per-step", type=int, default=128)
    ap.add_argument("--volume-scale", type=float, default=1.0, help="resolution factor of the synthetic volume (1 = the named shape)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling block (fixed total jobs of BASELINE configs 3 and 5 split over the N GPUs)")
    ap.add_argument("--strong-waves-4k", type=int, default=512, help="waves of the 3840x2160 job in the strong block (BASELINE config 5 asks 4096 spp: 165 s on one GPU)")
    ap.add_argument("--strong-volume-scale", type=float, default=1.0, help="resolution factor of the cloud volume of the strong block")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


def load_workload(args, rank=0, world=1, barrier=None):
    """Synthesise the named procedural volume. With several ranks, rank 0 synthesises once and the others read the
    `.nvdb` it wrote (the product's own writer/reader), instead of every rank burning the host cores on the same work."""
    from volume_path_tracer_b200 import api, scenes
    scene, spec = scenes.load_config(args.config)
    if args.volume_scale != 1.0:
        spec = spec.scaled(args.volume_scale)
    t = time.time()
    if world == 1:
        grids = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
    else:
        shm = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
        path = shm / f"vpt_bench_{os.environ.get('MASTER_PORT', '0')}_{args.config}_{args.volume_scale}.nvdb"
        if rank == 0:
            api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor).write(path)
        barrier()
        grids = api.VolumeGrids.read_from_file(path)
        barrier()
        if rank == 0:
            path.unlink(missing_ok=True)
    return scene, spec, grids, time.time() - t


def workload_config(args, scene, spec, grids):
    """The workload, described identically by both arms (the driver compares the two `config` objects)."""
    cfg = {
        "workload": f"{args.config}: synthetic {spec.kind} volume {spec.dims[0]}x{spec.dims[1]}x{spec.dims[2]} voxels (seed {spec.seed}), "
                    f"{scene.width}x{scene.height}, {scene.num_waves} spp job rendered in steps of {args.waves_per_step} waves",
        "waves_per_step": args.waves_per_step,
        "pixel_samples_per_step_per_gpu": scene.width * scene.height * args.waves_per_step,
        "tile_size": [int(scene.tile_size[0]), int(scene.tile_size[1])],
        "jobs_per_step_per_gpu": scene.ntiles * args.waves_per_step,
        "nanovdb_grid_mb": round((grids.density.nbytes + (grids.temperature.nbytes if grids.temperature else 0)) / 1e6, 1),
        "l2": "inputs larger than L2 (no flush)" if grids.density.nbytes > 200e6 else "volume fits in L2 (hot-L2 number; no flush)",
        "rng": "reference streams: pcg32_fast per (tile, wave) job, pixels serial within a job",
    }
    return cfg


# ---------------------------------------------------------------- clocks (B200_PROFILING.md recipe)
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.idx, self.rows, self.proc = device_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.idx)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); smax.append(float(r[2])); power.append(float(r[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "power_w_max": max(power), "samples": len(sm), "reasons": sorted(reasons)}


def hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------- CPU legs (the ONLY places bench.py touches oracle/)
def cpu_leg(scene, grids, first_wave: int, target_seconds: float, threads: int | None = None, steps: int = 3, warmup: int = 1):
    """Time the reference's CPU implementation of the path on the host cores, on a bounded sample of the workload: WHOLE waves of the full
    frame (every tile, `nw` consecutive waves per step), `steps` timed steps after `warmup`, sized so the leg takes ~target_seconds.
    Implementation, in order of preference:
      "reference": oracle/_ref/libvpt_reference.so — the reference's own src/*.cpp compiled unchanged (vpt::run on num_workers jthreads
                   over its TileProvider, src/main.cpp:63-87);
      "port":      the oracle's restatement (oracle/vpt_restated*.hpp), when that library is absent.
    Returns dict(value = pixel-samples/s over the timed steps, ...)."""
    import numpy as np
    from oracle import oracle_lib
    threads = threads or os.cpu_count() or 1
    t0 = time.time()
    if oracle_lib.REFSRC_PATH.exists():
        kind, vol = "reference", oracle_lib.Reference().volume(grids)
        note = "the reference's own translation units (worker/volume/majorant_transmittance_sampler/camera/tile_provider/... .cpp) compiled unchanged into oracle/_ref/libvpt_reference.so"
        def run(first, nw):
            return vol.render(scene, film, first_wave=first, n_waves=nw, threads=threads)
        checker = None
    else:
        flavor = "ref" if oracle_lib.REF_PATH.exists() else "own"
        if flavor == "own" and not oracle_lib.OWN_PATH.exists():
            oracle_lib.build(ref=False)
        kind, vol = "port", oracle_lib.Oracle(flavor).volume(grids)
        note = "CPU restatement of the reference integrator (oracle/), " + ("on the reference's vendored NanoVDB + pcg-cpp headers" if flavor == "ref" else "self-contained build")
        def run(first, nw):
            return vol.render_sample(scene, film, first, nw, 1, threads)[0]
    setup = time.time() - t0
    film = np.zeros((scene.height, scene.width, 4), np.float32)
    px_wave = scene.width * scene.height
    secs = run(first_wave, 1)  # calibration wave (also warms caches and the thread pool's pages)
    per_step = max(target_seconds / max(steps + warmup, 1), 1.0)
    nw = int(max(1, min(64, round(per_step / max(secs, 1e-3)))))
    times = []
    for i in range(warmup + steps):
        s = run(first_wave + 1 + i * nw, nw)
        if i >= warmup:
            times.append(s)
    total_s = sum(times)
    return {"value": px_wave * nw * steps / total_s, "seconds": total_s, "steps": steps, "waves_per_step": nw, "threads": threads, "kind": kind, "setup_s": setup,
            "sample": f"{nw} full wave(s) of the {scene.width}x{scene.height} frame per step ({px_wave * nw} pixel-samples/step), {steps} timed step(s) after {warmup} warm-up, {total_s:.1f} s",
            "note": note}


def checker_volume(grids):
    """The CPU checker for the parity blocks: the compiled reference when present (films only), else the restated oracle."""
    from oracle import oracle_lib
    if oracle_lib.REFSRC_PATH.exists():
        v = oracle_lib.Reference().volume(grids)
        return "the reference's own compiled translation units (oracle/_ref/libvpt_reference.so)", lambda scene, film, first, n, threads: v.render(scene, film, first_wave=first, n_waves=n, threads=threads)
    flavor = "ref" if oracle_lib.REF_PATH.exists() else "own"
    if flavor == "own" and not oracle_lib.OWN_PATH.exists():
        oracle_lib.build(ref=False)
    v = oracle_lib.Oracle(flavor).volume(grids)
    return "the CPU oracle (restated reference)", lambda scene, film, first, n, threads: v.render_sample(scene, film, first, n, 1, threads)


def film_stats(got, want):
    import numpy as np
    d = got[..., :3].astype(np.float64) - want[..., :3].astype(np.float64)
    ref2 = float((want[..., :3].astype(np.float64) ** 2).mean())
    return {"sample_counts_equal": bool(np.array_equal(want[..., 3], got[..., 3])), "bit_equal_pixels": float((got[..., :3] == want[..., :3]).all(axis=-1).mean()),
            "rmse": float(np.sqrt((d ** 2).mean())), "relMSE": float((d ** 2).mean() / max(ref2, 1e-300)),
            "max_rel_diff": float((np.abs(d) / np.maximum(np.abs(want[..., :3].astype(np.float64)), 1e-12)).max())}


def parity_check(renderer, grids, scene, threads: int):
    """BASELINE.json's "RMSE vs reference": wave 1 of the full frame rendered on the CPU and through the C ABI on the GPU, same RNG
    streams. One sample per pixel, so the two films must be EQUAL (no add-order slack); reported as counts / RMSE / relMSE / max rel. diff."""
    import numpy as np
    who, render = checker_volume(grids)
    cpu = np.zeros((scene.height, scene.width, 4), np.float32)
    render(scene, cpu, 1, 1, threads)
    gpu = np.zeros_like(cpu)
    renderer.run(scene, gpu, first_wave=1, n_waves=1)
    return {"against": who + ", wave 1 of the full frame, identical pcg32_fast streams", "pixels": int(scene.width * scene.height), **film_stats(gpu, cpu)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    scene, spec, grids, _ = load_workload(args)
    r = cpu_leg(scene, grids, 1, target_seconds=min(150.0, 8.0 * (args.steps + min(args.warmup, 2))), steps=args.steps, warmup=min(args.warmup, 2))
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, scene, spec, grids),
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["threads"], "kind": r["kind"], "sample": r["sample"], "note": r["note"]},
        "reference_step": r["sample"],
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------- the GPU arm
def ncu_constants(config):
    """DRAM traffic and issued warp instructions per pixel-sample from the committed ncu captures of this kernel on this workload
    (profiles/r02_ncu_constants.json, else round 1's); None for other configs."""
    for name in ("r02_ncu_constants.json", "r01_ncu_constants.json"):
        try:
            with open(ROOT / "profiles" / name) as f:
                c = json.load(f).get(config)
            if c:
                return c, name
        except (OSError, ValueError):
            pass
    return None, None


def strong_scaling(args, r, rank, world, local, barrier):
    """A FIXED total job split over the N ranks, timed end to end (render + film combine + rank-0 device->host read), beside the same job on
    rank 0 alone, so the line carries its own efficiency: BASELINE config 3 (wdas_cloud, its 128 spp, wave-sharded + NCCL reduce) and
    config 5 (wdas_cloud_4k, tile-stripe-sharded + NCCL gather; --strong-waves-4k of its 4096 spp)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from volume_path_tracer_b200 import api, scenes
    from volume_path_tracer_b200.distributed import gather_film_stripes, reduce_film, shard_tile_rows, shard_waves
    dev = f"cuda:{local}"
    sub = argparse.Namespace(**{**vars(args), "config": "wdas_cloud", "volume_scale": args.strong_volume_scale})
    _, spec, grids, synth_s = load_workload(sub, rank, world, lambda: dist.barrier() if world > 1 else None)
    t0 = time.time()
    r.set_volume(grids)
    bake_s = time.time() - t0
    out = {"volume": f"synthetic cloud {spec.dims[0]}x{spec.dims[1]}x{spec.dims[2]} voxels, {grids.density.nbytes / 1e6:.0f} MB NanoVDB grid replicated per GPU",
           "synth_s": round(synth_s, 2), "bake_upload_s": round(bake_s, 2), "jobs": []}
    for cfg, total, mode in (("wdas_cloud", None, "waves"), ("wdas_cloud_4k", args.strong_waves_4k, "tiles")):
        scene, _ = scenes.load_config(cfg)
        total = int(total or scene.num_waves)
        H, W = scene.height, scene.width
        film = torch.zeros((H, W, 4), dtype=torch.float32, device=dev)
        pinned = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory() if rank == 0 else None
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

        def job(n_ranks, me):
            """This rank's share when the job is split over n_ranks (me < n_ranks), then the combine onto rank 0 and the host read."""
            film.zero_()
            coll_ms = 0.0
            if mode == "waves":
                first, count, stride = shard_waves(1, total, me, n_ranks)
                if count:
                    r.run_device(scene, film.data_ptr(), first, count, stride)
            else:
                tb, tc, stride, _ = shard_tile_rows(scene, me, n_ranks)
                if tc:
                    r.run_device_range(scene, film.data_ptr(), 1, total, 1, tile_begin=tb, tile_count=tc, tile_row_stride=stride)
            if n_ranks > 1:
                ev[0].record()
                if mode == "waves":
                    reduce_film(film)
                else:
                    gather_film_stripes(film, scene, me, n_ranks)
                ev[1].record()
                torch.cuda.synchronize()
                coll_ms = ev[0].elapsed_time(ev[1])
            if me == 0:
                pinned.copy_(film, non_blocking=False)
            return coll_ms

        r.run_device(scene, film.data_ptr(), 1, 2, 1)  # warm-up of this frame size
        if world > 1:  # ... and of the combine: NCCL sets up its point-to-point connections at the first gather (1.6 s at N = 8), like the W warm-up steps of the main loop
            if mode == "waves":
                reduce_film(film)
            else:
                gather_film_stripes(film, scene, rank, world)
            torch.cuda.synchronize()
        # the whole job on rank 0 alone
        barrier()
        t1 = None
        if rank == 0:
            t = time.perf_counter(); job(1, 0); torch.cuda.synchronize(); t1 = time.perf_counter() - t
            counts_ok = bool((pinned[..., 3] == total).all())
            ref_sum = pinned[..., :3].double().sum().item()
        barrier()
        # ... and split over all ranks
        if world > 1:
            t = time.perf_counter(); coll_ms = job(world, rank); barrier(); tn = time.perf_counter() - t
        else:
            tn, coll_ms = t1, 0.0
        tt = torch.tensor([tn, coll_ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        if rank == 0:
            tn, coll_ms = float(tt[0]), float(tt[1])
            px = W * H * total
            out["jobs"].append({"config": cfg, "frame": [W, H], "total_waves": total, "sharding": "waves + NCCL reduce" if mode == "waves" else "interleaved rows of tiles + NCCL gather",
                                "pixel_samples": px, "t_n1_s": t1, "t_s": tn, "value_n1": px / t1, "value": px / tn, "speedup_vs_n1": t1 / tn, "efficiency_vs_n1": t1 / (tn * world),
                                "collective_ms": coll_ms, "collective_share": coll_ms * 1e-3 / tn, "waves_per_gpu": total / world if mode == "waves" else total,
                                "sample_counts_equal": counts_ok and bool((pinned[..., 3] == total).all()),
                                "film_sum_rel_diff_vs_n1": abs(pinned[..., :3].double().sum().item() - ref_sum) / max(abs(ref_sum), 1e-30)})
        del film
    return out


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from volume_path_tracer_b200 import api
    from volume_path_tracer_b200.distributed import reduce_film

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU baseline)")
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch N>1 with torch.distributed.run")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def host_barrier():
        if world > 1:
            dist.barrier()
    scene, spec, grids, synth_s = load_workload(args, rank, world, host_barrier)
    K, Wu, M = args.steps, args.warmup, args.waves_per_step
    W, H = scene.width, scene.height
    px_step = W * H * M

    r = api.Renderer(local)  # raises if libvptcuda.so is missing or no device: no fallback
    t0 = time.time()
    r.set_volume(grids)
    bake_s = time.time() - t0
    device_volume_mb = round(r.device_volume_bytes() / 1e6, 1)
    d_film = torch.zeros((H, W, 4), dtype=torch.float32, device=dev)           # the accumulated film (meaningful on rank 0)
    d_step = torch.zeros_like(d_film) if world > 1 else None                     # this step's partial film, reduced onto rank 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # wave addressing: step i, rank r renders waves 1 + (i*world + r)*M ... + M  (disjoint across ranks and steps)
    def step_device(i):
        first = 1 + (i * world + rank) * M
        if world == 1:
            r.run_device(scene, d_film.data_ptr(), first, M, 1)
        else:  # reduce THIS STEP's partial films, then accumulate on rank 0 (reducing the accumulated film would re-add earlier steps)
            d_step.zero_()
            r.run_device(scene, d_step.data_ptr(), first, M, 1)
            reduce_film(d_step)
            if rank == 0:
                d_film.add_(d_step)
        return r.last_render_ms()

    for i in range(Wu):
        step_device(i)
    # one counted step (work counters cost atomics: outside the timed region) for the algorithmic bytes
    r.enable_counters(True)
    r.run_device(scene, (d_step if world > 1 else d_film).data_ptr(), 1 + (Wu * world + rank) * M, M, 1)
    cnt = r.counters()
    r.enable_counters(False)
    alg_bytes_per_sample = cnt.algorithmic_bytes() / max(cnt.pixel_samples, 1)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    d_film.zero_()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    kernel_ms = []
    for i in range(K):
        kernel_ms.append(step_device(Wu + 1 + i))
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    value = px_step * world * K / (total_ms * 1e-3)
    counts_ok = bool((d_film[..., 3] == K * M * world).all()) if rank == 0 else True  # every timed sample arrived exactly once

    # ---- e2e: the reference-facing call with HOST buffers, same steps ----
    host_film = np.zeros((H, W, 4), np.float32)
    pinned = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory() if world > 1 else None

    def step_e2e(i):
        first = 1 + (i * world + rank) * M
        if world == 1:
            r.run(scene, host_film, first_wave=first, n_waves=M)  # H2D scene POD, kernel, D2H film, host accumulate
        else:
            d_step.zero_()
            r.run_device(scene, d_step.data_ptr(), first, M, 1)
            reduce_film(d_step)
            if rank == 0:
                pinned.copy_(d_step, non_blocking=False)
                np.add(host_film, pinned.numpy(), out=host_film)
    for i in range(min(Wu, 2)):
        step_e2e(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        step_e2e(Wu + 1 + i)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = px_step * world * K / float(e2e_s.item())

    # ---- the same host-film job handed over in SMALL calls (16 waves each, asynchronous): what a progressive renderer does ----
    small = None
    if world == 1:
        host_film[:] = 0
        per = min(16, M)
        calls = [(1 + Wu * M + i, min(per, K * M - i)) for i in range(0, K * M, per)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for f, n in calls:
            r.run_async(scene, host_film, f, n)
        r.wait()
        dt = time.perf_counter() - t0
        small = {"value": px_step * K / dt, "unit": UNIT, "waves_per_call": per, "calls": len(calls),
                 "sample_counts_equal": bool((host_film[..., 3] == K * M).all()),
                 "note": "vptc_render_async x calls + vptc_wait: each call's range is published to the device ring and the running kernel carries on into it"}
    clk = clocks.stop() if rank == 0 else None

    # ---- N > 1: the NCCL-reduced film against rank 0 rendering the same waves alone (outside the timed region) ----
    multi_parity = None
    if world > 1:
        Mp = 4
        d_step.zero_()
        r.run_device(scene, d_step.data_ptr(), 1 + rank * Mp, Mp, 1)
        reduce_film(d_step)
        if rank == 0:
            alone = torch.zeros_like(d_step)
            r.run_device(scene, alone.data_ptr(), 1, Mp * world, 1)
            torch.cuda.synchronize()
            a, b = d_step.cpu().numpy(), alone.cpu().numpy()
            multi_parity = {"against": f"rank 0 alone rendering the same {Mp * world} waves through vptc_render_device", "waves": Mp * world,
                            "allclose_rtol_1e-5": bool(np.allclose(a, b, rtol=1e-5, atol=1e-7)), **film_stats(a, b)}
            del alone
        barrier()

    strong = None
    if not args.no_strong:
        try:
            strong = strong_scaling(args, r, rank, world, local, barrier)
        except Exception as e:  # the block must never cost the bench line
            strong = {"error": repr(e)}
            if world > 1:
                raise

    if rank == 0:
        peak, peak_src = hbm_peak()
        k_ms = sum(kernel_ms) / len(kernel_ms)
        achieved = alg_bytes_per_sample * px_step / (k_ms * 1e-3) / 1e9
        traffic = issue = None
        ncu_c, ncu_file = ncu_constants(args.config)
        if ncu_c:
            traffic = ncu_c["dram_bytes_per_pixel_sample"] * px_step
            sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
            peak_issue = torch.cuda.get_device_properties(local).multi_processor_count * 4 * sm_mhz * 1e6  # 4 schedulers/SM, 1 warp instr/clk each
            ach_issue = ncu_c["warp_instructions_per_pixel_sample"] * px_step / (k_ms * 1e-3)
            lanes = ncu_c["threads_per_warp_instruction"]
            issue = {"bound": "warp-instruction issue (the binding limit of this kernel)", "achieved": ach_issue / 1e9, "peak": peak_issue / 1e9,
                     "unit": "G warp-instr/s", "frac": ach_issue / peak_issue, "threads_per_warp_instruction": lanes,
                     "useful": ach_issue / peak_issue * lanes / 32.0,  # share of the SMs' THREAD-instruction rate doing work: frac x lanes / 32
                     "source": f"warp instructions per pixel-sample from profiles/{ncu_file} x live kernel time"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wu, "ms_per_step": total_ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, scene, spec, grids),
            "run": {"parallelism": f"wave-sharded x{world}, volume replicated, 1 NCCL film reduce per step" if world > 1 else "single GPU",
                    "device_volume_mb": device_volume_mb, "synth_s": round(synth_s, 2), "bake_upload_s": round(bake_s, 2),
                    "timed_sample_counts_exact": counts_ok},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 200 + W * H * 16, "d2h_bytes_per_step": W * H * 16,
                    "note": "vptc_render: scene POD (200 B) in, this call's device film cleared by a copy-engine upload of zeros, float4 film out and accumulated "
                            "on the host; the volume is staged to HBM once, like the reference loads its grids once"},
            "gpu_launches": K,
            "kernel": "vpt_render_sched_kernel (variant 2; 1 launch per step per GPU)",
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_src, "kernel_ms": k_ms,
                         "algorithmic_bytes_per_pixel_sample": alg_bytes_per_sample,
                         "per_pixel_sample": {"density_evals": cnt.density_evals / cnt.pixel_samples, "temperature_evals": cnt.temperature_evals / cnt.pixel_samples,
                                              "majorant_fetches": cnt.majorant_fetches / cnt.pixel_samples, "rng_draws": cnt.rng_draws / cnt.pixel_samples,
                                              "scatter_events": cnt.scatter_events / cnt.pixel_samples}},
            "clocks": clk,
        }
        if issue:
            line["roofline"]["issue"] = issue
        if small:
            line["e2e_small_calls"] = small
        if multi_parity:
            line["parity"] = multi_parity
        if strong:
            line["strong"] = strong
        if world == 1:
            r.set_volume(grids)  # the strong block staged the cloud: back to this workload for the parity check
            if not args.no_cpu_baseline:
                c = cpu_leg(scene, grids, 1, args.cpu_seconds)
                line["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": c["threads"], "kind": c["kind"], "sample": c["sample"], "note": c["note"]}
            try:
                line["parity"] = parity_check(r, grids, scene, os.cpu_count() or 1)
            except Exception as e:  # the check must never cost the bench line
                line["parity"] = {"error": repr(e)}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    r.close()
    return 0


_RESULT_FD = None


def emit(line: dict):
    """The ONE JSON line of the contract, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    # Libraries write to fd 1 behind Python's back (NCCL prints "NCCL version ..." there when the box sets NCCL_DEBUG, and its
    # INFO log too): keep stdout for the result line alone by pointing fd 1 at stderr for the rest of the run.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
