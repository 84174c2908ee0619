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
roofline: counted algorithmic bytes (32*N_rho + 32*N_T + 4*N_maj + 32 per pixel-sample, N_* from
        the kernel's own work counters on one step) / mean kernel duration (CUDA events on the launching stream,
        vptc_last_render_ms) against MEASURED_PEAKS.json hbm_gbs.
cpu_baseline: the CPU oracle (restated reference code; on the reference's vendored NanoVDB when oracle/_ref is
        present) with all host threads on a bounded, frame-covering sample of the same workload.
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


def workload_config(args, scene, spec, grids, extra=None):
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
    if extra:
        cfg.update(extra)
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
def cpu_leg(scene, grids, first_wave: int, target_seconds: float, threads: int | None = None, steps: int = 1, warmup: int = 0):
    """Time the CPU oracle on a bounded, frame-covering sample: every `stride`-th (tile, wave) job of one wave
    per step, stride calibrated so one step takes ~target_seconds / steps. Returns dict(value, seconds per step, ...)."""
    import numpy as np
    from oracle import oracle_lib
    flavor = "ref" if oracle_lib.REF_PATH.exists() else "own"
    if flavor == "own" and not oracle_lib.OWN_PATH.exists():
        oracle_lib.build(ref=False)
    orc = oracle_lib.Oracle(flavor)
    t0 = time.time()
    vol = orc.volume(grids)
    setup = time.time() - t0
    threads = threads or os.cpu_count() or 1
    film = np.zeros((scene.height, scene.width, 4), np.float32)
    ntiles = scene.ntiles
    pix_per_job = scene.width * scene.height / ntiles
    def plan(want_jobs):
        stride = max(1, ntiles // max(want_jobs, 1))
        while stride > 1 and ntiles % stride == 0:  # keep it coprime-ish so the sample covers the frame
            stride += 1
        nw = max(1, min(256, want_jobs // ntiles)) if stride == 1 else 1  # fast scenes: several whole waves per step
        return stride, nw
    # calibrate: grow the sample until it runs >= 0.5 s (thread start-up and cold caches dominate tiny runs)
    want_jobs = threads * 32
    while True:
        stride, nw = plan(want_jobs)
        secs, c = vol.render_sample(scene, film, first_wave, nw, stride, threads)
        if secs >= 0.5 or nw >= 256:
            break
        want_jobs *= 6
    rate = c.pixel_samples / max(secs, 1e-6)
    per_step = max(target_seconds / max(steps, 1), 0.5)
    stride, nw = plan(max(threads * 4, int(rate * per_step / pix_per_job)))
    times, samples, counters = [], [], None
    for i in range(warmup + steps):
        secs, c = vol.render_sample(scene, film, first_wave + i * nw, nw, stride, threads)
        if i >= warmup:
            times.append(secs); samples.append(c.pixel_samples); counters = c
    total_s, total_n = sum(times), sum(samples)
    return {"_vol": vol, "value": total_n / total_s, "seconds": total_s, "steps": steps, "samples_per_step": samples[0], "threads": threads, "flavor": flavor,
            "stride": stride, "setup_s": setup, "counters": counters.as_dict(),
            "sample": (f"every {stride}th of the {ntiles} tiles of one wave per step" if stride > 1 else f"{nw} full wave(s) per step")
                      + f" ({samples[0]} pixel-samples/step, {steps} step(s), {total_s:.1f} s)",
            "kind_note": "CPU restatement of the reference integrator (oracle/), " + ("on the reference's vendored NanoVDB + pcg-cpp headers (oracle/_ref)" if flavor == "ref" else "self-contained build")}


def parity_check(renderer, vol, scene, threads: int):
    """BASELINE.json's "RMSE vs reference": wave 1 of the frame rendered by the CPU oracle and through the C ABI on the GPU, same RNG
    streams. One sample per pixel, so the two films must be EQUAL (no add-order slack); reported as counts / RMSE / relMSE / max rel. diff."""
    import numpy as np
    cpu = np.zeros((scene.height, scene.width, 4), np.float32)
    vol.render_sample(scene, cpu, 1, 1, 1, threads)
    gpu = np.zeros_like(cpu)
    renderer.run(scene, gpu, first_wave=1, n_waves=1)
    d = gpu[..., :3].astype(np.float64) - cpu[..., :3].astype(np.float64)
    ref2 = float((cpu[..., :3].astype(np.float64) ** 2).mean())
    return {"against": "CPU oracle, wave 1 of the full frame, identical pcg32_fast streams", "pixels": int(scene.width * scene.height),
            "sample_counts_equal": bool(np.array_equal(cpu[..., 3], gpu[..., 3])), "bit_equal_pixels": float((gpu[..., :3] == cpu[..., :3]).all(axis=-1).mean()),
            "rmse": float(np.sqrt((d ** 2).mean())), "relMSE": float((d ** 2).mean() / max(ref2, 1e-300)),
            "max_rel_diff": float((np.abs(d) / np.maximum(np.abs(cpu[..., :3].astype(np.float64)), 1e-12)).max())}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    scene, spec, grids, _ = load_workload(args)
    r = cpu_leg(scene, grids, 1, target_seconds=min(120.0, 6.0 * args.steps), steps=args.steps, warmup=min(args.warmup, 2))
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * r["seconds"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, scene, spec, grids, {"reference_step": r["sample"]}),
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["threads"], "kind": "port", "sample": r["sample"], "note": r["kind_note"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------- the GPU arm
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
    d_film = torch.zeros((H, W, 4), dtype=torch.float32, device=f"cuda:{local}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # wave addressing: step i, rank r renders waves base + (i*world + r)*M ... + M  (disjoint across ranks and steps)
    def step_device(i):
        first = 1 + (i * world + rank) * M
        r.run_device(scene, d_film.data_ptr(), first, M, 1)
        if world > 1:
            reduce_film(d_film)
        return r.last_render_ms()

    for i in range(Wu):
        step_device(i)
    # one counted step (work counters cost atomics: outside the timed region) for the algorithmic bytes
    r.enable_counters(True)
    r.run_device(scene, d_film.data_ptr(), 1 + (Wu * world + rank) * M, M, 1)
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
    ms = torch.tensor([e0.elapsed_time(e1)], device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    value = px_step * world * K / (total_ms * 1e-3)

    # ---- e2e: the reference-facing call with host buffers (vptc_render), same steps ----
    host_film = np.zeros((H, W, 4), np.float32)
    pinned = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory() if world > 1 else None

    def step_e2e(i):
        first = 1 + (i * world + rank) * M
        if world == 1:
            r.run(scene, host_film, first_wave=first, n_waves=M)  # H2D scene POD, kernel, D2H film, host accumulate
        else:
            d_film.zero_()
            r.run_device(scene, d_film.data_ptr(), first, M, 1)
            reduce_film(d_film)
            if rank == 0:
                pinned.copy_(d_film, non_blocking=False)
                np.add(host_film, pinned.numpy(), out=host_film)
    for i in range(min(Wu, 2)):
        step_e2e(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        step_e2e(Wu + 1 + i)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = px_step * world * K / float(e2e_s.item())
    clk = clocks.stop() if rank == 0 else None

    if rank == 0:
        peak, peak_src = hbm_peak()
        k_ms = sum(kernel_ms) / len(kernel_ms)
        achieved = alg_bytes_per_sample * px_step / (k_ms * 1e-3) / 1e9
        # DRAM traffic and issued warp instructions per pixel-sample come from the committed ncu --set full capture of this
        # kernel on this workload (profiles/r01_ncu_constants.json); scaled to this launch's pixel-samples. None for other configs.
        traffic = issue = None
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_ncu_constants.json")) as f:
                ncu_c = json.load(f).get(args.config)
            if ncu_c:
                traffic = ncu_c["dram_bytes_per_pixel_sample"] * px_step
                sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
                peak_issue = torch.cuda.get_device_properties(local).multi_processor_count * 4 * sm_mhz * 1e6  # 4 schedulers/SM, 1 warp instr/clk each
                ach_issue = ncu_c["warp_instructions_per_pixel_sample"] * px_step / (k_ms * 1e-3)
                issue = {"bound": "warp-instruction issue (the binding limit of this kernel)", "achieved": ach_issue / 1e9, "peak": peak_issue / 1e9,
                         "unit": "G warp-instr/s", "frac": ach_issue / peak_issue, "threads_per_warp_instruction": ncu_c["threads_per_warp_instruction"],
                         "source": "warp instructions per pixel-sample from profiles/r01_ncu_constants.json x live kernel time"}
        except (OSError, ValueError, KeyError):
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wu, "ms_per_step": total_ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, scene, spec, grids, {
                "parallelism": f"wave-sharded x{world}, volume replicated, 1 NCCL film reduce per step" if world > 1 else "single GPU",
                "device_volume_mb": round(r.device_volume_bytes() / 1e6, 1), "synth_s": round(synth_s, 2), "bake_upload_s": round(bake_s, 2)}),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 200, "d2h_bytes_per_step": W * H * 16,
                    "note": "vptc_render: scene POD (200 B) in, float4 film out and accumulated on the host; the volume is staged to HBM once, "
                            "like the reference loads its grids once"},
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
        if not args.no_cpu_baseline and world == 1:
            c = cpu_leg(scene, grids, 1, args.cpu_seconds)
            line["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": c["threads"], "kind": "port", "sample": c["sample"], "note": c["kind_note"]}
            try:
                line["parity"] = parity_check(r, c["_vol"], scene, c["threads"])
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
