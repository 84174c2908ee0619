"""The benchmark / parity configurations of BASELINE.json as (Scene, synthetic-volume spec) pairs.

Scene parameters of C1-C3 are the values of the reference's shipped scene files
(scenes/fire.json, scenes/fire_lowscattering.json, scenes/wdas_cloud.json);
C4/C5 are the two synthetic stress configurations BASELINE.json adds (schema-identical). The
reference ships no volumes (`volumes/` is git-ignored), so every configuration renders a seeded
procedural NanoVDB volume (api.synth_volume) placed where the scene's camera looks.
"""
from __future__ import annotations

import json
from dataclasses import dataclass

from .api import Scene, scene_from_json


@dataclass(frozen=True)
class VolumeSpec:
    kind: str
    dims: tuple
    voxel_size: float
    seed: int = 1
    anchor: tuple = (0.0, 0.0, 0.0)

    def scaled(self, factor: float) -> "VolumeSpec":
        """Same world-space extent with `factor` x the resolution (for fast parity cases)."""
        dims = tuple(max(16, int(round(d * factor))) for d in self.dims)
        return VolumeSpec(self.kind, dims, self.voxel_size / factor, self.seed, self.anchor)


def _scene_dict(output_size, inf_xyz, inf_mult, dl_mult, dl_dir, max_depth, sigma_s, sigma_a, g, le_scale, t_off, t_scale, seed, num_waves,
                cam_pos, cam_look, cam_up, vfov, sp_coord):
    return {
        "output_size": list(output_size),
        "worker_parameters": {
            "single_pixel": {"enabled": False, "coord": list(sp_coord)},
            "infinite_light": {"xyz": list(inf_xyz), "multiplier": inf_mult},
            "distant_light": {"xyz": [0.95047, 1.0, 1.08883], "inv_direction": list(dl_dir), "multiplier": dl_mult},
            "use_jitter": True,
            "max_depth": max_depth,
        },
        "volume_parameters": {"sigma_s": sigma_s, "sigma_a": sigma_a, "henyey_greenstein_g": g, "le_scale": le_scale,
                              "temperature_offset": t_off, "temperature_scale": t_scale},
        "seed": seed, "tile_size": [8, 8], "num_waves": num_waves, "num_workers": 12,
        "volume_path": "synthetic",
        "camera_parameters": {"position": list(cam_pos), "look": list(cam_look), "up": list(cam_up), "vfov_deg": vfov, "imaging_ratio": 1e-1},
    }


_FIRE = dict(output_size=(700, 1000), inf_xyz=(0.25, 0.25, 0.5), inf_mult=10, dl_mult=20, dl_dir=(0.5, 1, 0), max_depth=1000000,
             sigma_a=2, g=0.7, le_scale=4e-8, t_off=300.0, t_scale=43.0, seed=500, cam_pos=(120, 30, 0), cam_look=(0, 30, 0),
             cam_up=(0, 1, 0), vfov=37, sp_coord=(239, 879))
_CLOUD = dict(inf_xyz=(4.382, 3.509, 17.603), inf_mult=0.14, dl_mult=50, dl_dir=(0.5826, 0.7660, 0.2717), max_depth=100, sigma_s=0.15,
              sigma_a=0.0, g=0.4, le_scale=0.0, t_off=300.0, t_scale=40.0, seed=10, cam_pos=(648.064, -82.473, -63.856),
              cam_look=(6.021, 100.043, -43.679), cam_up=(0.273, 0.962, -0.009), vfov=35, sp_coord=(550, 450))

_WDAS_ANCHOR = (6.021, 100.043, -43.679)  # the scene's look-at point

CONFIGS = {
    # C1: scenes/fire.json — 700x1000, 1024 waves
    "fire": (_scene_dict(sigma_s=0.9, num_waves=1024, **_FIRE), VolumeSpec("fire", (200, 240, 200), 0.25, 1, (0.0, 0.0, 0.0))),
    # C2: scenes/fire_lowscattering.json — emission dominated
    "fire_lowscattering": (_scene_dict(sigma_s=0.09, num_waves=256, **_FIRE), VolumeSpec("fire", (200, 240, 200), 0.25, 1, (0.0, 0.0, 0.0))),
    # C3: scenes/wdas_cloud.json — 1920x1080, 128 waves, WDAS-shaped sparse cloud (~2k^3 extent)
    "wdas_cloud": (_scene_dict(output_size=(1920, 1080), num_waves=128, **_CLOUD), VolumeSpec("cloud", (2000, 1350, 2450), 0.3, 1, _WDAS_ANCHOR)),
    # C4: dense 512^3 procedural grid, 1920x1080 at 1024 spp. wdas_cloud's lights/camera; sigma_s scaled for a
    # mean free path of ~40 voxels at the mean density (0.5): sigma_t * 0.5 * 40 voxels * voxel_size = 1
    "dense512": (_scene_dict(output_size=(1920, 1080), num_waves=1024, **{**_CLOUD, "sigma_s": 0.0625}),
                 VolumeSpec("dense", (512, 512, 512), 0.8, 1, _WDAS_ANCHOR)),
    # C5: WDAS-shaped, 3840x2160 at 4096 spp, sharded over 2/4/8 GPUs
    "wdas_cloud_4k": (_scene_dict(output_size=(3840, 2160), num_waves=4096, **_CLOUD), VolumeSpec("cloud", (2000, 1350, 2450), 0.3, 1, _WDAS_ANCHOR)),
}


def config_json(name: str) -> str:
    return json.dumps(CONFIGS[name][0], indent=2)


def load_config(name: str):
    """-> (Scene, VolumeSpec). The scene goes through the product's own JSON reader."""
    scene, _ = scene_from_json(config_json(name))
    return scene, CONFIGS[name][1]


def small_variant(name: str, output_size, num_waves: int, volume_factor: float):
    """A scaled-down version of a configuration for parity tests (same camera, lights, medium)."""
    scene, spec = load_config(name)
    scene.output_size[0], scene.output_size[1] = output_size
    scene.num_waves = num_waves
    return scene, spec.scaled(volume_factor)
