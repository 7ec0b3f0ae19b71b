"""Synthetic scenes of the benchmark configs.  The reference ships no scene file (`*.toml` is
git-ignored, .gitignore:7-9) and has no generator; these emit scenes in its TOML schema through
Scene.to_toml(), one `[materials.m<i>]` per sphere."""
from __future__ import annotations

import numpy as np

from .scene import Material, Scene, Sphere


def three_spheres() -> Scene:
    """S1 (config C1): ground + diffuse / glass / metal unit spheres."""
    mats = {
        "ground": Material.lambertian((0.5, 0.5, 0.5)),
        "diffuse": Material.lambertian((0.4, 0.2, 0.1)),
        "metal": Material.metal((0.7, 0.6, 0.5), 0.0),
        "glass": Material.dielectric(1.5, False),
    }
    objs = [
        Sphere((0.0, -1000.0, 0.0), 1000.0, "ground"),
        Sphere((-4.0, 1.0, 0.0), 1.0, "diffuse"),
        Sphere((4.0, 1.0, 0.0), 1.0, "metal"),
        Sphere((0.0, 1.0, 0.0), 1.0, "glass"),
    ]
    return Scene.from_list(objs, mats)


def _grid_scene(half: int, seed: int, p_lambert: float, p_metal: float, spacing: float = 1.0, big: bool = True) -> Scene:
    rng = np.random.default_rng(seed)
    mats = {"ground": Material.lambertian((0.5, 0.5, 0.5))}
    objs = [Sphere((0.0, -1000.0, 0.0), 1000.0, "ground")]
    k = 0
    for a in range(-half, half):
        for b in range(-half, half):
            choose = rng.random()
            cx = a * spacing + 0.9 * rng.random()
            cz = b * spacing + 0.9 * rng.random()
            u3 = rng.random(3)
            u4 = rng.random(4)
            if big and np.sqrt((cx - 4.0) ** 2 + (cz - 0.0) ** 2) <= 0.9:
                continue
            name = f"m{k}"
            k += 1
            if choose < p_lambert:
                mats[name] = Material.lambertian(tuple(float(x) for x in (u3 * u3)))
            elif choose < p_lambert + p_metal:
                mats[name] = Material.metal(tuple(float(x) for x in (0.5 + 0.5 * u4[:3])), float(0.5 * u4[3]))
            else:
                mats[name] = Material.dielectric(1.5, False)
            objs.append(Sphere((float(cx), 0.2, float(cz)), 0.2, name))
    if big:
        mats["big_glass"] = Material.dielectric(1.5, False)
        mats["big_diffuse"] = Material.lambertian((0.4, 0.2, 0.1))
        mats["big_metal"] = Material.metal((0.7, 0.6, 0.5), 0.0)
        objs.append(Sphere((0.0, 1.0, 0.0), 1.0, "big_glass"))
        objs.append(Sphere((-4.0, 1.0, 0.0), 1.0, "big_diffuse"))
        objs.append(Sphere((4.0, 1.0, 0.0), 1.0, "big_metal"))
    return Scene.from_list(objs, mats)


def random_spheres(seed: int = 20240607) -> Scene:
    """S2 (configs C2/C3/T): RTIOW-style 22x22 grid of r=0.2 spheres (80 % lambertian / 15 % metal / 5 % glass),
    a ground sphere and three big spheres: about 485 spheres."""
    return _grid_scene(11, seed, 0.80, 0.15)


def ten_k_spheres(seed: int = 20240608) -> Scene:
    """S3 (config C4): 100x100 jittered grid, same material mix, + ground = 10 001 spheres."""
    return _grid_scene(50, seed, 0.80, 0.15, big=False)


def forty_k_spheres(seed: int = 20240610) -> Scene:
    """Beyond the BASELINE configs: 200x200 jittered grid + ground = 40 001 spheres, the scene size at which the stage-1
    data no longer fit in shared memory at once (the streamed trace kernel)."""
    return _grid_scene(100, seed, 0.80, 0.15, big=False)


def dielectric_heavy(seed: int = 20240609) -> Scene:
    """S4 (config C5): S2 layout with 10 % lambertian / 10 % metal / 80 % dielectric."""
    return _grid_scene(11, seed, 0.10, 0.10)
