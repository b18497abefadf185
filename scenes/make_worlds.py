#!/usr/bin/env python3
"""Writes the benchmark/parity scene files scenes/*.json in the reference's world*.json schema
(keys: /root/reference/src/procedural/cube_world.cc:44-61,64,71-107,117-138,142-143,178-190).

The parameter sets below describe the same worlds as the reference's shipped world1/2/4/8/16,
world8_stress and config files (grid size, materials, lights, camera); tests/test_loader.py checks
that the loader output for each is bit-identical to the reference loader's output for its own
file (tests/golden/scene_dumps.json). Larger `grid*` worlds are stress scenes with no reference
counterpart."""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))

DIRT = {"Ka": [100, 100, 100, 255], "Kd": [210, 105, 0, 30]}
GRASS = {"Ka": [0, 255, 0, 255], "Kd": [124, 252, 0, 255]}
LIGHTS = {"point": [{"pos": [0.0, 10.0, 0.0], "col": [255, 255, 255, 255]}],
          "directional": [{"dir": [0.0, -1.0, 1.0], "col": [128, 128, 128, 255]}]}


def world(grid, cubes, unit_length=200, depth=2, **extra):
    w = {"seed": 42, "grid_size": grid, "amplitude": 4.0, "width": 640, "height": 480, "fov": 45,
         "unit_length": unit_length, "depth": depth, "ambience": [0.5, 0.5, 0.5, 1],
         "distance_attenuation": {"constant_term": 0.1, "linear_term": 0.4, "quadratic_term": 0.8},
         "atlas": "assets/sus.png", "cubes": cubes, "lights": LIGHTS}
    w.update(extra)
    return w


GLASS = dict(GRASS, Kt=[0.6, 0.6, 0.6, 1.0], eta=0.8)
STRESS_GRASS = {"Ka": [60, 120, 0, 255], "Kd": [124, 252, 0, 255]}
MIRROR = {"Ka": [120, 120, 120, 255], "Kr": [0.8, 0.8, 0.8, 1], "Ks": [255, 255, 255, 255], "alpha": 0.7}

WORLDS = {
    "world1": world(1, [DIRT, GLASS]),
    "world2": world(2, [DIRT, GRASS]),
    "world4": world(4, [DIRT, GRASS]),
    "world8": world(8, [DIRT, GRASS]),
    "world16": world(16, [DIRT, GRASS]),
    "world8_stress": world(8, [DIRT, STRESS_GRASS, MIRROR], unit_length=1000),
    "config": world(16, [dict(DIRT, Kd=[210, 105, 0, 255], alpha=0.8), GRASS]),
    # stress content beyond the reference's shipped files (same schema)
    "grid64": world(64, [DIRT, GRASS]),
    "grid64_mixed": world(64, [DIRT, STRESS_GRASS, MIRROR, GLASS], depth=9),
    "grid128": world(128, [DIRT, GRASS]),
    # a large world with real secondary rays that stays benchable and checkable at depth 9: the mirror layer of
    # world8_stress on top of a 64 x 64 terrain (no stacked glass, so none of grid64_mixed's Light::attenuate restarts)
    "grid64_mirror": world(64, [DIRT, STRESS_GRASS, MIRROR], depth=9),
    # analytic sphere hitables (schema extension "spheres"): matte, mirror and glass spheres over world8
    "spheres8": world(8, [DIRT, GRASS], depth=5, spheres=[
        dict(radius=1.5, pos=[-2.0, 7.0, 1.0], Ka=[120, 30, 30, 255], Kd=[220, 40, 40, 255], Ks=[255, 255, 255, 255], alpha=20.0),
        dict(MIRROR, radius=2.0, pos=[2.5, 8.0, 3.0]),
        dict(radius=1.0, pos=[0.0, 9.0, -1.0], Ka=[20, 20, 40, 255], Kd=[30, 30, 60, 255], Kt=[0.8, 0.8, 0.9, 1.0], eta=0.8),
        dict(radius=0.75, pos=[-3.25, 4.5, -2.0], Ka=[40, 40, 120, 255], Kd=[60, 60, 230, 255]),
        dict(radius=0.3, pos=[1.0, 10.5, -3.0], Ke=[255, 220, 120, 255]),
    ]),
}

def write_atlas(path, tiles=4, tile=16):
    """Deterministic atlas of tiles x tiles square tiles as a binary PPM: every tile has its own base colour, a darker
    checker pattern, a bright border and one marked corner, so that orientation and clamping errors show in a frame."""
    n = tiles * tile
    rows = bytearray()
    for y in range(n):
        for x in range(n):
            tx, ty, lx, ly = x // tile, y // tile, x % tile, y % tile
            base = (40 + 60 * tx, 230 - 50 * ty, 60 + 45 * ((tx + 2 * ty) % 4))
            k = 1.0 if ((lx // 4) + (ly // 4)) % 2 == 0 else 0.55
            c = [int(v * k) for v in base]
            if lx in (0, tile - 1) or ly in (0, tile - 1):
                c = [250, 250, 250]
            if lx < 4 and ly < 4:
                c = [255, 0, 0]
            rows += bytes(c)
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (n, n))
        f.write(bytes(rows))


WORLDS["textured8"] = dict(world(8, [dict(DIRT, texture=[0, 0, 16]), dict(GRASS, texture=[16, 32, 16])], spheres=[
    dict(radius=1.5, pos=[-2.0, 7.0, 1.0], Ka=[120, 30, 30, 255], Kd=[220, 40, 40, 255], Ks=[255, 255, 255, 255], alpha=20.0),
]), atlas_ppm="atlas64.ppm")

if __name__ == "__main__":
    write_atlas(os.path.join(HERE, "atlas64.ppm"))
    for name, w in WORLDS.items():
        with open(os.path.join(HERE, name + ".json"), "w") as f:
            json.dump(w, f, indent=1)
            f.write("\n")
        print("wrote", name)
