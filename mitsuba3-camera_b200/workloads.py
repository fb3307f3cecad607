"""Deterministic synthetic workloads of BASELINE.json's configs as scene dictionaries for
mi.load_dict. No datasets: every mesh / lens is generated from a seed."""
from __future__ import annotations

import numpy as np

try:
    from .transform import Transform4f
except ImportError:      # loaded by path (tools/ref_bench.py: the reference arm must not import the product package)
    from transform import Transform4f

f32 = np.float32


def _value_noise(n, rng, octaves=5, base=6):
    """[n,n] multi-octave bilinear value noise in [-1,1]."""
    out = np.zeros((n, n), np.float64)
    amp, tot = 1.0, 0.0
    xs = np.linspace(0, 1, n)
    for o in range(octaves):
        k = base * 2 ** o
        lat = rng.random((k + 1, k + 1))
        g = xs * k
        i0 = np.minimum(g.astype(np.int64), k - 1)
        f = g - i0
        f = f * f * (3 - 2 * f)
        a = lat[i0][:, i0] * np.outer(1 - f, 1 - f) + lat[i0 + 1][:, i0] * np.outer(f, 1 - f) \
            + lat[i0][:, i0 + 1] * np.outer(1 - f, f) + lat[i0 + 1][:, i0 + 1] * np.outer(f, f)
        out += amp * (2 * a - 1)
        tot += amp
        amp *= 0.5
    return out / tot


def heightfield_scene(n_side=708, width=3840, height=2160, spp=16, lens_design="gauss", max_depth=6, seed=1,
                      extent=10.0, amplitude=1.2, patch=16, n_diffuse=8, sensor="realisticlens", sensor_extra=None):
    """Config 5: ~1M-triangle displaced grid (n_side=708 -> 999,698 triangles), 50 % diffuse (albedo U[0.2,0.8]) /
    50 % conductor (Cu) in square patches, one large area light + 4 emissive quads, realistic lens."""
    rng = np.random.default_rng(seed)
    n = int(n_side)
    h = _value_noise(n, rng) * amplitude
    xs = np.linspace(-extent, extent, n)
    X, Z = np.meshgrid(xs, xs, indexing="xy")
    verts = np.stack([X, h, Z], -1).reshape(-1, 3).astype(f32)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="xy")
    v00 = (j * n + i).reshape(-1)
    v10, v01, v11 = v00 + 1, v00 + n, v00 + n + 1
    # counter-clockwise seen from +y so that face normals point up
    faces = np.stack([np.stack([v00, v01, v11], 1), np.stack([v00, v11, v10], 1)], 1).reshape(-1, 3).astype(np.int32)
    cell_patch = ((j // patch) * ((n - 1 + patch - 1) // patch) + (i // patch)).reshape(-1)
    n_patches = int(cell_patch.max()) + 1
    patch_mat = rng.integers(0, 2 * n_diffuse, n_patches)       # < n_diffuse: diffuse k ; else conductor
    cell_mat = np.where(patch_mat[cell_patch] < n_diffuse, patch_mat[cell_patch], n_diffuse)
    tri_mat = np.repeat(cell_mat, 2)
    albedo = rng.uniform(0.2, 0.8, (n_diffuse, 3))
    T = Transform4f
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": max_depth, "rr_depth": 5}}
    for m in range(n_diffuse + 1):
        sel = faces[tri_mat == m]
        if len(sel) == 0:
            continue
        used, inv = np.unique(sel.reshape(-1), return_inverse=True)
        bsdf = ({"type": "diffuse", "reflectance": {"type": "rgb", "value": albedo[m].tolist()}} if m < n_diffuse
                else {"type": "conductor", "material": "Cu"})
        d[f"field{m}"] = {"type": "mesh", "vertices": verts[used], "faces": inv.reshape(-1, 3).astype(np.int32), "bsdf": bsdf}
    black = {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.0, 0.0, 0.0]}}
    d["sun"] = {"type": "rectangle", "to_world": T.translate([0, 9.0, 0]).rotate([1, 0, 0], 90).scale([5.0, 5.0, 1.0]),
                "bsdf": black, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [6.0, 5.5, 5.0]}}}
    quads = [([-7, 4.0, -7], [9.0, 3.0, 1.0]), ([7, 4.5, -6], [1.0, 6.0, 9.0]), ([-6, 5.0, 6], [2.0, 8.0, 3.0]), ([6, 3.5, 7], [8.0, 7.0, 1.0])]
    for q, (pos, rad) in enumerate(quads):
        d[f"quad{q}"] = {"type": "rectangle", "to_world": T.translate(pos).rotate([1, 0, 0], 90).scale([0.8, 0.8, 1.0]),
                         "bsdf": black, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": rad}}}
    origin, target = [0.0, 6.5, 15.0], [0.0, 0.0, 0.0]
    tw = T.look_at(origin=origin, target=target, up=[0, 1, 0])
    film = {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "gaussian"}, "pixel_format": "rgb"}
    smp = {"type": "independent", "sample_count": spp}
    if sensor == "realisticlens":
        d["sensor"] = {"type": "realisticlens", "lens_design": lens_design, "object_distance": 16.0, "to_world": tw,
                       "near_clip": 1e-2, "far_clip": 1e3, "film": film, "sampler": smp}
        d["sensor"].update(sensor_extra or {})
    else:
        d["sensor"] = {"type": "perspective", "fov": 40.0, "to_world": tw, "near_clip": 1e-2, "far_clip": 1e3, "film": film, "sampler": smp}
    return d


def cornell_config1(res=256, spp=64, max_depth=6):
    """Config 1: mi.cornell_box(), 256x256, 64 spp, path max_depth=6."""
    from .scene import cornell_box
    d = cornell_box()
    d["integrator"]["max_depth"] = max_depth
    d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = res
    d["sensor"]["sampler"]["sample_count"] = spp
    return d


def lowpass_texture(n=2048, seed=0, passes=3, radius=None):
    """Config 4's texture: numpy.random.default_rng(seed).random((n, n, 3)) low-passed (repeated box blur, wrap-around) and
    stretched back to [0.05, 0.95]; float32 [n, n, 3]."""
    rng = np.random.default_rng(seed)
    a = rng.random((n, n, 3), dtype=np.float32).astype(np.float64)
    r = radius if radius is not None else max(1, n // 64)
    for _ in range(passes):
        for axis in (0, 1):
            c = np.cumsum(np.concatenate([a.take(range(n - r - 1, n), axis), a, a.take(range(0, r), axis)], axis), axis)
            hi = c.take(range(2 * r + 1, n + 2 * r + 1), axis); lo = c.take(range(0, n), axis)
            a = (hi - lo) / (2 * r + 1)
    a -= a.min(axis=(0, 1), keepdims=True); a /= np.maximum(a.max(axis=(0, 1), keepdims=True), 1e-12)
    return (0.05 + 0.9 * a).astype(f32)


def textured_boxes_scene(n_boxes=64, width=1920, height=1080, spp=64, lens_design="exp2a", max_depth=6, seed=0, tex_size=2048,
                         sensor="realisticlens", sensor_extra=None):
    """Config 4: ground plane + n_boxes boxes (cube shapes: per-vertex normals and uvs), one RGB bitmap texture (bilinear,
    repeat) on the ground and on every other box, one area light, imaged through the Nikon `exp2a` 22-interface stack with
    exit-pupil sampling (rear-element sampling vignettes ~99 % of the samples of this design on a 35 mm film)."""
    rng = np.random.default_rng(seed)
    T = Transform4f
    tex = {"type": "bitmap", "data": lowpass_texture(tex_size, seed), "filter_type": "bilinear", "wrap_mode": "repeat", "raw": True}
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": max_depth, "rr_depth": 5},
         "ground": {"type": "rectangle", "to_world": T.rotate([1, 0, 0], -90).scale(12.0),
                    "bsdf": {"type": "diffuse", "reflectance": tex}},
         "lamp": {"type": "rectangle", "to_world": T.translate([0, 9.0, 2.0]).rotate([1, 0, 0], 90).scale([4.0, 4.0, 1.0]),
                  "bsdf": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.0, 0.0, 0.0]}},
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [9.0, 8.5, 8.0]}}}}
    side = int(np.ceil(np.sqrt(n_boxes)))
    for b in range(n_boxes):
        gx, gz = b % side, b // side
        pos = [(gx - (side - 1) / 2) * 2.2 + rng.uniform(-0.3, 0.3), 0.0, (gz - (side - 1) / 2) * 2.2 + rng.uniform(-0.3, 0.3)]
        sx, sy, sz = rng.uniform(0.35, 0.8), rng.uniform(0.3, 1.6), rng.uniform(0.35, 0.8)
        pos[1] = sy
        if b % 2 == 0:
            bsdf = {"type": "diffuse", "reflectance": tex}
        elif b % 8 == 1:
            bsdf = {"type": "conductor", "material": "Cu"}
        else:
            bsdf = {"type": "diffuse", "reflectance": {"type": "rgb", "value": rng.uniform(0.2, 0.8, 3).tolist()}}
        d[f"box{b:02d}"] = {"type": "cube", "to_world": T.translate(pos).rotate([0, 1, 0], float(rng.uniform(0, 90))).scale([sx, sy, sz]),
                            "bsdf": bsdf}
    tw = T.look_at(origin=[0.0, 7.0, 16.0], target=[0.0, 0.5, 0.0], up=[0, 1, 0])
    film = {"type": "hdrfilm", "width": width, "height": height, "rfilter": {"type": "gaussian"}, "pixel_format": "rgb"}
    smp = {"type": "independent", "sample_count": spp}
    if sensor == "realisticlens":
        d["sensor"] = {"type": "realisticlens", "lens_design": lens_design, "object_distance": 17.0, "to_world": tw, "near_clip": 1e-2,
                       "far_clip": 1e3, "sample_exit_pupil": True, "shutter_open": 0.0, "shutter_close": 1.0, "film": film, "sampler": smp}
        d["sensor"].update(sensor_extra or {})
    else:
        d["sensor"] = {"type": "perspective", "fov": 40.0, "to_world": tw, "near_clip": 1e-2, "far_clip": 1e3, "film": film, "sampler": smp}
    return d
