"""Scene dictionaries shared by the reference golden generator (tests/golden/gen_ref_integrator.py, run with the
reference's own scalar_rgb build on PYTHONPATH) and by the parity tests (run with lensrt / the oracle).

Every transform is a float32 4x4 matrix computed HERE (not by either side's Transform class) and wrapped with the
caller's `T(matrix)`, so that both sides see bit-identical matrices. Meshes and textures are small files under
tests/golden/assets/ that both loaders read (binary PLY with normals + uvs; uncompressed float EXR)."""
from __future__ import annotations

import os

import numpy as np

f32 = np.float32
ASSETS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "assets")


# ------------------------------------------------------------------------------------------------ float32 matrices
def m_translate(v):
    m = np.eye(4, dtype=f32); m[:3, 3] = np.asarray(v, f32); return m


def m_scale(v):
    v = np.asarray(v, f32)
    if v.ndim == 0:
        v = np.repeat(v, 3)
    m = np.eye(4, dtype=f32); m[0, 0], m[1, 1], m[2, 2] = v; return m


def m_rotate(axis, deg):
    a = np.asarray(axis, np.float64); a = a / np.linalg.norm(a)
    t = np.radians(deg); c, s = np.cos(t), np.sin(t)
    x, y, z = a
    m = np.eye(4, dtype=np.float64)
    m[:3, :3] = [[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s],
                 [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s],
                 [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c)]]
    return m.astype(f32)


def m_look_at(origin, target, up):
    o, t, u = (np.asarray(x, np.float64) for x in (origin, target, up))
    d = t - o; d /= np.linalg.norm(d)
    left = np.cross(u, d); left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    m = np.eye(4, dtype=np.float64)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, o
    return m.astype(f32)


def mm(*ms):
    out = np.eye(4, dtype=np.float64)
    for m in ms:
        out = out @ m.astype(np.float64)
    return out.astype(f32)


# ------------------------------------------------------------------------------------------------ assets
def icosphere(subdiv=2):
    """Unit icosphere with vertex normals (= positions) and spherical uvs."""
    t = (1 + 5 ** 0.5) / 2
    v = [[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1],
         [-t, 0, -1], [-t, 0, 1]]
    f = [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
         [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]]
    v = [np.asarray(p, np.float64) / np.linalg.norm(p) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                p = v[a] + v[b]; v.append(p / np.linalg.norm(p)); cache[k] = len(v) - 1
            return cache[k]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        f = nf
    v = np.asarray(v, np.float64)
    uv = np.stack([np.arctan2(v[:, 2], v[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(v[:, 1], -1, 1)) / np.pi], 1)
    return v.astype(f32), np.asarray(f, np.int32), v.astype(f32), uv.astype(f32)


def write_ply(path, verts, faces, normals=None, uvs=None):
    """binary_little_endian PLY (x y z [nx ny nz] [u v], uchar-count int faces)."""
    cols = [verts]
    head = ["ply", "format binary_little_endian 1.0", f"element vertex {len(verts)}", "property float x", "property float y",
            "property float z"]
    if normals is not None:
        cols.append(normals); head += ["property float nx", "property float ny", "property float nz"]
    if uvs is not None:
        cols.append(uvs); head += ["property float u", "property float v"]
    head += [f"element face {len(faces)}", "property list uchar int vertex_indices", "end_header"]
    with open(path, "wb") as fh:
        fh.write(("\n".join(head) + "\n").encode())
        fh.write(np.ascontiguousarray(np.concatenate(cols, 1), "<f4").tobytes())
        rec = np.zeros(len(faces), dtype=[("n", "u1"), ("i", "<i4", 3)])
        rec["n"] = 3; rec["i"] = faces
        fh.write(rec.tobytes())


def texture_array(n=32, seed=7):
    """Smooth random RGB texture in [0.05, 0.95], float32 [n, n, 3]."""
    rng = np.random.default_rng(seed)
    a = rng.random((n // 4 + 1, n // 4 + 1, 3))
    xs = np.linspace(0, n // 4, n, endpoint=False)
    i0 = xs.astype(int); fr = xs - i0
    a = a[i0] * (1 - fr)[:, None, None] + a[i0 + 1] * fr[:, None, None]
    a = a[:, i0] * (1 - fr)[None, :, None] + a[:, i0 + 1] * fr[None, :, None]
    return (0.05 + 0.9 * a).astype(f32)


MESH_SMOOTH = os.path.join(ASSETS, "icosphere_n_uv.ply")     # normals + uvs
MESH_PLAIN = os.path.join(ASSETS, "icosphere_plain.ply")     # positions only (loaders compute vertex normals)
TEXTURE = os.path.join(ASSETS, "noise32.exr")


# ------------------------------------------------------------------------------------------------ scenes
def _film(res, rfilter="gaussian"):
    return {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": rfilter}, "pixel_format": "rgb",
            "component_format": "float32"}


def _box_walls(T, white, green, red):
    R = m_rotate
    return {
        "floor": {"type": "rectangle", "to_world": T(mm(m_translate([0, -1, 0]), R([1, 0, 0], -90))), "bsdf": white},
        "ceiling": {"type": "rectangle", "to_world": T(mm(m_translate([0, 1, 0]), R([1, 0, 0], 90))), "bsdf": white},
        "back": {"type": "rectangle", "to_world": T(m_translate([0, 0, -1])), "bsdf": white},
        "green-wall": {"type": "rectangle", "to_world": T(mm(m_translate([1, 0, 0]), R([0, 1, 0], -90))), "bsdf": green},
        "red-wall": {"type": "rectangle", "to_world": T(mm(m_translate([-1, 0, 0]), R([0, 1, 0], 90))), "bsdf": red},
    }


def scene_dicts(T):
    """name -> (scene dict, spp). `T(matrix4x4 float32)` builds the caller's transform type."""
    white = {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.885809, 0.698859, 0.666422]}}
    green = {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.105421, 0.37798, 0.076425]}}
    red = {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.570068, 0.0430135, 0.0443706]}}
    light = {"type": "rectangle", "to_world": T(mm(m_translate([0.0, 0.99, 0.01]), m_rotate([1, 0, 0], 90), m_scale([0.23, 0.19, 0.19]))),
             "bsdf": white, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [18.387, 13.9873, 6.75357]}}}
    cam = T(m_look_at([0, 0, 3.9], [0, 0, 0], [0, 1, 0]))
    out = {}

    # Cornell box through the cemented doublet (the fork's sensor): rear-element sampling, heavy vignetting
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 6},
         "sensor": {"type": "realisticlens", "lens_design": "doublet", "object_distance": 3.9, "to_world": cam,
                    "near_clip": 1e-3, "far_clip": 100.0, "sampler": {"type": "independent", "sample_count": 32},
                    "film": _film(96)},
         "light": light}
    d.update(_box_walls(T, white, green, red))
    d["small-box"] = {"type": "cube", "to_world": T(mm(m_translate([0.335, -0.7, 0.38]), m_rotate([0, 1, 0], -17), m_scale(0.3))), "bsdf": white}
    out["cbox_doublet"] = (d, 32)

    # conductor + dielectric + twosided + two emitters, Russian roulette active (max_depth 10 > rr_depth 3), tent filter
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 10, "rr_depth": 3},
         "sensor": {"type": "perspective", "fov": 39.3077, "fov_axis": "smaller", "near_clip": 1e-3, "far_clip": 100.0, "to_world": cam,
                    "sampler": {"type": "independent", "sample_count": 64}, "film": _film(96, "tent")},
         "light": light,
         "light2": {"type": "rectangle", "to_world": T(mm(m_translate([-0.99, 0.2, 0.3]), m_rotate([0, 1, 0], 90), m_scale([0.15, 0.25, 1.0]))),
                    "bsdf": white, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [2.0, 6.0, 9.0]}}}}
    d.update(_box_walls(T, white, green, red))
    # (the glass box floats 1 cm above the floor: coplanar faces would make the closest hit a tie that every accelerator breaks its own way)
    d["small-box"] = {"type": "cube", "to_world": T(mm(m_translate([0.335, -0.69, 0.38]), m_rotate([0, 1, 0], -17), m_scale(0.3))),
                      "bsdf": {"type": "dielectric", "int_ior": 1.5046, "ext_ior": 1.000277}}
    d["large-box"] = {"type": "cube", "to_world": T(mm(m_translate([-0.33, -0.4, -0.28]), m_rotate([0, 1, 0], 18.25), m_scale([0.3, 0.61, 0.3]))),
                      "bsdf": {"type": "conductor", "material": "Cu"}}
    d["panel"] = {"type": "rectangle", "to_world": T(mm(m_translate([0.45, 0.1, -0.5]), m_rotate([0, 1, 0], -40), m_scale([0.3, 0.4, 1.0]))),
                  "bsdf": {"type": "twosided", "material": green}}
    out["cbox_materials"] = (d, 64)

    # smooth-shaded + textured meshes (vertex normals, uvs, bilinear bitmap), box filter
    tex = {"type": "bitmap", "filename": TEXTURE, "filter_type": "bilinear", "wrap_mode": "repeat", "raw": True}
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 5},
         "sensor": {"type": "perspective", "fov": 45.0, "near_clip": 1e-2, "far_clip": 100.0,
                    "to_world": T(m_look_at([0.3, 1.4, 3.2], [0, 0.2, 0], [0, 1, 0])),
                    "sampler": {"type": "independent", "sample_count": 32}, "film": _film(96, "box")},
         "ground": {"type": "rectangle", "to_world": T(mm(m_rotate([1, 0, 0], -90), m_scale(3.0))),
                    "bsdf": {"type": "diffuse", "reflectance": tex}},
         "ball": {"type": "ply", "filename": MESH_SMOOTH, "to_world": T(mm(m_translate([-0.6, 0.6, 0]), m_scale(0.6))),
                  "bsdf": {"type": "diffuse", "reflectance": tex}},
         "ball2": {"type": "ply", "filename": MESH_PLAIN, "to_world": T(mm(m_translate([0.8, 0.45, 0.3]), m_scale([0.45, 0.45, 0.6]))),
                   "bsdf": {"type": "conductor", "material": "Cu"}},
         "lamp": {"type": "rectangle", "to_world": T(mm(m_translate([0, 3.0, 0.5]), m_rotate([1, 0, 0], 90), m_scale([0.8, 0.8, 1.0]))),
                  "bsdf": white, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [12.0, 11.0, 10.0]}}}}
    out["textured"] = (d, 32)

    # area lights on general meshes (faces sampled by area): a cube (vertex normals + uvs), a smooth icosphere, next to a rectangle
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 5},
         "sensor": {"type": "perspective", "fov": 39.3077, "fov_axis": "smaller", "near_clip": 1e-3, "far_clip": 100.0, "to_world": cam,
                    "sampler": {"type": "independent", "sample_count": 32}, "film": _film(96)},
         "light": light,
         "cube-light": {"type": "cube", "to_world": T(mm(m_translate([-0.5, -0.75, 0.2]), m_rotate([0, 1, 0], 30), m_scale([0.2, 0.25, 0.15]))),
                        "bsdf": white, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [1.0, 4.0, 2.0]}}},
         "ball-light": {"type": "ply", "filename": MESH_SMOOTH, "to_world": T(mm(m_translate([0.45, -0.2, -0.2]), m_scale(0.25))),
                        "bsdf": white, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [5.0, 2.0, 1.0]}}}}
    d.update(_box_walls(T, white, green, red))
    out["mesh_lights"] = (d, 32)
    return out


def path_b_scene(T, spp=64, res=64, source="rectangle"):
    """The notebooks' "Path B" in miniature (notebooks/scripts/lens_design/design_problem.py:46-120,355-428): a lens made of
    MESHES (the two caps of a singlet: icosphere caps whose vertex z follows the sag, smooth normals computed by the loader,
    `dielectric` BSDF), a `directionalarea` field source tilted by two degrees, a white diffuse film plane at z = 0, a
    perspective sensor looking at the film plane, the particle tracer. -> scene dict for `T(matrix)` transforms."""
    assets = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "assets")
    white = {"type": "diffuse", "reflectance": {"type": "rgb", "value": [1.0, 1.0, 1.0]}}
    glass = {"type": "dielectric", "int_ior": 1.5168, "ext_ior": 1.000277}
    # source="disk": the notebooks' own choice (design_problem.py:355-371), an elliptic one here (dp_du != dp_dv)
    src_scale = [0.004, 0.004, 1.0] if source == "rectangle" else [0.0045, 0.0035, 1.0]
    return {
        "type": "scene", "integrator": {"type": "ptracer", "max_depth": 4, "hide_emitters": False},
        "sensor": {"type": "perspective", "fov": 30.0, "near_clip": 1e-3, "far_clip": 10.0,
                   "to_world": T(m_look_at([0, 0, 0.03], [0, 0, 0], [0, 1, 0])),
                   "sampler": {"type": "independent", "sample_count": spp},
                   "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "tent"}, "pixel_format": "rgb",
                            "component_format": "float32"}},
        "film_plane": {"type": "rectangle", "to_world": T(m_scale([0.008, 0.008, 1.0])), "bsdf": white},
        # cap normals face +z; behind the rear cap's normal (film side) is air, behind the front cap's is glass
        "lens_rear": {"type": "ply", "filename": os.path.join(assets, "cap_rear.ply"), "face_normals": False,
                      "bsdf": {"type": "dielectric", "int_ior": 1.000277, "ext_ior": 1.5168}},
        "lens_front": {"type": "ply", "filename": os.path.join(assets, "cap_front.ply"), "face_normals": False, "bsdf": glass},
        "field_source": {"type": source,
                         "to_world": T(mm(m_translate([0.0008, 0.0, 0.08]), m_rotate([0, 1, 0], 2.0), m_rotate([1, 0, 0], 180.0), m_scale(src_scale))),
                         "emitter": {"type": "directionalarea", "radiance": {"type": "rgb", "value": [3.0, 2.0, 1.0]}}},
    }


def disk_light_scene(T, spp=16, res=64):
    """An `area` light on an elliptic, tilted `disk` (src/shapes/disk.cpp + src/emitters/area.cpp) over a diffuse floor and a
    back wall, next to a rectangle light; perspective sensor, `path` integrator."""
    grey = {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.7, 0.6, 0.5]}}
    return {
        "type": "scene", "integrator": {"type": "path", "max_depth": 3},
        "sensor": {"type": "perspective", "fov": 50.0, "near_clip": 1e-2, "far_clip": 100.0,
                   "to_world": T(m_look_at([0, 1.2, 3.2], [0, 0.5, 0], [0, 1, 0])),
                   "sampler": {"type": "independent", "sample_count": spp},
                   "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "tent"}, "pixel_format": "rgb",
                            "component_format": "float32"}},
        "floor": {"type": "rectangle", "to_world": T(mm(m_rotate([1, 0, 0], -90.0), m_scale([2.0, 2.0, 1.0]))), "bsdf": grey},
        "wall": {"type": "rectangle", "to_world": T(mm(m_translate([0, 1.0, -1.5]), m_scale([2.0, 1.0, 1.0]))), "bsdf": grey},
        "lamp": {"type": "disk",
                 "to_world": T(mm(m_translate([0.3, 1.6, 0.2]), m_rotate([0, 0, 1], 12.0), m_rotate([1, 0, 0], 90.0), m_scale([0.5, 0.3, 1.0]))),
                 "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [6.0, 5.0, 4.0]}}},
        "panel": {"type": "rectangle", "to_world": T(mm(m_translate([-1.2, 0.8, 0.0]), m_rotate([0, 1, 0], 90.0), m_scale([0.3, 0.3, 1.0]))),
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [2.0, 3.0, 4.0]}}},
    }
