"""Scene assembly: mi.load_dict for the plugin set of the hot path (src/python/python/xml.py, src/core/xml.cpp
-> PluginManager::create_object src/core/plugin.cpp:177-210 in the reference) and Scene (src/render/scene.cpp).

Shapes become one flat triangle list + shape/BSDF/emitter tables (include/lensrt.h lrt_scene_desc):
  rectangle (src/shapes/rectangle.cpp), cube (src/shapes/cube.cpp), `mesh` from arrays (vertices/faces),
  obj / ply files (ascii obj, ascii|binary_little_endian ply).
BSDFs: diffuse, conductor, dielectric, twosided(diffuse|conductor). Emitters: area (on rectangles)."""
from __future__ import annotations

import ctypes as C
import os
import struct

import numpy as np

from . import _capi as capi
from .film import FilmDesc, ReconstructionFilter, develop_rgb
from .sensors import Film, PerspectiveCamera, RealisticLensCamera
from .transform import Transform4f

f32 = np.float32


class TextureDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("wrap", C.c_int32), ("rgb", C.POINTER(C.c_float)),
                ("to_uv", C.c_float * 6), ("kind", C.c_int32), ("color0", C.c_float * 3), ("color1", C.c_float * 3),
                ("linewidth", C.c_float), ("spokes", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("n_vertices", C.c_int32), ("vertices", C.POINTER(C.c_float)),
                ("n_triangles", C.c_int64), ("indices", C.POINTER(C.c_int32)), ("tri_shape", C.POINTER(C.c_int32)),
                ("n_shapes", C.c_int32), ("shape_bsdf", C.POINTER(C.c_int32)), ("shape_emitter", C.POINTER(C.c_int32)),
                ("n_bsdfs", C.c_int32), ("bsdfs", C.POINTER(C.c_float)),
                ("n_emitters", C.c_int32), ("emitters", C.POINTER(C.c_float)),
                ("vertex_normals", C.POINTER(C.c_float)), ("vertex_uvs", C.POINTER(C.c_float)),
                ("shape_flags", C.POINTER(C.c_int32)),
                ("n_textures", C.c_int32), ("textures", C.POINTER(TextureDesc)),
                ("n_spheres", C.c_int32), ("spheres", C.POINTER(C.c_float))]


class RenderParams(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("spp", C.c_int32), ("max_depth", C.c_int32), ("rr_depth", C.c_int32),
                ("shard_index", C.c_int32), ("shard_count", C.c_int32), ("chunks", C.c_int32),
                ("rng_mode", C.c_int32), ("block_size", C.c_int32), ("base_seed", C.c_uint32), ("hide_emitters", C.c_int32)]


# ------------------------------------------------------------------------------------------------ plugins
def _rgb(v, default):
    if v is None:
        return np.array(default, f32)
    if isinstance(v, dict):
        v = v.get("value", default)
    a = np.asarray(v, f32).reshape(-1)
    return np.repeat(a, 3) if a.size == 1 else a


# the reference's conductor table (resources/data/ior/*.spd converted to RGB by its scalar_rgb build: values read back
# through mi.traverse(mi.load_dict({'type': 'conductor', 'material': m})); float32)
CONDUCTORS = {"Cu": ([0.2010054737329483, 0.9237495064735413, 1.1022154092788696], [3.913262128829956, 2.453045129776001, 2.14208984375]),
              "Au": ([0.14303597807884216, 0.3753070831298828, 1.4420453310012817], [3.982996940612793, 2.38555645942688, 1.6033592224121094]),
              "Al": ([1.6575008630752563, 0.8804047107696533, 0.521244466304779], [9.223811149597168, 6.26950216293335, 4.8370041847229]),
              "Ag": ([0.15527619421482086, 0.11672795563936234, 0.13838763535022736], [4.828354358673096, 3.122222423553467, 2.1469011306762695]),
              "Cr": ([4.3696675300598145, 2.9165053367614746, 1.6549046039581299], [5.206546783447266, 4.231306076049805, 3.7546987533569336]),
              "W": ([4.370816707611084, 3.300246477127075, 2.9984073638916016], [3.500640630722046, 2.6048738956451416, 2.27323317527771]),
              "none": ([0.0, 0.0, 0.0], [1.0, 1.0, 1.0])}
IOR = {"vacuum": 1.0, "air": 1.000277, "water": 1.3330, "bk7": 1.5046, "diamond": 2.419, "acrylic glass": 1.49, "pyrex": 1.470}


WRAP_MODES = {"repeat": 0, "mirror": 1, "clamp": 2}


# procedural textures: `checkerboard` (src/textures/checkerboard.cpp) and the fork's line patterns `checkerlines`, `comb2d`,
# `polarlines` (src/textures/{checkerlines,comb2d,polarlines}.cpp): kind, default color0 / color1 / linewidth / spokes
PROCEDURAL_TEXTURES = {"checkerboard": (1, .4, .2, 0.0, 0), "checkerlines": (2, .4, .05, .05, 0), "comb2d": (3, .4, .05, .01, 0),
                       "polarlines": (4, .4, .05, .05, 16)}
TEXTURE_TYPES = ("bitmap",) + tuple(PROCEDURAL_TEXTURES)


def _to_uv(d):
    tu = d.get("to_uv")
    m = np.eye(3, dtype=f32) if tu is None else np.asarray(getattr(tu, "matrix", tu), f32)
    if m.shape == (4, 4):      # "A 4x4 matrix can also be provided. In that case, the last row and columns will be ignored"
        m = np.array([[m[0, 0], m[0, 1], m[0, 3]], [m[1, 0], m[1, 1], m[1, 3]], [0, 0, 1]], f32)
    return np.ascontiguousarray(m.reshape(3, 3)[:2].reshape(-1), f32)


def load_texture(d: dict) -> dict:
    """`bitmap` texture plugin (src/textures/bitmap.cpp:115-310) -> {'rgb': [h,w,3] float32 linear, 'wrap', 'to_uv'}.
    Sources: 'filename' (.exr / .pfm natively; 8-bit formats through OpenCV when present, sRGB -> linear unless raw)
    or 'data' ([h,w,3] array). Only the bilinear filter of RGB textures is on the hot path.
    Procedural patterns (PROCEDURAL_TEXTURES) with constant colours add 'kind', 'color0', 'color1', 'linewidth', 'spokes'."""
    if d.get("type") in PROCEDURAL_TEXTURES:
        kind, c0, c1, lw, spokes = PROCEDURAL_TEXTURES[d["type"]]
        for k in ("color0", "color1"):
            if isinstance(d.get(k), dict) and d[k].get("type") not in (None, "rgb", "spectrum", "uniform"):
                raise ValueError(f"{d['type']} texture: only constant colours are built")
        return {"rgb": np.zeros((1, 1, 3), f32), "wrap": 0, "to_uv": _to_uv(d), "kind": kind, "color0": _rgb(d.get("color0"), [c0] * 3),
                "color1": _rgb(d.get("color1"), [c1] * 3), "linewidth": f32(d.get("linewidth", lw)), "spokes": int(d.get("spokes", spokes))}
    if d.get("type") != "bitmap":
        raise ValueError(f"texture plugin '{d.get('type')}' is outside the hot path of this build")
    if d.get("filter_type", "bilinear") != "bilinear":
        raise ValueError("bitmap textures: only filter_type='bilinear' is built")
    if "data" in d:
        img = np.asarray(d["data"], f32)
    else:
        fn = d["filename"]
        ext = fn.lower().rsplit(".", 1)[-1]
        if ext in ("exr", "pfm"):
            from . import imageio
            img = (imageio.read_exr if ext == "exr" else imageio.read_pfm)(fn).astype(f32)
        else:
            try:
                import cv2
            except ImportError as e:
                raise ValueError(f"reading '{fn}' needs OpenCV; .exr and .pfm are read natively") from e
            raw = cv2.imread(fn, cv2.IMREAD_UNCHANGED)
            if raw is None:
                raise ValueError(f"cannot read '{fn}'")
            if raw.ndim == 2:
                raw = raw[..., None].repeat(3, -1)
            img = raw[..., 2::-1].astype(f32) / f32(255 if raw.dtype == np.uint8 else 65535)
            if not d.get("raw", False):    # Bitmap::convert undoes the sRGB gamma of 8-bit images
                img = np.where(img <= 0.04045, img / 12.92, ((img + 0.055) / 1.055) ** 2.4).astype(f32)
    if img.ndim == 2:
        img = img[..., None].repeat(3, -1)
    if img.ndim != 3 or img.shape[2] < 3:
        raise ValueError("bitmap texture: expected an RGB image")
    return {"rgb": np.ascontiguousarray(img[..., :3], f32), "wrap": WRAP_MODES[d.get("wrap_mode", "repeat")], "to_uv": _to_uv(d), "kind": 0}


def bsdf_row(d: dict, textures=None) -> np.ndarray:
    """One BSDF dict -> the 16-float row of lrt_scene_desc. A `bitmap` reflectance is appended to `textures` and
    referenced from column 15 (texture id + 1)."""
    r = np.zeros(16, f32)
    t = d["type"]
    if t == "twosided":
        inner = next(v for v in d.values() if isinstance(v, dict) and "type" in v)
        r = bsdf_row(inner, textures); r[1] = 1
        return r
    if t == "diffuse":
        refl = d.get("reflectance")
        if isinstance(refl, dict) and refl.get("type") in TEXTURE_TYPES:
            if textures is None:
                raise ValueError("a textured BSDF needs a scene to live in")
            hit = next((i for i, t in enumerate(textures) if t.get("_src") is refl), None)    # one dictionary shared by many BSDFs = one texture
            if hit is None:
                textures.append(dict(load_texture(refl), _src=refl)); hit = len(textures) - 1
            r[15] = hit + 1
            refl = None
        r[0] = 0; r[2:5] = _rgb(refl, [0.5, 0.5, 0.5])
    elif t == "conductor":
        r[0] = 1; r[2:5] = _rgb(d.get("specular_reflectance"), [1, 1, 1])
        if "material" in d:
            eta, k = CONDUCTORS[d["material"]]
        else:
            eta, k = _rgb(d.get("eta"), [0, 0, 0]), _rgb(d.get("k"), [1, 1, 1])
        r[5:8], r[8:11] = eta, k
    elif t == "dielectric":
        def ior(v, dflt):
            v = d.get(v, dflt)
            return IOR[v] if isinstance(v, str) else float(v)
        r[0] = 2; r[2:5] = _rgb(d.get("specular_reflectance"), [1, 1, 1]); r[11:14] = _rgb(d.get("specular_transmittance"), [1, 1, 1])
        r[14] = f32(ior("int_ior", "bk7")) / f32(ior("ext_ior", "air"))
    elif t == "dispersive":
        # the fork's DispersiveDielectric (src/bsdfs/dispersive_dielectric.cpp:209-224,425-431): Cauchy fit through
        # (n_d, V_d). The *_rgb variants evaluate it at 589.3 nm, where A + B / 0.5893^2 collapses to n_d (up to float32
        # rounding, reproduced here), so the surface behaves as a `dielectric` with eta = int_ior(589.3) / ext_ior(589.3).
        def eval_ior(n_d, V_d, wl=f32(0.5893)):
            n_d, V_d = f32(n_d), f32(V_d)
            B = f32(0) if not V_d > f32(5.9604644775390625e-08) else f32(f32(n_d - f32(1)) * f32(f32(1) / f32(V_d * f32(
                f32(f32(1) / f32(f32(0.48613) * f32(0.48613))) - f32(f32(1) / f32(f32(0.65627) * f32(0.65627)))))))
            A = f32(n_d - f32(B * f32(f32(1) / f32(f32(0.5893) * f32(0.5893)))))
            return f32(A + f32(B * f32(f32(1) / f32(wl * wl))))
        r[0] = 2; r[2:5] = _rgb(d.get("specular_reflectance"), [1, 1, 1]); r[11:14] = _rgb(d.get("specular_transmittance"), [1, 1, 1])
        r[14] = f32(eval_ior(d.get("int_ior", 1.5046), d.get("int_V_d", 64.17)) *
                    f32(f32(1) / eval_ior(d.get("ext_ior", 1.0), d.get("ext_V_d", 0.0))))
    else:
        raise ValueError(f"BSDF plugin '{t}' is outside the hot path of this build")
    return r


def _rect_mesh(tw: Transform4f):
    """rectangle.cpp as two triangles carrying the rectangle's own parameterisation uv = (x, y) / 2 + 1/2 (:556-557), so
    that dp_du (hence the shading frame, interaction.h:245-255) and si.uv are those of the analytic shape."""
    v = tw.transform_affine(np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], f32))
    return v, np.array([[0, 1, 2], [0, 2, 3]], np.int32), None, np.array([[0, 0], [1, 0], [1, 1], [0, 1]], f32)


def _transform_normals(tw: Transform4f, n):
    """Transform * Normal (transform.h:146-155: inverse transpose), normalised as the mesh loaders do."""
    it = np.linalg.inv(tw.matrix[:3, :3].astype(np.float64)).T
    out = np.asarray(n, np.float64) @ it.T
    return (out / np.maximum(np.linalg.norm(out, axis=1, keepdims=True), 1e-30)).astype(f32)


def _cube_mesh(tw: Transform4f):
    """src/shapes/cube.cpp:107-133: 24 vertices with per-face normals and a [0,1]^2 parameterisation of every face."""
    v = np.array([[1, -1, -1], [1, -1, 1], [-1, -1, 1], [-1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, 1, 1], [1, 1, 1],
                  [1, -1, -1], [1, 1, -1], [1, 1, 1], [1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1], [-1, -1, 1],
                  [-1, -1, 1], [-1, 1, 1], [-1, 1, -1], [-1, -1, -1], [1, 1, -1], [1, -1, -1], [-1, -1, -1], [-1, 1, -1]], f32)
    n = np.repeat(np.array([[0, -1, 0], [0, 1, 0], [1, 0, 0], [0, 0, 1], [-1, 0, 0], [0, 0, -1]], f32), 4, 0)
    uv = np.tile(np.array([[0, 1], [1, 1], [1, 0], [0, 0]], f32), (6, 1))
    f = np.array([[0, 1, 2], [3, 0, 2], [4, 5, 6], [7, 4, 6], [8, 9, 10], [11, 8, 10], [12, 13, 14], [15, 12, 14],
                  [16, 17, 18], [19, 16, 18], [20, 21, 22], [23, 20, 22]], np.int32)
    return tw.transform_affine(v), f, _transform_normals(tw, n), uv


def vertex_normals(verts, faces):
    """Mesh::recompute_vertex_normals (src/render/mesh.cpp:321-375): face normals weighted by the face angle at the
    vertex (Thuermer & Wuethrich); what the obj / ply loaders do for files without normals unless face_normals=true."""
    v = np.asarray(verts, np.float64); f = np.asarray(faces, np.int64)
    p = [v[f[:, k]] for k in range(3)]
    n = np.cross(p[1] - p[0], p[2] - p[0])
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    ok = ln[:, 0] > 0
    n = np.where(ln > 0, n / np.maximum(ln, 1e-300), 0.0)
    out = np.zeros_like(v)
    for k in range(3):
        d0 = p[(k + 1) % 3] - p[k]; d1 = p[(k + 2) % 3] - p[k]
        d0 /= np.maximum(np.linalg.norm(d0, axis=1, keepdims=True), 1e-300); d1 /= np.maximum(np.linalg.norm(d1, axis=1, keepdims=True), 1e-300)
        # unit_angle (math.h): numerically robust angle between unit vectors
        dp = (d0 * d1).sum(1)
        ang = np.where(dp < 0, np.pi - 2 * np.arcsin(np.clip(0.5 * np.linalg.norm(d1 + d0, axis=1), -1, 1)),
                       2 * np.arcsin(np.clip(0.5 * np.linalg.norm(d1 - d0, axis=1), -1, 1)))
        np.add.at(out, f[:, k], n * (ang * ok)[:, None])
    ln = np.linalg.norm(out, axis=1, keepdims=True)
    out = np.where(ln > 0, out / np.maximum(ln, 1e-300), np.array([1.0, 0.0, 0.0]))
    return out.astype(f32)


def load_obj(path):
    """Wavefront OBJ (src/shapes/obj.cpp): positions, optional vt / vn; vertices are split per (v, vt, vn) triple.
    Returns (verts, faces, normals | None, uvs | None)."""
    vs, vts, vns, fs, key_of, out_v, out_t, out_n = [], [], [], [], {}, [], [], []
    with open(path) as f:
        for line in f:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                vs.append([float(x) for x in p[1:4]])
            elif p[0] == "vt":
                vts.append([float(x) for x in p[1:3]])
            elif p[0] == "vn":
                vns.append([float(x) for x in p[1:4]])
            elif p[0] == "f":
                idx = []
                for tok in p[1:]:
                    q = (tok.split("/") + ["", ""])[:3]
                    key = tuple((int(x) - 1 if int(x) > 0 else n + int(x)) if x else -1 for x, n in zip(q, (len(vs), len(vts), len(vns))))
                    if key not in key_of:
                        key_of[key] = len(out_v); out_v.append(vs[key[0]])
                        out_t.append(vts[key[1]] if key[1] >= 0 else [0.0, 0.0]); out_n.append(vns[key[2]] if key[2] >= 0 else [0.0, 0.0, 0.0])
                    idx.append(key_of[key])
                for k in range(1, len(idx) - 1):
                    fs.append([idx[0], idx[k], idx[k + 1]])
    uv = np.array(out_t, f32).reshape(-1, 2) if vts else None
    if uv is not None:
        uv[:, 1] = 1 - uv[:, 1]      # obj.cpp flips v by default (flip_tex_coords = true)
    return (np.array(out_v, f32).reshape(-1, 3), np.array(fs, np.int32).reshape(-1, 3),
            np.array(out_n, f32).reshape(-1, 3) if vns else None, uv)


def load_ply(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"ply"
        fmt, elems, cur = None, [], None
        while True:
            line = f.readline().decode().strip()
            if line == "end_header":
                break
            p = line.split()
            if p[0] == "format":
                fmt = p[1]
            elif p[0] == "element":
                cur = {"name": p[1], "count": int(p[2]), "props": []}; elems.append(cur)
            elif p[0] == "property":
                cur["props"].append(p[1:])
        tmap = {"float": "f", "float32": "f", "double": "d", "float64": "d", "uchar": "B", "uint8": "B", "char": "b", "int": "i",
                "int32": "i", "uint": "I", "uint32": "I", "short": "h", "ushort": "H", "int16": "h", "uint16": "H"}
        verts, faces, normals, uvs = None, [], None, None
        for e in elems:
            if fmt == "ascii":
                rows = [f.readline().split() for _ in range(e["count"])]
                if e["name"] == "vertex":
                    names = [p[-1] for p in e["props"]]
                    ix = [names.index(c) for c in "xyz"]
                    verts = np.array([[float(r[i]) for i in ix] for r in rows], f32)
                    if "nx" in names:
                        normals = np.array([[float(r[names.index(c)]) for c in ("nx", "ny", "nz")] for r in rows], f32)
                    for cu, cv in (("u", "v"), ("texture_u", "texture_v"), ("s", "t")):
                        if cu in names and cv in names:
                            uvs = np.array([[float(r[names.index(cu)]), float(r[names.index(cv)])] for r in rows], f32); break
                elif e["name"] == "face":
                    for r in rows:
                        n = int(r[0]); idx = [int(x) for x in r[1:1 + n]]
                        faces += [[idx[0], idx[k], idx[k + 1]] for k in range(1, n - 1)]
            else:
                assert fmt == "binary_little_endian", fmt
                if e["name"] == "vertex":
                    dt = np.dtype([(p[-1], "<" + tmap[p[0]]) for p in e["props"]])
                    a = np.frombuffer(f.read(dt.itemsize * e["count"]), dt)
                    verts = np.stack([a["x"], a["y"], a["z"]], 1).astype(f32)
                    if "nx" in dt.names:
                        normals = np.stack([a["nx"], a["ny"], a["nz"]], 1).astype(f32)
                    for cu, cv in (("u", "v"), ("texture_u", "texture_v"), ("s", "t")):
                        if cu in dt.names and cv in dt.names:
                            uvs = np.stack([a[cu], a[cv]], 1).astype(f32); break
                elif e["name"] == "face":
                    p = e["props"][0]   # list <count type> <index type> vertex_indices
                    ct, it = tmap[p[1]], tmap[p[2]]
                    for _ in range(e["count"]):
                        n = struct.unpack("<" + ct, f.read(struct.calcsize(ct)))[0]
                        idx = struct.unpack("<" + it * n, f.read(struct.calcsize(it) * n))
                        faces += [[idx[0], idx[k], idx[k + 1]] for k in range(1, n - 1)]
                else:
                    raise ValueError("unsupported ply element " + e["name"])
    return verts, np.array(faces, np.int32).reshape(-1, 3), normals, uvs


class SceneData:
    """Flat arrays of lrt_scene_desc, plus names for mi.traverse()."""

    def __init__(self):
        self.vertices, self.indices, self.tri_shape = [], [], []
        self.shape_bsdf, self.shape_emitter, self.shape_names = [], [], []
        self.bsdf_rows, self.bsdf_names, self.emitter_rows = [], [], []
        self.normals, self.uvs, self.shape_flags, self.textures = [], [], [], []
        self.rectangles = []      # (shape id, to_world 4x4) of `rectangle` shapes (the oracle can intersect them analytically)
        self.spheres = []         # analytic `sphere` shapes: (shape id, centre, radius, flip_normals, to_world 4x4, to_object 4x4)
        self._nv = 0

    def add_environment(self, radiance):
        """`constant` emitter (src/emitters/constant.cpp): row kind 2; its bounding sphere is filled in by arrays()."""
        if any(r[13] == 2 for r in self.emitter_rows):
            raise ValueError("Only one environment emitter can be specified per scene.")
        row = np.zeros(16, f32); row[0:3] = radiance; row[3] = -1; row[13] = 2
        self.emitter_rows.append(row)

    def add_sphere(self, name, to_world, bsdf_id, flip_normals=False):
        """`sphere` shape (src/shapes/sphere.cpp:148-174): centre and radius extracted from to_world in float32."""
        sid = self.add_shape(name, np.zeros((0, 3), f32), np.zeros((0, 3), np.int32), bsdf_id)
        m = to_world.matrix
        col = m[:3, 0]
        radius = np.sqrt(f32(f32(col[0] * col[0]) + f32(col[1] * col[1])) + f32(col[2] * col[2]), dtype=f32)
        self.spheres.append((sid, m[:3, 3].copy(), f32(radius), bool(flip_normals), m.copy(), to_world.inverse_matrix.copy()))
        return sid

    def add_disk(self, name, to_world, bsdf_id, emitter=None):
        """`disk` shape (src/shapes/disk.cpp): the unit disk of object space's xy plane, an analytic shape next to the spheres
        (record: unnormalised frame normal, -, shape, -, kind 1, -, to_world rows, to_object rows). `emitter`: {"radiance", "directional"}."""
        sid = self.add_shape(name, np.zeros((0, 3), f32), np.zeros((0, 3), np.int32), bsdf_id)
        m, inv = to_world.matrix, to_world.inverse_matrix
        n = inv[2, :3].astype(f32)             # to_world * Normal3f(0, 0, 1): the inverse transpose's third column; normalised by the library
        self.spheres.append((sid, n, f32(0), False, m.copy(), inv.copy(), 1))
        if emitter is not None:
            row = np.zeros(16, f32); row[0:3] = emitter["radiance"]; row[3] = sid
            row[4:7], row[7:10], row[10:13] = m[:3, 3], m[:3, 0], m[:3, 1]
            row[13] = 3; row[14] = 1.0 if emitter.get("directional", False) else 0.0
            row[15] = 1.0 if float(np.dot(np.cross(m[:3, 0].astype(np.float64), m[:3, 1].astype(np.float64)), n)) >= 0 else -1.0
            self.emitter_rows.append(row); self.shape_emitter[sid] = len(self.emitter_rows) - 1
        return sid

    def add_bsdf(self, name, row):
        self.bsdf_rows.append(np.asarray(row, f32)); self.bsdf_names.append(name)
        return len(self.bsdf_rows) - 1

    def add_shape(self, name, verts, faces, bsdf_id, emitter=None, normals=None, uvs=None, rectangle=False):
        sid = len(self.shape_bsdf)
        verts = np.ascontiguousarray(verts, f32).reshape(-1, 3); faces = np.ascontiguousarray(faces, np.int32).reshape(-1, 3)
        self.vertices.append(verts); self.indices.append(faces + self._nv); self._nv += len(verts)
        self.normals.append(np.zeros((len(verts), 3), f32) if normals is None else np.ascontiguousarray(normals, f32).reshape(len(verts), 3))
        self.uvs.append(np.zeros((len(verts), 2), f32) if uvs is None else np.ascontiguousarray(uvs, f32).reshape(len(verts), 2))
        self.shape_flags.append((1 if normals is not None else 0) | (2 if uvs is not None else 0) | (4 if rectangle else 0))
        self.tri_shape.append(np.full(len(faces), sid, np.int32))
        self.shape_bsdf.append(bsdf_id); self.shape_names.append(name)
        if emitter is not None:
            row = np.zeros(16, f32); row[0:3] = emitter["radiance"]; row[3] = sid
            row[4:7], row[7:10], row[10:13] = emitter["center"], emitter["ex"], emitter["ey"]
            row[13] = emitter.get("kind", 0)
            row[14] = 1.0 if emitter.get("directional", False) else 0.0
            self.emitter_rows.append(row); self.shape_emitter.append(len(self.emitter_rows) - 1)
        else:
            self.shape_emitter.append(-1)
        return sid

    def bounding_sphere(self):
        """Scene bbox (scene.cpp:43) -> BoundingBox::bounding_sphere (bbox.h:113-115,330-333), padded as constant.cpp:77-82; float32."""
        lo, hi = np.full(3, np.inf, f32), np.full(3, -np.inf, f32)
        for v in self.vertices:
            if len(v):
                lo, hi = np.minimum(lo, v.min(0)), np.maximum(hi, v.max(0))
        for sp in self.spheres:
            c, r, m = sp[1], sp[2], sp[4]
            if len(sp) > 6:                                      # disk.cpp:149-161
                c, r = m[:3, 3], np.sqrt(m[:3, 0] * m[:3, 0] + m[:3, 1] * m[:3, 1], dtype=f32)
            lo, hi = np.minimum(lo, (c - r).astype(f32)), np.maximum(hi, (c + r).astype(f32))
        if not np.all(lo <= hi):
            return np.zeros(3, f32), f32(1)
        c = ((hi + lo) * f32(.5)).astype(f32)
        d = (c - hi).astype(f32)
        r = np.sqrt(f32(f32(d[0] * d[0]) + f32(d[1] * d[1])) + f32(d[2] * d[2]), dtype=f32)
        eps = f32(8.94069671630859375e-05)
        return c, max(eps, f32(r * f32(f32(1) + eps)))

    def arrays(self):
        cat = lambda xs, dt, shape: (np.ascontiguousarray(np.concatenate(xs), dt) if xs else np.zeros(shape, dt))
        for row in self.emitter_rows:
            if row[13] == 2:
                row[4:7], row[7] = self.bounding_sphere()
        sph = np.zeros((len(self.spheres), 32), f32)
        for i, sp in enumerate(self.spheres):
            sid, c, r, flip, m, minv = sp[:6]
            sph[i, 0:3], sph[i, 3], sph[i, 4], sph[i, 5] = c, r, sid, float(flip)
            sph[i, 6] = float(sp[6]) if len(sp) > 6 else 0.0      # 1: a disk (then [0:3] is its frame normal)
            sph[i, 8:20] = m[:3].reshape(-1)
            sph[i, 20:32] = minv[:3].reshape(-1)      # m_to_object = m_to_world.inverse(): the transform's own float32 inverse
        return dict(spheres=sph, vertices=cat(self.vertices, f32, (0, 3)), indices=cat(self.indices, np.int32, (0, 3)),
                    tri_shape=cat(self.tri_shape, np.int32, (0,)),
                    shape_bsdf=np.array(self.shape_bsdf, np.int32), shape_emitter=np.array(self.shape_emitter, np.int32),
                    bsdfs=np.ascontiguousarray(np.stack(self.bsdf_rows) if self.bsdf_rows else np.zeros((0, 16), f32), f32),
                    emitters=np.ascontiguousarray(np.stack(self.emitter_rows) if self.emitter_rows else np.zeros((0, 16), f32), f32),
                    vertex_normals=cat(self.normals, f32, (0, 3)), vertex_uvs=cat(self.uvs, f32, (0, 2)),
                    shape_flags=np.array(self.shape_flags, np.int32), textures=list(self.textures),
                    rect_shapes=np.array([r[0] for r in self.rectangles], np.int32),
                    rect_to_world=np.ascontiguousarray(np.stack([r[1] for r in self.rectangles]) if self.rectangles
                                                       else np.zeros((0, 4, 4), f32), f32),
                    rect_to_object=np.ascontiguousarray(np.stack([r[2] for r in self.rectangles]) if self.rectangles
                                                        else np.zeros((0, 4, 4), f32), f32))


# ------------------------------------------------------------------------------------------------ objects
class PathIntegrator:
    """`path` plugin (src/integrators/path.cpp): max_depth (-1 = inf), rr_depth (5), hide_emitters."""

    def __init__(self, max_depth=-1, rr_depth=5, hide_emitters=False):
        self.max_depth, self.rr_depth, self.hide_emitters = int(max_depth), int(rr_depth), bool(hide_emitters)

    def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True):
        return render(scene, sensor=sensor, integrator=self, seed=seed, spp=spp, develop=develop)


class PRBIntegrator(PathIntegrator):
    """`prb` plugin (src/python/python/ad/integrators/prb.py): the `path` estimator with a reverse-mode derivative.
    render() is PathIntegrator's; render_backward() is RBIntegrator.render_backward (common.py:631-796): the gradient image is
    pulled through develop (division by the filter-weight sum, recomputed for the differential phase's own sample positions)
    and the splat, then lrt_render_backward walks the same paths (sampling and Russian roulette detached)."""

    def render_backward(self, scene, params, grad_in, sensor=0, seed=0, spp=0):
        import torch
        s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
        spp = int(spp) if spp else s.sampler.sample_count()
        grad_in = torch.as_tensor(grad_in, dtype=torch.float32, device="cuda")
        # W of the differential phase: a max_depth = 0 render splats weight 1 (and RGB 0) at the very positions it samples
        wsum = render(scene, sensor=s, integrator=PathIntegrator(0, self.rr_depth), seed=seed, spp=spp, develop=False)[..., 3]
        gs = torch.zeros((s.film.crop_h, s.film.crop_w, 4), dtype=torch.float32, device="cuda")
        gs[..., :3] = grad_in / torch.where(wsum == 0, torch.ones_like(wsum), wsum)[..., None]     # hdrfilm.cpp:390-392, transposed
        gb, ge = render_backward(scene, gs, sensor=s, integrator=self, seed=seed, spp=spp)
        grads = params.gradients(gb, ge)
        for k, g in grads.items():
            t = params[k]
            if getattr(t, "requires_grad", False):
                g = g.to(t.device, t.dtype)
                t.grad = g.clone() if t.grad is None else t.grad + g
        return grads

    def render_forward(self, scene, params, sensor=0, seed=0, spp=0):
        """Forward-mode derivative image for the tangents stored in `params[key].grad`-style `.tangent` attributes is not part
        of the hot path; the notebooks and the prb tests of the path use reverse mode only."""
        raise ValueError("render_forward: forward-mode differentiation is outside the hot path of this build (use render_backward)")


class PtraceParams(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("spp", C.c_int32), ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("hide_emitters", C.c_int32),
                ("shard_index", C.c_int32), ("shard_count", C.c_int32), ("rng_mode", C.c_int32), ("n_threads", C.c_int32),
                ("base_seed", C.c_uint32), ("to_camera", C.c_float * 12)]


class PTracerIntegrator(PathIntegrator):
    """`ptracer` plugin (src/integrators/ptracer.cpp under AdjointIntegrator::render, src/render/integrator.cpp:544-760):
    max_depth, rr_depth, hide_emitters. The image is the RGB of the film's storage as it is (no division by a weight)."""

    def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True, shard=(0, 1), rng_mode=0, n_threads=0):
        """rng_mode=1 / n_threads: the scalar AdjointIntegrator's seeding convention for that many worker threads (parity mode)."""
        import torch
        s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
        spp = int(spp) if spp else s.sampler.sample_count()
        fd = film_desc(s.film, coalesce=False, normalize=True)
        storage = torch.zeros((s.film.crop_h, s.film.crop_w, 4), dtype=torch.float32, device="cuda")
        tc = (C.c_float * 12)(*[float(x) for x in np.asarray(getattr(s, "to_camera", np.zeros((4, 4), f32)), f32)[:3].reshape(-1)])
        pp = PtraceParams(int(seed) & 0xffffffff, spp, self.max_depth, self.rr_depth, int(self.hide_emitters), shard[0], shard[1], int(rng_mode),
                          int(n_threads) if n_threads else (os.cpu_count() or 1), s.sampler.base_seed & 0xffffffff, tc)
        capi.check(capi.lib.lrt_ptrace(scene.handle, s.handle, C.byref(fd), C.byref(pp), capi.dev_ptr(storage), capi.stream_ptr()))
        return storage[..., :3].contiguous() if develop else storage

    def render_backward_lens(self, scene, grad_image, sensor=0, seed=0, spp=0, shard=(0, 1)):
        """d(loss)/d(interface table) [n_if, 16] for grad_image = d(loss)/d(image) [H, W, 3] of render(seed, spp): the same light
        paths replayed, every lens connection differentiated (lrt_ptrace_backward). Sum the result over ranks when sharded."""
        import torch
        s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
        spp = int(spp) if spp else s.sampler.sample_count()
        if not hasattr(s, "interfaces"):
            raise ValueError("render_backward_lens: the lens adjoint needs a realistic lens sensor")
        fd = film_desc(s.film, coalesce=False, normalize=True)
        gs = torch.zeros((s.film.crop_h, s.film.crop_w, 4), dtype=torch.float32, device="cuda")
        gs[..., :3] = torch.as_tensor(grad_image, dtype=torch.float32, device="cuda")
        n_if = len(s.interfaces())
        rows = torch.zeros((n_if, 16), dtype=torch.float32, device="cuda")
        tc = (C.c_float * 12)(*[float(x) for x in np.asarray(getattr(s, "to_camera", np.zeros((4, 4), f32)), f32)[:3].reshape(-1)])
        pp = PtraceParams(int(seed) & 0xffffffff, spp, self.max_depth, self.rr_depth, int(self.hide_emitters), shard[0], shard[1], 0, 1,
                          s.sampler.base_seed & 0xffffffff, tc)
        capi.check(capi.lib.lrt_ptrace_backward(scene.handle, s.handle, C.byref(fd), C.byref(pp), capi.dev_ptr(gs), capi.dev_ptr(rows),
                                                capi.stream_ptr()))
        return rows


class IndependentSampler:
    """`independent` plugin (src/samplers/independent.cpp): sample_count, seed (base seed)."""

    def __init__(self, sample_count=4, seed=0):
        self.sample_count_, self.base_seed = int(sample_count), int(seed)
        self._seed, self._n, self._cursor = 0, 1, 0

    def sample_count(self):
        return self.sample_count_

    def seed(self, seed, wavefront_size=1):
        self._seed, self._n, self._cursor = (self.base_seed + int(seed)) & 0xffffffff, int(wavefront_size), 0

    def _draw(self, k):
        """next k numbers of every lane (re-derives the streams; unit-test helper, not the render path)."""
        import torch
        out = torch.empty((self._n, self._cursor + k), dtype=torch.float32, device="cuda")
        capi.check(capi.lib.lrt_sampler_draw(C.c_uint32(self._seed), C.c_uint32(0), C.c_int64(self._n),
                                             C.c_int32(self._cursor + k), capi.dev_ptr(out), capi.stream_ptr()))
        res = out[:, self._cursor:]
        self._cursor += k
        return res

    def next_1d(self):
        return self._draw(1)[:, 0]

    def next_2d(self):
        return self._draw(2)


class Scene:
    def __init__(self, data: SceneData, sensors, integrator):
        self.data, self._sensors, self._integrator = data, sensors, integrator
        self.arrays = data.arrays()
        self.handle = None
        self._upload()

    def _upload(self):
        a = self.arrays
        opt = lambda arr, conv: conv(arr.reshape(-1)) if arr.size else None
        flags = a.get("shape_flags")
        has_attr = flags is not None and flags.size and bool(flags.any())
        tex = a.get("textures") or []
        tdesc = (TextureDesc * max(len(tex), 1))()
        for i, t in enumerate(tex):
            h, w, _ = t["rgb"].shape
            f3 = lambda v: (C.c_float * 3)(*[float(x) for x in v])   # noqa: E731
            tdesc[i] = TextureDesc(w, h, int(t["wrap"]), capi.fp(t["rgb"].reshape(-1)), (C.c_float * 6)(*[float(x) for x in t["to_uv"]]),
                                   int(t.get("kind", 0)), f3(t.get("color0", [0, 0, 0])), f3(t.get("color1", [0, 0, 0])),
                                   float(t.get("linewidth", 0.0)), float(t.get("spokes", 0)))
        d = SceneDesc(len(a["vertices"]), opt(a["vertices"], capi.fp), len(a["indices"]), opt(a["indices"], capi.ip),
                      opt(a["tri_shape"], capi.ip), len(a["shape_bsdf"]), opt(a["shape_bsdf"], capi.ip),
                      opt(a["shape_emitter"], capi.ip), len(a["bsdfs"]), opt(a["bsdfs"], capi.fp),
                      len(a["emitters"]), opt(a["emitters"], capi.fp),
                      opt(a["vertex_normals"], capi.fp) if has_attr else None, opt(a["vertex_uvs"], capi.fp) if has_attr else None,
                      opt(flags, capi.ip) if has_attr else None, len(tex), tdesc if tex else None,
                      len(a.get("spheres", ())), opt(a["spheres"], capi.fp) if len(a.get("spheres", ())) else None)
        h = C.c_void_p()
        capi.check(capi.lib.lrt_scene_create(C.byref(d), C.byref(h)))
        if self.handle:
            capi.lib.lrt_scene_destroy(self.handle)
        self.handle = h

    def __del__(self):
        if getattr(self, "handle", None):
            if capi is not None and getattr(capi, "lib", None) is not None:     # (module globals are gone at interpreter exit)
                capi.lib.lrt_scene_destroy(self.handle)
            self.handle = None

    def sensors(self):
        return self._sensors

    def integrator(self):
        return self._integrator

    def info(self):
        nn, nt, dep, sec = C.c_int64(), C.c_int64(), C.c_int32(), C.c_double()
        capi.check(capi.lib.lrt_scene_info(self.handle, C.byref(nn), C.byref(nt), C.byref(dep), C.byref(sec)))
        return {"bvh_nodes": nn.value, "triangles": nt.value, "bvh_depth": dep.value, "build_seconds": sec.value}

    def ray_intersect(self, o, d, maxt=None, any_hit=False):
        """Scene::ray_intersect_preliminary / ray_test for [n,3] origins/directions -> (t, u, v, prim)."""
        import torch
        o = torch.as_tensor(o, dtype=torch.float32, device="cuda").reshape(-1, 3)
        d = torch.as_tensor(d, dtype=torch.float32, device="cuda").reshape(-1, 3)
        n = o.shape[0]
        mt = (torch.full((n, 1), 3.402823466e+38, device="cuda") if maxt is None
              else torch.as_tensor(maxt, dtype=torch.float32, device="cuda").reshape(-1, 1))
        od = torch.cat([o, d, mt], 1).contiguous()
        out = torch.empty((n, 4), dtype=torch.float32, device="cuda")
        capi.check(capi.lib.lrt_scene_intersect(self.handle, C.c_int32(int(any_hit)), C.c_int64(n), capi.dev_ptr(od),
                                                capi.dev_ptr(out), capi.stream_ptr()))
        return out[:, 0], out[:, 1], out[:, 2], out[:, 3].to(torch.int64)

    def stats(self):
        out = (C.c_uint64 * 7)()
        capi.check(capi.lib.lrt_render_stats(self.handle, out))
        return {"samples": out[0], "lens_misses": out[1], "rays": out[2], "bounces": out[3], "bvh_nodes": out[4], "triangles": out[5],
                "stack_clamps": out[6]}


def film_desc(film: Film, coalesce=True, normalize=False) -> FilmDesc:
    rf = ReconstructionFilter(film.rfilter, getattr(film, "rfilter_param", 0.0))
    return FilmDesc(film.crop_w, film.crop_h, film.crop_x, film.crop_y, 0, 4, rf.kind, rf.param, int(coalesce), int(normalize))


WAVEFRONT_LIMIT = 0xffffffff     # integrator.cpp:248: one 32-bit sample index per wavefront lane


def tea32(v0, v1, rounds=4):
    """mi.sample_tea_32 (include/mitsuba/core/random.h:76-90) on host integers."""
    v0 &= 0xffffffff; v1 &= 0xffffffff; s = 0
    for _ in range(rounds):
        s = (s + 0x9e3779b9) & 0xffffffff
        v0 = (v0 + ((((v1 << 4) & 0xffffffff) + 0xa341316c) ^ (v1 + s) ^ ((v1 >> 5) + 0xc8013ea4))) & 0xffffffff
        v1 = (v1 + ((((v0 << 4) & 0xffffffff) + 0xad90777d) ^ (v0 + s) ^ ((v0 >> 5) + 0x7e95761e))) & 0xffffffff
    return v0, v1


def reference_block_size(width, height, n_threads, block_size=0):
    """The block size the reference's scalar render loop settles on (src/render/integrator.cpp:171-183): the integrator's
    `block_size` if set, else MI_BLOCK_SIZE = 32 halved until there is a block for every worker thread. Per-pixel sampler seeds
    depend on it (block id x block_size^2 + Morton index), so a sample-for-sample comparison of a SMALL image with the reference
    (rng_mode=1) has to pass the value the reference used: render(..., rng_mode=1, block_size=reference_block_size(w, h, threads))."""
    if block_size:
        return int(block_size)
    bs = 32
    while not (bs == 1 or ((width + bs - 1) // bs) * ((height + bs - 1) // bs) >= n_threads):
        bs //= 2
    return bs


def render_passes(n_pixels, spp, seed, wavefront_limit=WAVEFRONT_LIMIT):
    """[(pass seed, pass spp)] of one render: SamplingIntegrator::render splits W*H*spp > 2^32 - 1 samples into equal passes
    (src/render/integrator.cpp:247-265: spp_per_pass = spp / ceil(wavefront / limit)). The reference keeps the 2^32 sampler
    states alive between passes (64 GB of PCG32 state); here pass k > 0 starts fresh streams from the seed
    sample_tea_32(seed, k)[0] — statistically equivalent, not the same numbers — and a remainder pass keeps the total at
    exactly `spp` where the reference's integer division would drop samples."""
    wavefront = n_pixels * spp
    if wavefront <= wavefront_limit:
        return [(seed & 0xffffffff, spp)]
    per = spp // ((wavefront + wavefront_limit - 1) // wavefront_limit)
    if per < 1:
        raise ValueError("film too large: a single sample per pixel exceeds the 2^32 wavefront limit")
    out, done, k = [], 0, 0
    while done < spp:
        n = min(per, spp - done)
        out.append((seed & 0xffffffff if k == 0 else tea32(seed, k)[0], n))
        done += n; k += 1
    return out


def render(scene: Scene, params=None, sensor=0, integrator=None, seed=0, seed_grad=0, spp=0, spp_grad=0, develop=True,
           shard=(0, 1), storage=None, rng_mode=0, wavefront_limit=WAVEFRONT_LIMIT, block_size=0):
    """mi.render (src/python/python/util.py:410-525), forward part: returns the developed [H,W,3] image as a
    torch CUDA tensor (or the raw R,G,B,W storage with develop=False). rng_mode=1 selects the scalar variants' sampler
    convention (lensrt.h). Renders of more than 2^32 - 1 samples run as several passes (render_passes)."""
    import torch
    if params is not None:
        if not isinstance(params, SceneParameters):
            raise Exception("params should be an instance of mi.SceneParameter!")
        diff = [k for k in params if getattr(params[k], "requires_grad", False)]
        if diff and develop and torch.is_grad_enabled():
            if seed_grad == 0:
                seed_grad = tea32(int(seed), 1)[0]                # util.py:518-520
            elif seed_grad == seed:
                raise Exception("The primal and differential seed should be different to ensure unbiased gradient computation!")
            return _RenderOp.apply(scene, sensor, params, integrator or scene.integrator(), (int(seed), int(seed_grad)),
                                   (int(spp), int(spp_grad) if spp_grad else int(spp)), diff, *[params[k] for k in diff])
        params.update()
    s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
    it = integrator or scene.integrator()
    if isinstance(it, PTracerIntegrator):
        return it.render(scene, sensor=s, seed=seed, spp=spp, develop=develop, shard=shard, rng_mode=rng_mode, n_threads=block_size)
    spp = int(spp) if spp else s.sampler.sample_count()
    fd = film_desc(s.film)
    if storage is None:
        storage = torch.zeros((s.film.crop_h, s.film.crop_w, 4), dtype=torch.float32, device="cuda")
    if rng_mode == 1:
        rp = RenderParams(int(seed) & 0xffffffff, spp, it.max_depth, it.rr_depth, 0, 1, 0, 1, int(block_size), s.sampler.base_seed & 0xffffffff,
                          int(it.hide_emitters))
        if tuple(shard) != (0, 1):
            raise ValueError("rng_mode=1 (scalar sampler convention) is a single-GPU parity mode")
        capi.check(capi.lib.lrt_render(scene.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(storage), capi.stream_ptr()))
    else:
        for pseed, pspp in render_passes(s.film.crop_w * s.film.crop_h, spp, s.sampler.base_seed + int(seed), wavefront_limit):
            rp = RenderParams(pseed, pspp, it.max_depth, it.rr_depth, shard[0], shard[1], 0, 0, 0, 0, int(it.hide_emitters))
            capi.check(capi.lib.lrt_render(scene.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(storage), capi.stream_ptr()))
    return develop_rgb(storage) if develop else storage


def render_backward(scene: Scene, grad_storage, sensor=0, integrator=None, seed=0, spp=0, shard=(0, 1), wavefront_limit=WAVEFRONT_LIMIT):
    """Adjoint of render() for diffuse reflectance / conductor specular reflectance / area-light radiance
    (RBIntegrator.render_backward, src/python/python/ad/integrators/common.py:631-796 with prb.py): grad_storage = d(loss)/
    d(R,G,B,W storage) [H,W,4] -> (grad_bsdfs [n_bsdfs,16], grad_emitters [n_emitters,16]) torch CUDA tensors."""
    import torch
    s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
    it = integrator or scene.integrator()
    spp = int(spp) if spp else s.sampler.sample_count()
    fd = film_desc(s.film)
    gb = torch.zeros((max(len(scene.arrays["bsdfs"]), 1), 16), dtype=torch.float32, device="cuda")
    ge = torch.zeros((max(len(scene.arrays["emitters"]), 1), 16), dtype=torch.float32, device="cuda")
    gs = grad_storage.contiguous()
    for pseed, pspp in render_passes(s.film.crop_w * s.film.crop_h, spp, s.sampler.base_seed + int(seed), wavefront_limit):
        rp = RenderParams(pseed, pspp, it.max_depth, it.rr_depth, shard[0], shard[1], 0, 0, 0, 0, int(it.hide_emitters))
        capi.check(capi.lib.lrt_render_backward(scene.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(gs),
                                                capi.dev_ptr(gb), capi.dev_ptr(ge), capi.stream_ptr()))
    return gb[:len(scene.arrays["bsdfs"])], ge[:len(scene.arrays["emitters"])]


def _make_render_op():
    import torch

    class RenderOp(torch.autograd.Function):
        """_RenderOp (src/python/python/util.py:356-407) as a torch.autograd.Function: forward = integrator.render with the
        primal seed / spp, backward = integrator.render_backward with the differential seed / spp."""

        @staticmethod
        def forward(ctx, scene, sensor, params, integrator, seeds, spps, keys, *tensors):
            ctx.args = (scene, sensor, params, integrator, seeds, spps, keys)
            params.update()
            with torch.no_grad():
                return render(scene, None, sensor, integrator, seed=seeds[0], spp=spps[0])

        @staticmethod
        def backward(ctx, grad_out):
            scene, sensor, params, integrator, seeds, spps, keys = ctx.args
            if not hasattr(integrator, "render_backward"):
                raise Exception("the integrator has no reverse-mode derivative: use {'type': 'prb'}")
            s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
            saved = {k: params[k].grad for k in keys}
            for k in keys:
                params[k].grad = None
            integrator.render_backward(scene, params, grad_out.contiguous(), s, seeds[1], spps[1])
            out = []
            for k in keys:
                g = params[k].grad
                out.append(torch.zeros_like(params[k]) if g is None else g)
                params[k].grad = saved[k]
            return (None,) * 7 + tuple(out)

    return RenderOp


class _LazyRenderOp:
    _op = None

    @classmethod
    def apply(cls, *a):
        if cls._op is None:
            cls._op = _make_render_op()
        return cls._op.apply(*a)


_RenderOp = _LazyRenderOp


class SceneParameters(dict):
    """mi.traverse(scene) for the parameters the adjoint reaches: '<bsdf key>.reflectance.value' (diffuse),
    '<bsdf key>.specular_reflectance.value' (conductor), '<shape key>.emitter.radiance.value' — torch tensors of 3 floats.
    update() pushes them to the device tables (src/python/python/util.py:292-353)."""

    def __init__(self, scene):
        import torch
        super().__init__()
        self.scene, self._where = scene, {}
        a = scene.arrays
        for i, name in enumerate(scene.data.bsdf_names):
            t = int(a["bsdfs"][i, 0])
            if a["bsdfs"][i, 15] != 0:
                continue                  # a bitmap reflectance is not a differentiable parameter of this build
            key = f"{name}.{'reflectance' if t == 0 else 'specular_reflectance'}.value"
            self[key] = torch.tensor(a["bsdfs"][i, 2:5].copy()); self._where[key] = ("bsdf", i)
            if t == 2:
                key = f"{name}.specular_transmittance.value"
                self[key] = torch.tensor(a["bsdfs"][i, 11:14].copy()); self._where[key] = ("bsdf_t", i)
        for sid, e in enumerate(a["shape_emitter"]):
            if e >= 0:
                key = f"{scene.data.shape_names[sid]}.emitter.radiance.value"
                self[key] = torch.tensor(a["emitters"][e, 0:3].copy()); self._where[key] = ("emitter", int(e))

    def update(self, values=None):
        """params.update() / params.update(optimizer): push the current values to the device tables."""
        if values is not None:
            for k in values:
                if k in self._where:
                    self[k] = values[k]
        a = self.scene.arrays
        for key, (kind, i) in self._where.items():
            v = self[key].detach().cpu().numpy().astype(f32)
            if kind == "bsdf":
                a["bsdfs"][i, 2:5] = v
            elif kind == "bsdf_t":
                a["bsdfs"][i, 11:14] = v
            else:
                a["emitters"][i, 0:3] = v
        rad = np.ascontiguousarray(a["emitters"][:, 0:3], f32)
        capi.check(capi.lib.lrt_scene_set_params(self.scene.handle, capi.fp(a["bsdfs"].reshape(-1)) if a["bsdfs"].size else None,
                                                 capi.fp(rad.reshape(-1)) if rad.size else None))

    def gradients(self, grad_bsdfs, grad_emitters):
        """Split the adjoint's tables into per-key gradients."""
        pick = {"bsdf": lambda i: grad_bsdfs[i, 2:5], "bsdf_t": lambda i: grad_bsdfs[i, 11:14], "emitter": lambda i: grad_emitters[i, 0:3]}
        return {k: pick[kind](i) for k, (kind, i) in self._where.items()}


def traverse(scene):
    return SceneParameters(scene)


# ------------------------------------------------------------------------------------------------ load_dict
def _film_from(d):
    d = d or {"type": "hdrfilm"}
    rf = d.get("rfilter", {"type": "gaussian"})
    crop = None
    if "crop_width" in d:
        crop = (d["crop_width"], d["crop_height"], d.get("crop_offset_x", 0), d.get("crop_offset_y", 0))
    film = Film(d.get("width", 768), d.get("height", 576), crop=crop, diagonal=d.get("diagonal", 35.0), rfilter=rf["type"],
                pixel_format=d.get("pixel_format", "rgb"))
    film.rfilter_param = float(rf.get("stddev", rf.get("radius", 0.0))) if rf["type"] != "box" else 0.0
    return film


def _sensor_from(d):
    d = dict(d)
    t = d.pop("type")
    film = _film_from(d.pop("film", None))
    smp = d.pop("sampler", {"type": "independent"})
    sampler = IndependentSampler(smp.get("sample_count", 4), smp.get("seed", 0))
    d.pop("focus_distance", None)
    d.pop("id", None) if t == "perspective" else None
    if t == "perspective":
        s = PerspectiveCamera(film, fov=d.get("fov", 45.0), to_world=d.get("to_world"), near_clip=d.get("near_clip", 1e-2),
                              far_clip=d.get("far_clip", 1e4), fov_axis=d.get("fov_axis", "x"))
    elif t == "realisticlens":
        # the properties the plugin and its bases query (realisticlens.cpp:844-878, sensor.cpp, endpoint.cpp); anything else is
        # the reference's "unreferenced properties" error (xml.cpp:1100, e.g. `fov` on a realistic lens)
        known = {"id", "to_world", "near_clip", "far_clip", "lens_design", "object_distance", "lens_focal_length", "lens_diameter", "stop_ratio",
                 "ap_size", "ap_offset", "sample_exit_pupil", "shutter_open", "shutter_close", "pupil_render_pos", "pupil_render_wv", "srf",
                 "prescription"}
        extra = sorted(k for k in d if k not in known and not isinstance(d[k], dict))
        if extra:
            raise ValueError(f"unreferenced properties {extra} in sensor plugin of type \"realisticlens\"")
        d.pop("id", None)
        tw = d.pop("to_world", None)
        nc = d.pop("near_clip", 1e-2)
        fc = d.pop("far_clip", 1e4)
        s = RealisticLensCamera(film, d.pop("lens_design", "singlet"), to_world=tw, near_clip=nc, far_clip=fc, **d)
    else:
        raise ValueError(f"sensor plugin '{t}' is outside the hot path of this build")
    s.sampler = sampler
    return s


_SHAPES = ("rectangle", "cube", "mesh", "obj", "ply", "sphere", "disk")
_BSDFS = ("diffuse", "conductor", "dielectric", "dispersive", "twosided")


def natural_key(name: str):
    """Order of the reference's property map (src/core/properties.cpp:49-75 SortKey): names compare as strings except that
    a trailing run of digits compares as a number ('_arg_2' < '_arg_10'; 'lamp' < 'lamp2'). Children of a scene — hence its
    shapes and emitters, and the emitter a sample selects — come out in this order, for dictionaries and XML alike."""
    import functools

    def cmp(a, b):
        i = 0
        while i < len(a) and i < len(b) and a[i] == b[i]:
            i += 1
        while i > 0 and a[i - 1].isdigit():
            i -= 1
        ta, tb = a[i:], b[i:]
        if ta[:1].isdigit() and tb[:1].isdigit() and ta.isdigit() and tb.isdigit() and not (int(ta) == int(tb) and len(a) != len(b)):
            return -1 if int(ta) < int(tb) else (1 if int(ta) > int(tb) else 0)
        return -1 if ta < tb else (1 if ta > tb else 0)
    return functools.cmp_to_key(cmp)(name)


def _child(v: dict, key: str, types):
    """A shape's nested plugin: under its conventional key, or (as the reference does) whichever child has a matching type."""
    c = v.get(key)
    if isinstance(c, dict):
        return c
    return next((x for x in v.values() if isinstance(x, dict) and x.get("type") in types), None)


def scene_data_from_dict(d: dict):
    """The GPU-free half of load_dict('scene'): flat geometry/material tables (SceneData) plus the sensor and
    integrator sub-dictionaries. Also what the tests hand to the CPU oracle."""
    data = SceneData()
    named_bsdf, sensor_dicts, integ = {}, [], None
    default_bsdf = default_black = None
    for key, v in d.items():   # named BSDFs first: they are the targets of {'type': 'ref'}
        if isinstance(v, dict) and v.get("type") in _BSDFS:
            named_bsdf[v.get("id", key)] = data.add_bsdf(v.get("id", key), bsdf_row(v, data.textures))
    # child objects in the order of the reference's property map (natural_key): scene.shapes() / scene.emitters() come out in
    # this order, and the emitter index a sample selects depends on it
    for key, v in sorted(d.items(), key=lambda kv: natural_key(kv[0])):
        if not isinstance(v, dict):
            continue
        vt = v.get("type")
        name = v.get("id", key)
        if vt in ("perspective", "realisticlens"):
            sensor_dicts.append(v)
        elif vt in ("path", "prb", "ptracer"):
            integ = v
        elif vt == "constant":
            rad = v.get("radiance")
            if isinstance(rad, dict) and rad.get("type") not in (None, "rgb", "spectrum", "uniform"):
                raise ValueError("constant emitter: only a constant radiance is built")
            data.add_environment(_rgb(rad, [1, 1, 1]))
        elif vt in _SHAPES:
            tw = v.get("to_world")
            tw = tw if isinstance(tw, Transform4f) else Transform4f(tw)
            b = _child(v, "bsdf", _BSDFS + ("ref",))
            em = _child(v, "emitter", ("area", "directionalarea"))
            if b is None and em is not None:       # shape.cpp:66-73: the default BSDF of an emitter is a BLACK diffuse one
                if default_black is None:
                    default_black = data.add_bsdf("__default_emitter", bsdf_row({"type": "diffuse", "reflectance": 0.0}))
                bid = default_black
            elif b is None:
                if default_bsdf is None:
                    default_bsdf = data.add_bsdf("__default", bsdf_row({"type": "diffuse"}))
                bid = default_bsdf
            elif b.get("type") == "ref":
                bid = named_bsdf[b["id"]]
            else:
                bid = data.add_bsdf(f"{name}.bsdf", bsdf_row(b, data.textures))
            if vt == "sphere":
                if em is not None:
                    raise ValueError("area emitters on `sphere` shapes are outside the hot path of this build (use a mesh)")
                if "center" in v or "radius" in v:       # sphere.cpp:128-133: to_world * translate(center) * scale(radius)
                    c32, r32 = np.asarray(v.get("center", [0, 0, 0]), f32), f32(v.get("radius", 1.0))       # float32 operands, float32 products
                    tr = Transform4f.translate(c32); sc = Transform4f.scale(r32)
                    sc = Transform4f(sc.matrix, np.diag(np.array([f32(1) / r32] * 3 + [f32(1)], f32)))                   # dr::rcp in float32
                    tw = tw.matmul32(tr).matmul32(sc)
                data.add_sphere(name, tw, bid, v.get("flip_normals", False))
                continue
            if vt == "disk":
                if v.get("flip_normals", False):
                    tw = tw @ Transform4f(np.diag(np.array([1, 1, -1, 1], f32)))     # disk.cpp:99-101
                demit = None
                if em is not None:
                    if em.get("type") not in ("area", "directionalarea"):
                        raise ValueError(f"emitter `{em.get('type')}` on a disk: only `area` and `directionalarea` are built")
                    rad = em.get("radiance")
                    if isinstance(rad, dict) and rad.get("type") not in (None, "rgb", "spectrum", "uniform"):
                        raise ValueError("emitters: only a constant radiance is built")
                    demit = {"radiance": _rgb(rad, [1, 1, 1]), "directional": em.get("type") == "directionalarea"}
                data.add_disk(name, tw, bid, demit)
                continue
            if v.get("flip_normals", False) and vt == "rectangle":
                tw = tw @ Transform4f(np.diag(np.array([1, 1, -1, 1], f32)))     # rectangle.cpp:100-103
            if vt == "rectangle":
                verts, faces, normals, uvs = _rect_mesh(tw)
            elif vt == "cube":
                verts, faces, normals, uvs = _cube_mesh(tw)
            else:
                if vt == "mesh":     # in-memory mesh (mi.Mesh): flat-shaded unless the arrays carry normals
                    verts, faces = np.asarray(v["vertices"], f32), np.asarray(v["faces"], np.int32)
                    normals, uvs = v.get("vertex_normals"), v.get("vertex_texcoords")
                else:
                    verts, faces, normals, uvs = (load_obj if vt == "obj" else load_ply)(v["filename"])
                    if vt == "ply" and uvs is not None and v.get("flip_tex_coords", False):
                        uvs = uvs.copy(); uvs[:, 1] = 1 - uvs[:, 1]
                    if vt == "obj" and uvs is not None and not v.get("flip_tex_coords", True):
                        uvs = uvs.copy(); uvs[:, 1] = 1 - uvs[:, 1]
                verts = tw.transform_affine(verts)
                if v.get("face_normals", False):
                    normals = None
                elif normals is not None:
                    normals = _transform_normals(tw, normals)
                elif vt in ("obj", "ply"):     # ply.cpp:444-449 / obj.cpp:401-404
                    normals = vertex_normals(verts, faces)
            if v.get("flip_normals", False) and vt != "rectangle":
                faces = faces[:, ::-1]                              # mesh.cpp:1539-1542: si.n and sh_frame.n change sign
                if normals is not None:
                    normals = -np.asarray(normals, f32)
            emitter = None
            if em is not None:
                if em.get("type") not in ("area", "directionalarea"):
                    raise ValueError(f"emitter plugin '{em.get('type')}' is outside the hot path of this build (only `area`, `directionalarea`)")
                rad = em.get("radiance")
                if isinstance(rad, dict) and rad.get("type") not in (None, "rgb", "spectrum", "uniform"):
                    raise ValueError("area emitters: only a constant radiance is built")
                emitter = {"radiance": _rgb(rad, [1, 1, 1]), "center": tw.transform_affine(np.zeros(3, f32)),
                           "ex": tw.transform_vector(np.array([1, 0, 0], f32)), "ey": tw.transform_vector(np.array([0, 1, 0], f32)),
                           "kind": 0 if vt == "rectangle" else 1,      # 1: faces sampled by area (Mesh::sample_position)
                           # src/emitters/directionalarea.cpp: emits along the surface normal only (the notebooks' field sources,
                           # design_problem.py:46-120); seen by the particle tracer alone — `path` refuses such scenes
                           "directional": em.get("type") == "directionalarea"}
            sid = data.add_shape(name, verts, faces, bid, emitter, normals, uvs, rectangle=vt == "rectangle")
            if vt == "rectangle":
                data.rectangles.append((sid, tw.matrix.copy(), tw.inverse_matrix.copy()))
        elif vt in _BSDFS or vt is None:
            pass
        else:
            raise ValueError(f"plugin '{vt}' ('{key}') is outside the hot path of this build")
    return data, sensor_dicts, integ


def load_dict(d: dict):
    """mi.load_dict for: scene, realisticlens/perspective (+hdrfilm, independent, rfilters), path, diffuse/
    conductor/dielectric/twosided, rectangle/cube/mesh/obj/ply, area."""
    t = d.get("type")
    if t in ("perspective", "realisticlens"):
        return _sensor_from(d)
    if t == "hdrfilm":
        return _film_from(d)
    if t == "independent":
        return IndependentSampler(d.get("sample_count", 4), d.get("seed", 0))
    if t in ("gaussian", "tent", "box"):
        return ReconstructionFilter(t, float(d.get("stddev", d.get("radius", 0.0))) if t != "box" else 0.0)
    if t == "ptracer":
        return PTracerIntegrator(d.get("max_depth", -1), d.get("rr_depth", 5), d.get("hide_emitters", False))
    if t in ("path", "prb"):
        # hide_emitters: with an environment emitter, camera rays that escape yield no valid sample (path.cpp:113-114)
        return (PRBIntegrator if t == "prb" else PathIntegrator)(d.get("max_depth", -1), d.get("rr_depth", 5), d.get("hide_emitters", False))
    if t != "scene":
        raise ValueError(f"plugin '{t}' is outside the hot path of this build")
    data, sensor_dicts, integ = scene_data_from_dict(d)
    return Scene(data, [_sensor_from(v) for v in sensor_dicts], load_dict(integ) if integ else PathIntegrator())


def cornell_box():
    """mi.cornell_box() (src/python/python/util.py:565-700), same structure and values."""
    T = Transform4f
    white = {"type": "ref", "id": "white"}
    return {
        "type": "scene",
        "integrator": {"type": "path", "max_depth": 8},
        "sensor": {"type": "perspective", "fov_axis": "smaller", "near_clip": 0.001, "far_clip": 100.0, "focus_distance": 1000,
                   "fov": 39.3077, "to_world": T.look_at(origin=[0, 0, 3.90], target=[0, 0, 0], up=[0, 1, 0]),
                   "sampler": {"type": "independent", "sample_count": 64},
                   "film": {"type": "hdrfilm", "width": 256, "height": 256, "rfilter": {"type": "gaussian"},
                            "pixel_format": "rgb", "component_format": "float32"}},
        "white": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.885809, 0.698859, 0.666422]}},
        "green": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.105421, 0.37798, 0.076425]}},
        "red": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.570068, 0.0430135, 0.0443706]}},
        "light": {"type": "rectangle", "to_world": T.translate([0.0, 0.99, 0.01]).rotate([1, 0, 0], 90).scale([0.23, 0.19, 0.19]),
                  "bsdf": white, "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [18.387, 13.9873, 6.75357]}}},
        "floor": {"type": "rectangle", "to_world": T.translate([0.0, -1.0, 0.0]).rotate([1, 0, 0], -90), "bsdf": white},
        "ceiling": {"type": "rectangle", "to_world": T.translate([0.0, 1.0, 0.0]).rotate([1, 0, 0], 90), "bsdf": white},
        "back": {"type": "rectangle", "to_world": T.translate([0.0, 0.0, -1.0]), "bsdf": white},
        "green-wall": {"type": "rectangle", "to_world": T.translate([1.0, 0.0, 0.0]).rotate([0, 1, 0], -90),
                       "bsdf": {"type": "ref", "id": "green"}},
        "red-wall": {"type": "rectangle", "to_world": T.translate([-1.0, 0.0, 0.0]).rotate([0, 1, 0], 90),
                     "bsdf": {"type": "ref", "id": "red"}},
        "small-box": {"type": "cube", "to_world": T.translate([0.335, -0.7, 0.38]).rotate([0, 1, 0], -17).scale(0.3), "bsdf": white},
        "large-box": {"type": "cube", "to_world": T.translate([-0.33, -0.4, -0.28]).rotate([0, 1, 0], 18.25).scale([0.3, 0.61, 0.3]),
                      "bsdf": white},
    }
