"""Scene assembly: mi.load_dict for the plugin set of the hot path (src/python/python/xml.py, src/core/xml.cpp
-> PluginManager::create_object src/core/plugin.cpp:177-210 in the reference) and Scene (src/render/scene.cpp).

Shapes become one flat triangle list + shape/BSDF/emitter tables (include/lensrt.h lrt_scene_desc):
  rectangle (src/shapes/rectangle.cpp), cube (src/shapes/cube.cpp), `mesh` from arrays (vertices/faces),
  obj / ply files (ascii obj, ascii|binary_little_endian ply).
BSDFs: diffuse, conductor, dielectric, twosided(diffuse|conductor). Emitters: area (on rectangles)."""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np

from . import _capi as capi
from .film import FilmDesc, ReconstructionFilter, develop_rgb
from .sensors import Film, PerspectiveCamera, RealisticLensCamera
from .transform import Transform4f

f32 = np.float32


class SceneDesc(C.Structure):
    _fields_ = [("n_vertices", C.c_int32), ("vertices", C.POINTER(C.c_float)),
                ("n_triangles", C.c_int64), ("indices", C.POINTER(C.c_int32)), ("tri_shape", C.POINTER(C.c_int32)),
                ("n_shapes", C.c_int32), ("shape_bsdf", C.POINTER(C.c_int32)), ("shape_emitter", C.POINTER(C.c_int32)),
                ("n_bsdfs", C.c_int32), ("bsdfs", C.POINTER(C.c_float)),
                ("n_emitters", C.c_int32), ("emitters", C.POINTER(C.c_float))]


class RenderParams(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("spp", C.c_int32), ("max_depth", C.c_int32), ("rr_depth", C.c_int32),
                ("shard_index", C.c_int32), ("shard_count", C.c_int32), ("chunks", C.c_int32)]


# ------------------------------------------------------------------------------------------------ plugins
def _rgb(v, default):
    if v is None:
        return np.array(default, f32)
    if isinstance(v, dict):
        v = v.get("value", default)
    a = np.asarray(v, f32).reshape(-1)
    return np.repeat(a, 3) if a.size == 1 else a


# a few entries of the reference's conductor table (src/render/ior.h / data/ior): RGB eta, k
CONDUCTORS = {"Cu": ([0.200438, 0.924033, 1.102212], [3.912949, 2.452848, 2.142188]),
              "Au": ([0.143119, 0.374957, 1.442479], [3.983160, 2.385721, 1.603215]),
              "Al": ([1.657460, 0.880369, 0.521229], [9.223869, 6.269523, 4.837001]),
              "Ag": ([0.155265, 0.116723, 0.138342], [4.828371, 3.122500, 2.147038]),
              "none": ([0.0, 0.0, 0.0], [1.0, 1.0, 1.0])}
IOR = {"vacuum": 1.0, "air": 1.000277, "water": 1.3330, "bk7": 1.5046, "diamond": 2.419, "acrylic glass": 1.49, "pyrex": 1.470}


def bsdf_row(d: dict) -> np.ndarray:
    """One BSDF dict -> the 16-float row of lrt_scene_desc."""
    r = np.zeros(16, f32)
    t = d["type"]
    if t == "twosided":
        inner = next(v for v in d.values() if isinstance(v, dict) and "type" in v)
        r = bsdf_row(inner); r[1] = 1
        return r
    if t == "diffuse":
        r[0] = 0; r[2:5] = _rgb(d.get("reflectance"), [0.5, 0.5, 0.5])
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
    v = tw.transform_affine(np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], f32))
    return v, np.array([[0, 1, 2], [0, 2, 3]], np.int32)


def _cube_mesh(tw: Transform4f):
    c = np.array([[x, y, z] for z in (-1, 1) for y in (-1, 1) for x in (-1, 1)], f32)
    quads = [(0, 2, 3, 1), (4, 5, 7, 6), (0, 1, 5, 4), (2, 6, 7, 3), (0, 4, 6, 2), (1, 3, 7, 5)]   # outward facing
    faces = []
    for a, b, cc, d in quads:
        faces += [[a, b, cc], [a, cc, d]]
    return tw.transform_affine(c), np.array(faces, np.int32)


def load_obj(path):
    vs, fs = [], []
    with open(path) as f:
        for line in f:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                vs.append([float(x) for x in p[1:4]])
            elif p[0] == "f":
                idx = [int(tok.split("/")[0]) for tok in p[1:]]
                idx = [i - 1 if i > 0 else len(vs) + i for i in idx]
                for k in range(1, len(idx) - 1):
                    fs.append([idx[0], idx[k], idx[k + 1]])
    return np.array(vs, f32).reshape(-1, 3), np.array(fs, np.int32).reshape(-1, 3)


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
        verts, faces = None, []
        for e in elems:
            if fmt == "ascii":
                rows = [f.readline().split() for _ in range(e["count"])]
                if e["name"] == "vertex":
                    names = [p[-1] for p in e["props"]]
                    ix = [names.index(c) for c in "xyz"]
                    verts = np.array([[float(r[i]) for i in ix] for r in rows], f32)
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
                elif e["name"] == "face":
                    p = e["props"][0]   # list <count type> <index type> vertex_indices
                    ct, it = tmap[p[1]], tmap[p[2]]
                    for _ in range(e["count"]):
                        n = struct.unpack("<" + ct, f.read(struct.calcsize(ct)))[0]
                        idx = struct.unpack("<" + it * n, f.read(struct.calcsize(it) * n))
                        faces += [[idx[0], idx[k], idx[k + 1]] for k in range(1, n - 1)]
                else:
                    raise ValueError("unsupported ply element " + e["name"])
    return verts, np.array(faces, np.int32).reshape(-1, 3)


class SceneData:
    """Flat arrays of lrt_scene_desc, plus names for mi.traverse()."""

    def __init__(self):
        self.vertices, self.indices, self.tri_shape = [], [], []
        self.shape_bsdf, self.shape_emitter, self.shape_names = [], [], []
        self.bsdf_rows, self.bsdf_names, self.emitter_rows = [], [], []
        self._nv = 0

    def add_bsdf(self, name, row):
        self.bsdf_rows.append(np.asarray(row, f32)); self.bsdf_names.append(name)
        return len(self.bsdf_rows) - 1

    def add_shape(self, name, verts, faces, bsdf_id, emitter=None):
        sid = len(self.shape_bsdf)
        verts = np.ascontiguousarray(verts, f32).reshape(-1, 3); faces = np.ascontiguousarray(faces, np.int32).reshape(-1, 3)
        self.vertices.append(verts); self.indices.append(faces + self._nv); self._nv += len(verts)
        self.tri_shape.append(np.full(len(faces), sid, np.int32))
        self.shape_bsdf.append(bsdf_id); self.shape_names.append(name)
        if emitter is not None:
            row = np.zeros(16, f32); row[0:3] = emitter["radiance"]; row[3] = sid
            row[4:7], row[7:10], row[10:13] = emitter["center"], emitter["ex"], emitter["ey"]
            self.emitter_rows.append(row); self.shape_emitter.append(len(self.emitter_rows) - 1)
        else:
            self.shape_emitter.append(-1)
        return sid

    def arrays(self):
        cat = lambda xs, dt, shape: (np.ascontiguousarray(np.concatenate(xs), dt) if xs else np.zeros(shape, dt))
        return dict(vertices=cat(self.vertices, f32, (0, 3)), indices=cat(self.indices, np.int32, (0, 3)),
                    tri_shape=cat(self.tri_shape, np.int32, (0,)),
                    shape_bsdf=np.array(self.shape_bsdf, np.int32), shape_emitter=np.array(self.shape_emitter, np.int32),
                    bsdfs=np.ascontiguousarray(np.stack(self.bsdf_rows) if self.bsdf_rows else np.zeros((0, 16), f32), f32),
                    emitters=np.ascontiguousarray(np.stack(self.emitter_rows) if self.emitter_rows else np.zeros((0, 16), f32), f32))


# ------------------------------------------------------------------------------------------------ objects
class PathIntegrator:
    """`path` plugin (src/integrators/path.cpp): max_depth (-1 = inf), rr_depth (5), hide_emitters."""

    def __init__(self, max_depth=-1, rr_depth=5, hide_emitters=False):
        self.max_depth, self.rr_depth, self.hide_emitters = int(max_depth), int(rr_depth), bool(hide_emitters)

    def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True):
        return render(scene, sensor=sensor, integrator=self, seed=seed, spp=spp, develop=develop)


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
        d = SceneDesc(len(a["vertices"]), opt(a["vertices"], capi.fp), len(a["indices"]), opt(a["indices"], capi.ip),
                      opt(a["tri_shape"], capi.ip), len(a["shape_bsdf"]), opt(a["shape_bsdf"], capi.ip),
                      opt(a["shape_emitter"], capi.ip), len(a["bsdfs"]), opt(a["bsdfs"], capi.fp),
                      len(a["emitters"]), opt(a["emitters"], capi.fp))
        h = C.c_void_p()
        capi.check(capi.lib.lrt_scene_create(C.byref(d), C.byref(h)))
        if self.handle:
            capi.lib.lrt_scene_destroy(self.handle)
        self.handle = h

    def __del__(self):
        if getattr(self, "handle", None):
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
        out = (C.c_uint64 * 6)()
        capi.check(capi.lib.lrt_render_stats(self.handle, out))
        return {"samples": out[0], "lens_misses": out[1], "rays": out[2], "bounces": out[3], "bvh_nodes": out[4], "triangles": out[5]}


def film_desc(film: Film, coalesce=True, normalize=False) -> FilmDesc:
    rf = ReconstructionFilter(film.rfilter, getattr(film, "rfilter_param", 0.0))
    return FilmDesc(film.crop_w, film.crop_h, film.crop_x, film.crop_y, 0, 4, rf.kind, rf.param, int(coalesce), int(normalize))


def render(scene: Scene, params=None, sensor=0, integrator=None, seed=0, seed_grad=0, spp=0, spp_grad=0, develop=True,
           shard=(0, 1), storage=None):
    """mi.render (src/python/python/util.py:410-525), forward part: returns the developed [H,W,3] image as a
    torch CUDA tensor (or the raw R,G,B,W storage with develop=False)."""
    import torch
    if params is not None:
        params.update()
    s = scene.sensors()[sensor] if isinstance(sensor, int) else sensor
    it = integrator or scene.integrator()
    spp = int(spp) if spp else s.sampler.sample_count()
    fd = film_desc(s.film)
    if storage is None:
        storage = torch.zeros((s.film.crop_h, s.film.crop_w, 4), dtype=torch.float32, device="cuda")
    rp = RenderParams((s.sampler.base_seed + int(seed)) & 0xffffffff, spp, it.max_depth, it.rr_depth, shard[0], shard[1], 0)
    capi.check(capi.lib.lrt_render(scene.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(storage), capi.stream_ptr()))
    return develop_rgb(storage) if develop else storage


def render_backward(scene: Scene, grad_storage, sensor=0, integrator=None, seed=0, spp=0, shard=(0, 1)):
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
    rp = RenderParams((s.sampler.base_seed + int(seed)) & 0xffffffff, spp, it.max_depth, it.rr_depth, shard[0], shard[1], 0)
    capi.check(capi.lib.lrt_render_backward(scene.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(grad_storage.contiguous()),
                                            capi.dev_ptr(gb), capi.dev_ptr(ge), capi.stream_ptr()))
    return gb[:len(scene.arrays["bsdfs"])], ge[:len(scene.arrays["emitters"])]


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
            if t in (0, 1):
                key = f"{name}.{'reflectance' if t == 0 else 'specular_reflectance'}.value"
                self[key] = torch.tensor(a["bsdfs"][i, 2:5].copy()); self._where[key] = ("bsdf", i)
        for sid, e in enumerate(a["shape_emitter"]):
            if e >= 0:
                key = f"{scene.data.shape_names[sid]}.emitter.radiance.value"
                self[key] = torch.tensor(a["emitters"][e, 0:3].copy()); self._where[key] = ("emitter", int(e))

    def update(self):
        a = self.scene.arrays
        for key, (kind, i) in self._where.items():
            v = self[key].detach().cpu().numpy().astype(f32)
            if kind == "bsdf":
                a["bsdfs"][i, 2:5] = v
            else:
                a["emitters"][i, 0:3] = v
        rad = np.ascontiguousarray(a["emitters"][:, 0:3], f32)
        capi.check(capi.lib.lrt_scene_set_params(self.scene.handle, capi.fp(a["bsdfs"].reshape(-1)) if a["bsdfs"].size else None,
                                                 capi.fp(rad.reshape(-1)) if rad.size else None))

    def gradients(self, grad_bsdfs, grad_emitters):
        """Split the adjoint's tables into per-key gradients."""
        return {k: (grad_bsdfs[i, 2:5] if kind == "bsdf" else grad_emitters[i, 0:3]) for k, (kind, i) in self._where.items()}


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
    if t == "perspective":
        s = PerspectiveCamera(film, fov=d.get("fov", 45.0), to_world=d.get("to_world"), near_clip=d.get("near_clip", 1e-2),
                              far_clip=d.get("far_clip", 1e4), fov_axis=d.get("fov_axis", "x"))
    elif t == "realisticlens":
        tw = d.pop("to_world", None)
        nc = d.pop("near_clip", 1e-2)
        fc = d.pop("far_clip", 1e4)
        s = RealisticLensCamera(film, d.pop("lens_design", "singlet"), to_world=tw, near_clip=nc, far_clip=fc, **d)
    else:
        raise ValueError(f"sensor plugin '{t}' is outside the hot path of this build")
    s.sampler = sampler
    return s


_SHAPES = ("rectangle", "cube", "mesh", "obj", "ply")
_BSDFS = ("diffuse", "conductor", "dielectric", "dispersive", "twosided")


def scene_data_from_dict(d: dict):
    """The GPU-free half of load_dict('scene'): flat geometry/material tables (SceneData) plus the sensor and
    integrator sub-dictionaries. Also what the tests hand to the CPU oracle."""
    data = SceneData()
    named_bsdf, sensor_dicts, integ = {}, [], None
    default_bsdf = None
    for key, v in d.items():   # named BSDFs first: they are the targets of {'type': 'ref'}
        if isinstance(v, dict) and v.get("type") in _BSDFS:
            named_bsdf[v.get("id", key)] = data.add_bsdf(key, bsdf_row(v))
    for key, v in d.items():
        if not isinstance(v, dict):
            continue
        vt = v.get("type")
        if vt in ("perspective", "realisticlens"):
            sensor_dicts.append(v)
        elif vt == "path":
            integ = v
        elif vt in _SHAPES:
            tw = v.get("to_world")
            tw = tw if isinstance(tw, Transform4f) else Transform4f(tw)
            if vt == "rectangle":
                verts, faces = _rect_mesh(tw)
            elif vt == "cube":
                verts, faces = _cube_mesh(tw)
            else:
                if vt == "mesh":
                    verts, faces = np.asarray(v["vertices"], f32), np.asarray(v["faces"], np.int32)
                else:
                    verts, faces = (load_obj if vt == "obj" else load_ply)(v["filename"])
                verts = tw.transform_affine(verts)
            if v.get("flip_normals", False):
                faces = faces[:, ::-1]
            b = v.get("bsdf")
            if b is None:
                if default_bsdf is None:
                    default_bsdf = data.add_bsdf("__default", bsdf_row({"type": "diffuse"}))
                bid = default_bsdf
            elif b.get("type") == "ref":
                bid = named_bsdf[b["id"]]
            else:
                bid = data.add_bsdf(f"{key}.bsdf", bsdf_row(b))
            em = v.get("emitter")
            emitter = None
            if em is not None:
                if em.get("type") != "area" or vt != "rectangle":
                    raise ValueError("this build supports `area` emitters on `rectangle` shapes")
                emitter = {"radiance": _rgb(em.get("radiance"), [1, 1, 1]), "center": tw.transform_affine(np.zeros(3, f32)),
                           "ex": tw.transform_vector(np.array([1, 0, 0], f32)), "ey": tw.transform_vector(np.array([0, 1, 0], f32))}
            data.add_shape(key, verts, faces, bid, emitter)
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
    if t == "path":
        return PathIntegrator(d.get("max_depth", -1), d.get("rr_depth", 5), d.get("hide_emitters", False))
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
