"""ctypes access to the CHECKERS: oracle/liboracle.so (our CPU restatement) and, when it was built in
the build container, oracle/_ref/libreflens.so (the reference's own lens/RNG code). Test-only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
sys.path.insert(0, ROOT)

f32 = np.float32
c_f = C.c_float
FP = C.POINTER(C.c_float)
IP = C.POINTER(C.c_int)
UP = C.POINTER(C.c_uint32)


def fp(a):
    return a.ctypes.data_as(FP)


def ip(a):
    return a.ctypes.data_as(IP)


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.startswith("oracle") and f.endswith((".cpp", ".h"))]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "port"], stdout=subprocess.DEVNULL)
    return so


_ORC = None
_REF = None


def load_lens_designs():
    """mitsuba3-camera_b200/lens_designs.py as a standalone module (pure numpy; no CUDA library needed)."""
    import importlib.util
    name = "_lrt_lens_designs"
    if name not in sys.modules:
        spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "mitsuba3-camera_b200", "lens_designs.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
    return sys.modules[name]


def oracle():
    global _ORC
    if _ORC is None:
        L = C.CDLL(build_oracle())
        L.orc_tea_float32.restype = c_f
        L.orc_material_ior.restype = c_f
        L.orc_cam_from_prescription.restype = C.c_void_p
        L.orc_cam_from_interfaces.restype = C.c_void_p
        L.orc_cam_focus_delta.restype = c_f
        for n in ("orc_cam_destroy", "orc_cam_num_interfaces", "orc_cam_focus_delta", "orc_cam_sample_to_film",
                  "orc_cam_film_to_sample", "orc_cam_interface", "orc_cam_sample_ray", "orc_cam_trace"):
            getattr(L, n).argtypes = None
        _ORC = L
    return _ORC


def ref():
    """The reference's own code (None on the GPU box if oracle/_ref was not shipped)."""
    global _REF
    if _REF is None:
        so = os.path.join(ORACLE_DIR, "_ref", "libreflens.so")
        if not os.path.exists(so):
            if os.path.isdir("/root/reference"):
                subprocess.check_call(["make", "-C", ORACLE_DIR, "ref"], stdout=subprocess.DEVNULL)
            else:
                return None
        L = C.CDLL(so)
        L.ref_tea_float32.restype = c_f
        L.ref_material_ior.restype = c_f
        L.ref_cam_create.restype = C.c_void_p
        L.ref_cam_film_z.restype = c_f
        _REF = L
    return _REF


EYE = np.eye(4, dtype=f32)


class FilmSpec:
    def __init__(self, width=256, height=256, crop=None, diagonal_mm=35.0):
        self.width, self.height = width, height
        self.crop = np.array(crop if crop else [width, height, 0, 0], np.int32)
        self.diagonal_mm = diagonal_mm


def orc_camera(design, film: FilmSpec, to_world=EYE, near=1e-2, far=1e4, object_distance=6.0):
    """design: lens_designs.Prescription | InterfaceList -> opaque oracle camera handle."""
    from importlib import import_module
    L = oracle()
    tw = np.ascontiguousarray(to_world, f32)
    if hasattr(design, "rows"):
        rows = np.ascontiguousarray(design.rows, f32)
        return C.c_void_p(L.orc_cam_from_interfaces(
            C.c_int(len(rows)), fp(rows), c_f(design.film_z), C.c_int(int(design.fine_focus)), c_f(object_distance),
            C.c_int(film.width), C.c_int(film.height), ip(film.crop), c_f(film.diagonal_mm), fp(tw), c_f(near), c_f(far)))
    p = design
    return C.c_void_p(L.orc_cam_from_prescription(
        C.c_int(len(p.kinds)), ip(p.kinds), fp(p.thickness), fp(p.radius), fp(p.semi_diameter), fp(p.kappa),
        fp(np.ascontiguousarray(p.ai)), fp(np.ascontiguousarray(p.materials)), c_f(p.scale), c_f(p.stop_ratio),
        C.c_int(int(p.fine_focus)), c_f(object_distance), C.c_int(film.width), C.c_int(film.height), ip(film.crop),
        c_f(film.diagonal_mm), fp(tw), c_f(near), c_f(far)))


def ref_camera(name, film: FilmSpec, to_world=EYE, near=1e-2, far=1e4, props=None, shutter_open_time=0.0):
    L = ref()
    props = props or {}
    names = (C.c_char_p * len(props))(*[k.encode() for k in props])
    vals = (C.c_double * len(props))(*[float(v) for v in props.values()])
    tw = np.ascontiguousarray(to_world, f32)
    h = L.ref_cam_create(name.encode(), C.c_int(film.width), C.c_int(film.height), ip(film.crop),
                         c_f(film.diagonal_mm), fp(tw), c_f(near), c_f(far), c_f(shutter_open_time), C.c_int(len(props)), names, vals)
    return C.c_void_p(h) if h else None


def cam_exit_pupil(L, prefix, h, shutter_open_time=1.0):
    """Enable exit-pupil sampling on a reference ('ref') or oracle ('orc') camera; returns the 64 x 4 LUT."""
    out = np.zeros((64, 4), f32)
    if prefix == "ref":
        L.ref_cam_compute_exit_pupil(h, fp(out))
    else:
        L.orc_cam_set_exit_pupil(h, C.c_int(1), c_f(shutter_open_time))
        L.orc_cam_get_exit_pupil(h, fp(out))
    return out


def radical_inverse(L, prefix, base_index, indices):
    f = getattr(L, prefix + "_radical_inverse"); f.restype = c_f
    return np.array([f(C.c_int(base_index), C.c_uint64(int(i))) for i in indices], f32)


def cam_interfaces(L, prefix, h):
    n = getattr(L, prefix + "_cam_num_interfaces")(h)
    out = np.zeros((n, 16), f32)
    for i in range(n):
        getattr(L, prefix + "_cam_interface")(h, C.c_int(i), fp(out[i]))
    return out


def cam_sample_ray(L, prefix, h, pos, ap):
    n = len(pos)
    out = np.zeros((n, 10), f32)
    getattr(L, prefix + "_cam_sample_ray")(h, C.c_int(n), fp(np.ascontiguousarray(pos, f32)),
                                           fp(np.ascontiguousarray(ap, f32)), fp(out))
    return out


def cam_trace(L, prefix, h, od, from_world):
    n = len(od)
    out = np.zeros((n, 7), f32)
    getattr(L, prefix + "_cam_trace")(h, C.c_int(int(from_world)), C.c_int(n), fp(np.ascontiguousarray(od, f32)), fp(out))
    return out


def bits(a):
    return np.ascontiguousarray(a, f32).view(np.uint32)


# ------------------------------------------------------------------------------------------------ scenes / render
def orc_scene(a, analytic_rectangles=False, hide_emitters=False):
    """a: dict of flat arrays (mitsuba3-camera_b200/scene.py SceneData.arrays()) -> oracle scene handle.
    analytic_rectangles: intersect `rectangle` shapes as the reference does (rectangle.cpp) instead of as two triangles.
    Analytic spheres, procedural textures and the environment emitter of `a` are registered too."""
    L = oracle()
    L.orc_scene_create2.restype = C.c_void_p
    keep = {k: np.ascontiguousarray(v) for k, v in a.items() if isinstance(v, np.ndarray)}
    flags = keep.get("shape_flags")
    has_attr = flags is not None and flags.size and bool(flags.any())
    tex = a.get("textures") or []
    dims = np.array([[t["rgb"].shape[1], t["rgb"].shape[0], t["wrap"]] for t in tex], np.int32).reshape(-1, 3)
    tdata = np.concatenate([t["rgb"].reshape(-1) for t in tex]).astype(f32) if tex else np.zeros(1, f32)
    tuv = np.concatenate([np.asarray(t["to_uv"], f32) for t in tex]) if tex else np.zeros(6, f32)
    h = C.c_void_p(L.orc_scene_create2(
        C.c_int(len(keep["vertices"])), fp(keep["vertices"]), fp(keep["vertex_normals"]) if has_attr else None,
        fp(keep["vertex_uvs"]) if has_attr else None, C.c_int(len(keep["indices"])), ip(keep["indices"]), ip(keep["tri_shape"]),
        C.c_int(len(keep["shape_bsdf"])), ip(keep["shape_bsdf"]), ip(keep["shape_emitter"]), ip(keep["shape_flags"]) if has_attr else None,
        C.c_int(len(keep["bsdfs"])), fp(keep["bsdfs"]), C.c_int(len(keep["emitters"])), fp(keep["emitters"]),
        C.c_int(len(tex)), ip(dims), fp(tdata), fp(tuv)))
    if analytic_rectangles and len(keep.get("rect_shapes", [])):
        tw = keep["rect_to_world"].astype(f32)
        inv = keep["rect_to_object"].astype(f32) if "rect_to_object" in keep else np.stack([np.linalg.inv(m.astype(np.float64)) for m in tw]).astype(f32)
        L.orc_scene_set_rectangles(h, C.c_int(len(tw)), ip(keep["rect_shapes"]), fp(np.ascontiguousarray(tw[:, :3, :].reshape(-1))),
                                   fp(np.ascontiguousarray(inv[:, :3, :].reshape(-1))))
    sph = keep.get("spheres")
    if sph is not None and len(sph):
        L.orc_scene_set_spheres(h, C.c_int(len(sph)), ip(np.ascontiguousarray(sph[:, 4].astype(np.int32))),
                                fp(np.ascontiguousarray(np.concatenate([sph[:, 0:4], sph[:, 5:6], sph[:, 6:7], np.zeros((len(sph), 2), f32)], 1), f32)),
                                fp(np.ascontiguousarray(sph[:, 8:20])), fp(np.ascontiguousarray(sph[:, 20:32])))
    if any(t.get("kind", 0) for t in tex):
        kinds = np.array([t.get("kind", 0) for t in tex], np.int32)
        par = np.zeros((len(tex), 8), f32)
        for i, t in enumerate(tex):
            if t.get("kind", 0):
                par[i, 0:3], par[i, 3:6], par[i, 6], par[i, 7] = t["color0"], t["color1"], t["linewidth"], t["spokes"]
        L.orc_scene_set_procedural(h, C.c_int(len(tex)), ip(kinds), fp(par))
    if hide_emitters:
        L.orc_scene_set_hide_emitters(h, C.c_int(1))
    return h


def orc_intersect(scene_h, o, d, maxt, any_hit=False):
    L = oracle()
    n = len(o)
    od = np.ascontiguousarray(np.concatenate([o, d, np.broadcast_to(np.asarray(maxt, f32).reshape(-1, 1), (n, 1))], 1), f32)
    out = np.zeros((n, 4), f32)
    L.orc_scene_intersect(scene_h, C.c_int(int(any_hit)), C.c_int64(n), fp(od), fp(out))
    return out


def orc_render(scene_h, film, spp, seed=0, max_depth=8, rr_depth=5, cam_h=None, fov=45.0, to_world=EYE, near=1e-2, far=1e4,
               filter_kind=2, filter_param=0.0, idx_range=None):
    """film: FilmSpec. Returns (storage float64 [ch, cw, 4], lens_miss_count)."""
    L = oracle()
    cw, ch, cx, cy = (int(v) for v in film.crop)
    st = np.zeros((ch, cw, 4), np.float64)
    miss = C.c_int64(0)
    tw = np.ascontiguousarray(to_world, f32)
    total = cw * ch * spp
    b, e = idx_range if idx_range else (0, total)
    L.orc_render(scene_h, cam_h, c_f(fov), fp(tw), c_f(near), c_f(far), C.c_int(film.width), C.c_int(film.height),
                 C.c_int(cw), C.c_int(ch), C.c_int(cx), C.c_int(cy), C.c_int(filter_kind), c_f(filter_param), C.c_uint32(seed),
                 C.c_int(spp), C.c_int(max_depth), C.c_int(rr_depth), C.c_int64(b), C.c_int64(e),
                 st.ctypes.data_as(C.POINTER(C.c_double)), C.byref(miss))
    return st, miss.value


def orc_render_scalar(scene_h, film, spp, seed=0, base_seed=0, max_depth=8, rr_depth=5, cam_h=None, fov=45.0, to_world=EYE, near=1e-2,
                      far=1e4, filter_kind=2, filter_param=0.0, block_size=32):
    """The reference's SCALAR render loop (per-pixel streams, spiral blocks, Morton order, discretised filter)."""
    L = oracle()
    cw, ch, cx, cy = (int(v) for v in film.crop)
    st = np.zeros((ch, cw, 4), np.float64)
    miss = C.c_int64(0)
    tw = np.ascontiguousarray(to_world, f32)
    L.orc_render_scalar(scene_h, cam_h, c_f(fov), fp(tw), c_f(near), c_f(far), C.c_int(film.width), C.c_int(film.height),
                        C.c_int(cw), C.c_int(ch), C.c_int(cx), C.c_int(cy), C.c_int(filter_kind), c_f(filter_param),
                        C.c_uint32(base_seed), C.c_uint32(seed), C.c_int(spp), C.c_int(max_depth), C.c_int(rr_depth), C.c_int(block_size),
                        st.ctypes.data_as(C.POINTER(C.c_double)), C.byref(miss))
    return st, miss.value


def develop(storage):
    w = storage[..., 3:4]
    return (storage[..., :3] / np.where(w == 0, 1.0, w)).astype(f32)


def rel_rmse(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-12))


# ------------------------------------------------------------------------------------------------ spot path
def orc_spot_render(cam_h, film: FilmSpec, src_p, intensity, seed, spp, filter_kind=1, filter_param=0.0, origin_offset=0.0):
    """float32 checker of lrt_spot_render -> (storage float64 [ch,cw,4], valid samples)."""
    L = oracle()
    L.orc_spot_render.restype = C.c_longlong
    cw, ch, cx, cy = (int(v) for v in film.crop)
    st = np.zeros((ch, cw, 4), np.float64)
    valid = L.orc_spot_render(cam_h, C.c_int(cw), C.c_int(ch), C.c_int(cx), C.c_int(cy), C.c_int(filter_kind), c_f(filter_param),
                              C.c_uint32(seed), C.c_int(spp), fp(np.ascontiguousarray(src_p, f32)),
                              fp(np.ascontiguousarray(intensity, f32)), c_f(origin_offset), st.ctypes.data_as(C.POINTER(C.c_double)))
    return st, valid


def orc_spot_grad(rows, f2s, normalization, near, far, film: FilmSpec, src_p, intensity, seed, spp, G, sel, filter_kind=1,
                  filter_param=0.0, delta=None, origin_offset=0.0):
    """float64 dual-number checker: (Phi, dPhi/d rows[sel]) for Phi = <G, storage(rows)> with frozen masks."""
    L = oracle()
    L.orc_spot_grad.restype = C.c_double
    cw, ch, cx, cy = (int(v) for v in film.crop)
    sel = np.ascontiguousarray(sel, np.int32).reshape(-1, 2)
    out = np.zeros(len(sel), np.float64)
    dl = None if delta is None else np.ascontiguousarray(delta, np.float64)
    phi = L.orc_spot_grad(C.c_int(len(rows)), fp(np.ascontiguousarray(rows, f32)), fp(np.ascontiguousarray(f2s, f32).ravel()),
                          c_f(normalization), c_f(near), c_f(far), C.c_int(cw), C.c_int(ch), C.c_int(cx), C.c_int(cy),
                          C.c_int(filter_kind), c_f(filter_param), C.c_uint32(seed), C.c_int(spp),
                          fp(np.ascontiguousarray(src_p, f32)), fp(np.ascontiguousarray(intensity, f32)), c_f(origin_offset),
                          fp(np.ascontiguousarray(G, f32)), C.c_int(len(sel)), sel.ctypes.data_as(IP),
                          dl.ctypes.data_as(C.POINTER(C.c_double)) if dl is not None else None,
                          out.ctypes.data_as(C.POINTER(C.c_double)))
    return phi, out


# ------------------------------------------------------------------------------------------------ particle tracer
def orc_ptracer_render(scene_h, cam_h, film, spp, to_world, to_camera=None, seed=0, base_seed=0, max_depth=-1, rr_depth=5, hide_emitters=False,
                       shutter_open_time=0.0, filter_kind=2, filter_param=0.0, scalar=True, n_threads=8):
    """AdjointIntegrator::render + ParticleTracerIntegrator (oracle_ptracer_impl.h). Returns storage float64 [ch, cw, 4];
    the image is its RGB channels (particle-tracer films are not divided by a weight)."""
    L = oracle()
    cw, ch, cx, cy = (int(v) for v in film.crop)
    st = np.zeros((ch, cw, 4), np.float64)
    tw = np.ascontiguousarray(to_world, f32)
    tc = None if to_camera is None else np.ascontiguousarray(to_camera, f32)
    L.orc_ptracer_render(scene_h, cam_h, fp(tw), None if tc is None else fp(tc), C.c_int(film.width), C.c_int(film.height), C.c_int(cw), C.c_int(ch), C.c_int(cx), C.c_int(cy),
                         C.c_int(filter_kind), c_f(filter_param), C.c_uint32(base_seed), C.c_uint32(seed), C.c_int(spp), C.c_int(max_depth),
                         C.c_int(rr_depth), C.c_int(int(hide_emitters)), c_f(shutter_open_time), C.c_int(int(scalar)), C.c_int(n_threads),
                         st.ctypes.data_as(C.POINTER(C.c_double)))
    return st
