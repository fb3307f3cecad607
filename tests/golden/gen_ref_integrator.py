"""Generates tests/golden/ref_integrator.npz and tests/golden/ref_renders.npz with the REFERENCE's own scalar_rgb build
(tools/build_reference.sh pinned -> baseline/_ref/pinned). Run in the build container only:

    R=baseline/_ref/pinned/python; LD_LIBRARY_PATH=$R/mitsuba:$R/drjit PYTHONPATH=$R python tests/golden/gen_ref_integrator.py

Function-level vectors (perspective sample_ray, BSDF sample / eval_pdf, emitter sampling, ray_intersect with the full
surface interaction, bitmap texture lookups) pin rows P0 / B1 / I2 / T1 of SURVEY.md 8(a); the renders pin I1 and the scalar
seeding convention R3 (integrator.cpp:195-200,388-402). Nothing of the reference is stored except its outputs."""
import importlib.util
import os
import sys

import mitsuba as mi   # must be imported before numpy (SURVEY.md 8c gotcha)

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ref_scenes as rs  # noqa: E402

f32 = np.float32
T = lambda m: mi.ScalarTransform4f(np.asarray(m, f32))   # noqa: E731


def _imageio():
    spec = importlib.util.spec_from_file_location("_lrt_imageio", os.path.join(HERE, "..", "..", "mitsuba3-camera_b200", "imageio.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    return mod


def write_assets():
    os.makedirs(rs.ASSETS, exist_ok=True)
    v, f, n, uv = rs.icosphere(2)
    rs.write_ply(rs.MESH_SMOOTH, v, f, n, uv)
    rs.write_ply(rs.MESH_PLAIN, v, f)
    _imageio().write_exr(rs.TEXTURE, rs.texture_array())


def A(x):
    return np.array(x, dtype=f32)


def gen_functions(out):
    rng = np.random.default_rng(2024)
    # ---- P0: perspective sample_ray (two sensors: Cornell's, and a cropped non-square one with fov_axis x)
    sens = {"persp_cbox": {"type": "perspective", "fov_axis": "smaller", "near_clip": 0.001, "far_clip": 100.0, "fov": 39.3077,
                           "to_world": T(rs.m_look_at([0, 0, 3.9], [0, 0, 0], [0, 1, 0])),
                           "film": {"type": "hdrfilm", "width": 256, "height": 256}},
            "persp_crop": {"type": "perspective", "fov": 50.0, "near_clip": 0.05, "far_clip": 500.0,
                           "to_world": T(rs.m_look_at([1, 2, 3], [0.2, 0.1, -0.4], [0, 1, 0])),
                           "film": {"type": "hdrfilm", "width": 320, "height": 200, "crop_width": 100, "crop_height": 60,
                                    "crop_offset_x": 37, "crop_offset_y": 21}}}
    pos = rng.random((1024, 2)).astype(f32)
    out["persp_pos"] = pos
    for name, d in sens.items():
        s = mi.load_dict(d)
        rows = np.zeros((len(pos), 7), f32)
        for i, p in enumerate(pos):
            r, w = s.sample_ray(0.0, 0.0, [float(p[0]), float(p[1])], [0.5, 0.5])
            rows[i] = [*r.o, *r.d, r.maxt]
        out[name] = rows
        out[name + "_to_world"] = A(d["to_world"].matrix)

    # ---- B1: BSDF sample / eval_pdf in the local frame
    n = 1024
    wi = rng.standard_normal((n, 3)); wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    wi[:16, 2] = np.abs(wi[:16, 2]) * 1e-3            # grazing
    wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    wo = rng.standard_normal((n, 3)); wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    s1, s2 = rng.random(n).astype(f32), rng.random((n, 2)).astype(f32)
    out["bsdf_wi"], out["bsdf_wo"], out["bsdf_s1"], out["bsdf_s2"] = wi.astype(f32), wo.astype(f32), s1, s2
    bsdfs = {"diffuse": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.5, 0.3, 0.2]}},
             "conductor": {"type": "conductor", "material": "Cu",
                           "specular_reflectance": {"type": "rgb", "value": [0.9, 0.8, 0.7]}},
             "dielectric": {"type": "dielectric", "int_ior": 1.5046, "ext_ior": 1.000277,
                            "specular_reflectance": {"type": "rgb", "value": [0.9, 0.8, 0.7]},
                            "specular_transmittance": {"type": "rgb", "value": [0.6, 0.7, 0.8]}},
             "twosided": {"type": "twosided", "material": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.5, 0.3, 0.2]}}},
             "dispersive": {"type": "dispersive", "int_ior": 1.62, "int_V_d": 36.0, "ext_ior": 1.0, "ext_V_d": 0.0}}
    ctx = mi.BSDFContext()
    for name, d in bsdfs.items():
        try:
            b = mi.load_dict(d)
        except Exception as e:   # a plugin the build lacks must not take the other vectors down
            print("skipping bsdf", name, e); continue
        rows = np.zeros((n, 13), f32)   # eval rgb, pdf, sample wo xyz, spdf, eta, weight rgb, sampled_type
        for i in range(n):
            si = mi.SurfaceInteraction3f()
            si.wi = [float(x) for x in wi[i].astype(f32)]
            si.sh_frame = mi.Frame3f([0, 0, 1]); si.n = [0, 0, 1]; si.t = 1.0
            si.wavelengths = []
            val, pdf = b.eval_pdf(ctx, si, [float(x) for x in wo[i].astype(f32)])
            bs, wgt = b.sample(ctx, si, float(s1[i]), [float(s2[i, 0]), float(s2[i, 1])])
            rows[i] = [*val, pdf, *bs.wo, bs.pdf, bs.eta, *wgt, float(bs.sampled_type)]
        out["bsdf_" + name] = rows

    # ---- T1 + I2 + textures: ray_intersect with the full interaction, emitter sampling, texture lookups
    d, _ = rs.scene_dicts(T)["textured"]
    d = dict(d)
    d["lamp2"] = {"type": "rectangle", "to_world": T(rs.mm(rs.m_translate([2.5, 1.0, 0]), rs.m_rotate([0, 1, 0], -90), rs.m_scale([0.5, 0.7, 1]))),
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [1.0, 2.0, 3.0]}}}
    scene = mi.load_dict(d)
    n = 2048
    o = np.tile(A([0.3, 1.4, 3.2]), (n, 1)) + rng.standard_normal((n, 3)).astype(f32) * 0.05
    tgt = rng.uniform(-1.5, 1.5, (n, 3)).astype(f32); tgt[:, 1] = rng.uniform(-0.2, 1.2, n)
    dd = tgt - o; dd = (dd / np.linalg.norm(dd, axis=1, keepdims=True)).astype(f32)
    out["isect_o"], out["isect_d"] = o, dd
    rows = np.zeros((n, 25), f32)  # t, p3, n3, sh_n3, sh_s3, sh_t3, uv2, wi3 | shape index | reflectance rgb (diffuse shapes)
    shapes = list(scene.shapes()); shape_ids = [s.id() for s in shapes]
    em_s = rng.random((n, 2)).astype(f32)
    out["em_sample"] = em_s
    em_rows = np.zeros((n, 16), f32)  # ds.p3 n3 d3 dist pdf weight3 pdf_emitter_direction | em index
    emitters = list(scene.emitters()); em_ids = [e.shape().id() for e in emitters]
    for i in range(n):
        ray = mi.Ray3f([float(x) for x in o[i]], [float(x) for x in dd[i]])
        si = scene.ray_intersect(ray)
        if not si.is_valid():
            rows[i, 0] = np.inf; rows[i, 21] = -1
            continue
        rows[i, :21] = [si.t, *si.p, *si.n, *si.sh_frame.n, *si.sh_frame.s, *si.sh_frame.t, *si.uv, *si.wi]
        rows[i, 21] = shape_ids.index(si.shape.id())
        bsdf = si.bsdf()
        if "Diffuse" in str(type(bsdf)) or "diffuse" in bsdf.class_().name().lower():
            val = bsdf.eval_diffuse_reflectance(si)
            rows[i, 22:25] = [*val]
        ds, w = scene.sample_emitter_direction(si, [float(em_s[i, 0]), float(em_s[i, 1])], False)
        pdf2 = scene.pdf_emitter_direction(si, ds) if ds.pdf > 0 else 0.0
        em_rows[i] = [*ds.p, *ds.n, *ds.d, ds.dist, ds.pdf, *w, pdf2, em_ids.index(ds.emitter.shape().id()) if ds.pdf > 0 else -1]
    out["isect"] = rows
    out["isect_shape_ids"] = np.array(shape_ids)
    out["emitter_dir"] = em_rows
    out["emitter_ids"] = np.array(em_ids)

    # bitmap texture alone (repeat / clamp / mirror), bilinear
    uv = (rng.random((1024, 2)) * 3 - 1).astype(f32)
    out["tex_uv"] = uv
    for wrap in ("repeat", "clamp", "mirror"):
        tex = mi.load_dict({"type": "bitmap", "filename": rs.TEXTURE, "filter_type": "bilinear", "wrap_mode": wrap, "raw": True})
        vals = np.zeros((len(uv), 3), f32)
        for i in range(len(uv)):
            si = mi.SurfaceInteraction3f(); si.uv = [float(uv[i, 0]), float(uv[i, 1])]
            vals[i] = [*tex.eval(si)]
        out["tex_" + wrap] = vals

    # ---- I2 on general meshes: emitter sampling in the mesh_lights scene (cube with normals, smooth icosphere, rectangle)
    d, _ = rs.scene_dicts(T)["mesh_lights"]
    scene = mi.load_dict(d)
    emitters = list(scene.emitters()); em_ids = [e.shape().id() for e in emitters]
    n = 2048
    ref_p = rng.uniform(-0.9, 0.9, (n, 3)).astype(f32)
    smp = rng.random((n, 2)).astype(f32)
    rows = np.zeros((n, 16), f32)
    for i in range(n):
        it = mi.SurfaceInteraction3f(); it.p = [float(x) for x in ref_p[i]]; it.n = [0, 0, 1]; it.t = 1.0
        ds, w = scene.sample_emitter_direction(it, [float(smp[i, 0]), float(smp[i, 1])], False)
        pdf2 = scene.pdf_emitter_direction(it, ds) if ds.pdf > 0 else 0.0
        rows[i] = [*ds.p, *ds.n, *ds.d, ds.dist, ds.pdf, *w, pdf2, em_ids.index(ds.emitter.shape().id()) if ds.pdf > 0 else -1]
    out["mesh_em_ref_p"], out["mesh_em_sample"], out["mesh_em_dir"], out["mesh_em_ids"] = ref_p, smp, rows, np.array(em_ids)


def matrices_of(d, prefix, store):
    for k, v in d.items():
        if isinstance(v, dict):
            matrices_of(v, prefix + k + "/", store)
        elif isinstance(v, mi.ScalarTransform4f):
            store[prefix + k] = A(v.matrix)


def gen_renders(out):
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Warn)
    scenes = dict(rs.scene_dicts(T))
    cb = mi.cornell_box()
    cb["integrator"]["max_depth"] = 6
    scenes["cbox"] = (cb, 64)     # BASELINE.json configs[0]: mi.cornell_box() verbatim, 256^2, 64 spp, max_depth 6
    for name, (d, spp) in scenes.items():
        matrices_of(d, name + "/", out)
        scene = mi.load_dict(d)
        img = mi.render(scene, spp=spp, seed=0)
        out[name + "/image"] = A(img)
        img1 = mi.render(scene, spp=spp, seed=3)
        out[name + "/image_seed3_mean"] = A(np.array(img1).mean(axis=(0, 1)))
        print(name, np.array(img).shape, np.array(img).mean())
    # a 1-spp render exercises the per-pixel seeding alone (one sample per stream)
    scene = mi.load_dict(cb)
    out["cbox/image_1spp"] = A(mi.render(scene, spp=1, seed=0))


if __name__ == "__main__":
    write_assets()
    fn, rn = {}, {}
    gen_functions(fn)
    np.savez_compressed(os.path.join(HERE, "ref_integrator.npz"), **fn)
    gen_renders(rn)
    np.savez_compressed(os.path.join(HERE, "ref_renders.npz"), **rn)
    print("wrote", len(fn), "+", len(rn), "arrays")
