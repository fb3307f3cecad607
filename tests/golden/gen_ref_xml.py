"""Generates tests/golden/ref_xml.npz with the REFERENCE's own scalar_rgb build (tools/build_reference.sh pinned): the XML scenes
under tests/golden/assets/xml/ loaded by the reference's `mi.load_file` and rendered, plus function-level vectors for what those
scenes add to the path — analytic spheres (sphere.cpp), the constant environment emitter (constant.cpp), the procedural
textures (checkerboard.cpp and the fork's checkerlines / comb2d / polarlines). Run in the build container only:

    R=baseline/_ref/pinned/python; LD_LIBRARY_PATH=$R/mitsuba:$R/drjit PYTHONPATH=$R python tests/golden/gen_ref_xml.py

Nothing of the reference is stored except its outputs."""
import os

import mitsuba as mi   # must be imported before numpy

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
XML = os.path.join(HERE, "assets", "xml")
f32 = np.float32
A = lambda x: np.array(x, dtype=f32)   # noqa: E731

RENDERS = {"room": [{}], "charts": [{}, {"hide": "true", "spp": "4"}, {"lens": "gauss", "spp": "4"}], "studio": [{}]}


def tag(kw):
    return "".join(f"@{k}={v}" for k, v in sorted(kw.items()))


def emitter_name(e):
    return e.shape().id() if e.shape() is not None else "__env"


def interactions(scene, o, dd):
    n = len(o)
    rows = np.zeros((n, 25), f32)  # t, p3, n3, sh_n3, sh_s3, sh_t3, uv2, wi3 | shape index | reflectance rgb (diffuse shapes)
    shape_ids = [s.id() for s in scene.shapes()]
    for i in range(n):
        ray = mi.Ray3f([float(x) for x in o[i]], [float(x) for x in dd[i]])
        si = scene.ray_intersect(ray)
        if not si.is_valid():
            rows[i, 0] = np.inf; rows[i, 21] = -1
            continue
        rows[i, :21] = [si.t, *si.p, *si.n, *si.sh_frame.n, *si.sh_frame.s, *si.sh_frame.t, *si.uv, *si.wi]
        rows[i, 21] = shape_ids.index(si.shape.id())
        bsdf = si.bsdf()
        if "diffuse" in bsdf.class_().name().lower():
            rows[i, 22:25] = [*bsdf.eval_diffuse_reflectance(si)]
    return rows, np.array(shape_ids)


def main():
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Warn)
    out = {}
    rng = np.random.default_rng(77)
    for name, variants in RENDERS.items():
        path = os.path.join(XML, name + ".xml")
        for kw in variants:
            scene = mi.load_file(path, **kw)
            img = mi.render(scene, seed=0)
            out[f"{name}{tag(kw)}/image"] = A(img)
            print(name, kw, np.array(img).shape, float(np.array(img).mean()))
        scene = mi.load_file(path)
        out[name + "/shape_ids"] = np.array([s.id() for s in scene.shapes()])
        out[name + "/emitter_ids"] = np.array([emitter_name(e) for e in scene.emitters()])
        params = mi.traverse(scene)
        for k in params.keys():
            if k.endswith("to_world"):
                out[f"{name}/param/{k}"] = A(params[k].matrix)
        s = scene.sensors()[0]
        out[name + "/film_size"] = np.array(list(s.film().size()), np.int32)
        out[name + "/spp"] = np.array(s.sampler().sample_count(), np.int32)

    # ---- PathIntegrator::sample for individual camera rays, each with its own sampler seed (1000 + i)
    smp = mi.load_dict({"type": "independent"})
    for name in RENDERS:
        scene = mi.load_file(os.path.join(XML, name + ".xml"))
        sensor, integ = scene.sensors()[0], scene.integrator()
        n = 400
        o, d, L = np.zeros((n, 3), f32), np.zeros((n, 3), f32), np.zeros((n, 4), f32)
        for i in range(n):
            pos, ap = rng.random(2), rng.random(2)
            ray, _ = sensor.sample_ray(0.0, 0.0, [float(pos[0]), float(pos[1])], [float(ap[0]), float(ap[1])])
            o[i], d[i] = ray.o, ray.d
            smp.seed(1000 + i)
            res = integ.sample(scene, smp, mi.RayDifferential3f(ray), None)
            L[i, :3], L[i, 3] = res[0], float(res[1])
        out[name + "/path_o"], out[name + "/path_d"], out[name + "/path_L"] = o, d, L

    # ---- spheres + procedural textures: rays into charts.xml
    scene = mi.load_file(os.path.join(XML, "charts.xml"))
    n = 4096
    o = np.tile(A([0.0, 0.0, 2.5]), (n, 1)) + rng.standard_normal((n, 3)).astype(f32) * 0.2
    tgt = np.stack([rng.uniform(-2.0, 2.0, n), rng.uniform(-2.0, 2.0, n), np.full(n, -2.0)], 1).astype(f32)
    tgt[: n // 4] = A([0, 0, -1.2]) + rng.standard_normal((n // 4, 3)).astype(f32) * 0.2      # a quarter aimed at the sphere
    dd = tgt - o; dd = (dd / np.linalg.norm(dd, axis=1, keepdims=True)).astype(f32)
    out["charts/isect_o"], out["charts/isect_d"] = o, dd
    out["charts/isect"], out["charts/isect_shape_ids"] = interactions(scene, o, dd)
    # rays that start inside the sphere, and shadow-style rays with a finite maxt (ray_test)
    o2 = (A([0, 0, -1.2]) + rng.uniform(-0.15, 0.15, (512, 3))).astype(f32)
    d2 = rng.standard_normal((512, 3)).astype(f32); d2 = (d2 / np.linalg.norm(d2, axis=1, keepdims=True)).astype(f32)
    out["charts/inside_o"], out["charts/inside_d"] = o2, d2
    out["charts/inside"], _ = interactions(scene, o2, d2)
    maxt = rng.uniform(0.5, 4.5, n).astype(f32)
    occ = np.zeros(n, np.int32)
    for i in range(n):
        ray = mi.Ray3f([float(x) for x in o[i]], [float(x) for x in dd[i]]); ray.maxt = float(maxt[i])
        occ[i] = int(scene.ray_test(ray))
    out["charts/test_maxt"], out["charts/test_occluded"] = maxt, occ

    # ---- environment + mesh emitter: emitter sampling in studio.xml
    scene = mi.load_file(os.path.join(XML, "studio.xml"))
    em_ids = [emitter_name(e) for e in scene.emitters()]
    n = 2048
    ref_p = np.stack([rng.uniform(-1.5, 1.5, n), rng.uniform(0.0, 1.5, n), rng.uniform(-1.5, 1.5, n)], 1).astype(f32)
    ref_p[:64] *= 40.0           # reference points outside the bounding sphere (constant.cpp:121-122)
    smp = rng.random((n, 2)).astype(f32)
    rows = np.zeros((n, 16), f32)
    for i in range(n):
        it = mi.SurfaceInteraction3f(); it.p = [float(x) for x in ref_p[i]]; it.n = [0, 0, 1]; it.t = 1.0
        ds, w = scene.sample_emitter_direction(it, [float(smp[i, 0]), float(smp[i, 1])], False)
        pdf2 = scene.pdf_emitter_direction(it, ds) if ds.pdf > 0 else 0.0
        rows[i] = [*ds.p, *ds.n, *ds.d, ds.dist, ds.pdf, *w, pdf2, em_ids.index(emitter_name(ds.emitter)) if ds.pdf > 0 else -1]
    out["studio/em_ref_p"], out["studio/em_sample"], out["studio/em_dir"] = ref_p, smp, rows
    np.savez_compressed(os.path.join(HERE, "ref_xml.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
