"""Generates tests/golden/ref_disk.npz with the REFERENCE's own scalar_rgb build (tools/build_reference.sh pinned): the `disk`
shape (src/shapes/disk.cpp) as the notebooks use it — the carrier of a `directionalarea` field source
(design_problem.py:355-371) — in tests/ref_scenes.py path_b_scene(source="disk"):

  * `isect`: Scene::ray_intersect of rays aimed at and around the (elliptic, tilted) disk, from both sides — t, p, normals,
    shading frame, uv, wi, shape (the layout of ref_xml.npz's charts/isect);
  * `image_seed{0,3}`: the particle tracer's renders (Disk::sample_position through DirectionalArea::sample_ray);
  * `em_*`, `lamp_image_seed{0,3}`: an `area` light on a disk (ref_scenes.disk_light_scene) — Scene::sample_emitter_direction /
    pdf_emitter_direction on 2048 reference points and the path tracer's renders.

    R=baseline/_ref/pinned/python; LD_LIBRARY_PATH=$R/mitsuba:$R/drjit PYTHONPATH=$R python tests/golden/gen_ref_disk.py"""
import os
import sys

import mitsuba as mi   # must be imported before numpy

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import ref_scenes as rs  # noqa: E402
from gen_ref_xml import interactions  # noqa: E402

f32 = np.float32


def main():
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Error)
    T = lambda m: mi.ScalarTransform4f(np.asarray(m, f32))   # noqa: E731
    out = {"n_threads": np.array(os.cpu_count(), np.int32)}
    scene = mi.load_dict(rs.path_b_scene(T, source="disk"))
    rng = np.random.default_rng(5)
    n = 2048
    c = np.array([0.0008, 0.0, 0.08], f32)
    o = np.stack([rng.uniform(-0.01, 0.01, n), rng.uniform(-0.01, 0.01, n), rng.uniform(0.03, 0.075, n)], 1).astype(f32)
    o[n // 2:, 2] = rng.uniform(0.085, 0.12, n - n // 2).astype(f32)            # half from behind the source
    tgt = c + np.stack([rng.uniform(-0.006, 0.006, n), rng.uniform(-0.005, 0.005, n), np.zeros(n)], 1).astype(f32)
    dd = tgt - o; dd = (dd / np.linalg.norm(dd, axis=1, keepdims=True)).astype(f32)
    out["isect_o"], out["isect_d"] = o, dd
    out["isect"], ids = interactions(scene, o, dd)
    k = list(ids).index("field_source")
    print("rays on the disk:", int((out["isect"][:, 21] == k).sum()), "of", n, "| misses", int((~np.isfinite(out["isect"][:, 0])).sum()))
    for seed in (0, 3):
        img = np.array(mi.render(scene, seed=seed), dtype=f32)
        out[f"image_seed{seed}"] = img
        print("seed", seed, img.shape, float(img.mean()), float(img.max()), "lit pixels", int((img.sum(-1) > 0).sum()))
    # ---- an `area` light on a disk: Scene::sample_emitter_direction / pdf_emitter_direction and the path tracer's render
    scene = mi.load_dict(rs.disk_light_scene(T))
    emitters = list(scene.emitters()); em_ids = [e.shape().id() for e in emitters]
    n = 2048
    ref_p = np.stack([rng.uniform(-1.8, 1.8, n), rng.uniform(0.0, 2.4, n), rng.uniform(-1.4, 1.8, n)], 1).astype(f32)
    smp = rng.random((n, 2)).astype(f32)
    rows = np.zeros((n, 16), f32)
    for i in range(n):
        it = mi.SurfaceInteraction3f(); it.p = [float(x) for x in ref_p[i]]; it.n = [0, 0, 1]; it.t = 1.0
        ds, w = scene.sample_emitter_direction(it, [float(smp[i, 0]), float(smp[i, 1])], False)
        pdf2 = scene.pdf_emitter_direction(it, ds) if ds.pdf > 0 else 0.0
        rows[i] = [*ds.p, *ds.n, *ds.d, ds.dist, ds.pdf, *w, pdf2, em_ids.index(ds.emitter.shape().id()) if ds.pdf > 0 else -1]
    out["em_ref_p"], out["em_sample"], out["em_dir"], out["em_ids"] = ref_p, smp, rows, np.array(em_ids)
    print("emitter samples: on the disk", int((rows[:, 15] == em_ids.index("lamp")).sum()), "valid", int((rows[:, 15] >= 0).sum()))
    for seed in (0, 3):
        img = np.array(mi.render(scene, seed=seed), dtype=f32)
        out[f"lamp_image_seed{seed}"] = img
        print("lamp seed", seed, float(img.mean()), float(img.max()))
    np.savez_compressed(os.path.join(HERE, "ref_disk.npz"), **out)


if __name__ == "__main__":
    main()
