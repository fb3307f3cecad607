"""Generates tests/golden/ref_ptracer.npz with the REFERENCE's own scalar_rgb build (tools/build_reference.sh pinned): the
particle tracer (`ptracer`, src/integrators/ptracer.cpp) on XML scenes seen through the realistic lens. Run in the build
container only:

    R=baseline/_ref/pinned/python; LD_LIBRARY_PATH=$R/mitsuba:$R/drjit PYTHONPATH=$R python tests/golden/gen_ref_ptracer.py

The scalar AdjointIntegrator seeds one sampler per range of max(total / (4 * n_threads), 1) samples
(src/render/integrator.cpp:600-640), so its images depend on the machine's thread count: it is stored next to the images.
Nothing of the reference is stored except its outputs."""
import os

import mitsuba as mi   # must be imported before numpy

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
XML = os.path.join(HERE, "assets", "xml")
CASES = [("lens_room", {}), ("lens_room", {"max_depth": "2", "lens": "gauss"}), ("lens_room", {"max_depth": "-1", "spp": "4"}),
         ("charts", {"integrator": "ptracer"}), ("charts", {"integrator": "ptracer", "hide": "true", "spp": "4"}),
         ("lens_studio", {}),
         ("room", {"integrator": "ptracer"})]          # (its own PERSPECTIVE sensor: PerspectiveCamera::sample_direction)


def tag(kw):
    return "".join(f"@{k}={v}" for k, v in sorted(kw.items()))


def main():
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Error)
    out = {"n_threads": np.array(os.cpu_count(), np.int32)}
    for name, kw in CASES:
        scene = mi.load_file(os.path.join(XML, name + ".xml"), **kw)
        assert "ParticleTracer" in str(scene.integrator())
        img = np.array(mi.render(scene, seed=0), dtype=np.float32)
        out[f"{name}{tag(kw)}/image"] = img
        img3 = np.array(mi.render(scene, seed=3), dtype=np.float32)
        out[f"{name}{tag(kw)}/image_seed3"] = img3
        print(name, kw, img.shape, float(img.mean()), float(img3.mean()))
    np.savez_compressed(os.path.join(HERE, "ref_ptracer.npz"), **out)
    print("wrote", len(out), "arrays; threads", int(out["n_threads"]))


if __name__ == "__main__":
    main()
