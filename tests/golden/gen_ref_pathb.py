"""Generates tests/golden/ref_pathb.npz with the REFERENCE's own scalar_rgb build (tools/build_reference.sh pinned): the
notebooks' "Path B" in miniature (tests/ref_scenes.py path_b_scene: mesh lens + `directionalarea` field source + diffuse film
plane + perspective sensor + `ptracer`). Run in the build container only:

    R=baseline/_ref/pinned/python; LD_LIBRARY_PATH=$R/mitsuba:$R/drjit PYTHONPATH=$R python tests/golden/gen_ref_pathb.py

The scalar AdjointIntegrator's images depend on the machine's thread count (see gen_ref_ptracer.py): it is stored too."""
import os
import sys

import mitsuba as mi   # must be imported before numpy

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ref_scenes as rs  # noqa: E402


def main():
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Error)
    T = lambda m: mi.ScalarTransform4f(np.asarray(m, np.float32))   # noqa: E731
    out = {"n_threads": np.array(os.cpu_count(), np.int32)}
    for seed in (0, 3):
        scene = mi.load_dict(rs.path_b_scene(T))
        img = np.array(mi.render(scene, seed=seed), dtype=np.float32)
        out[f"image_seed{seed}"] = img
        print("seed", seed, img.shape, float(img.mean()), float(img.max()), "lit pixels", int((img.sum(-1) > 0).sum()))
    np.savez_compressed(os.path.join(HERE, "ref_pathb.npz"), **out)


if __name__ == "__main__":
    main()
