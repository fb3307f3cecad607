"""The fork's own scene files (notebooks/scenes/cboxes/*.xml), read IN PLACE from the reference checkout: the reference's
scalar_rgb build renders each one (subprocess under baseline/_ref/pinned), this repo's XML loader + the pinned oracle render the
same file in the scalar sampling convention, and the two images are compared. Build container only (needs /root/reference and
tools/build_reference.sh's output); writes profiles/<tag>_fork_scenes.json.

    python tools/check_fork_scenes.py r02n
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
SCENES = "/root/reference/notebooks/scenes/cboxes"
# (cbox_path.xml is stale in the fork itself: the reference refuses it — "unreferenced properties [fov, fov_axis] in sensor plugin
# of type realisticlens" — and so does this repo's loader; cbox_ptracer.xml is stale the same way: see the end of main())
CASES = [("cbox_real", {"integrator": "path", "spp": 4, "res": 64}), ("cbox_ref", {"integrator": "path", "spp": 4, "res": 64}),
         ("focus_near", {"spp": 2, "res": 96}), ("grid", {"spp": 2, "resx": 96, "resy": 96}), ("grid_SA_coma", {"spp": 2, "resx": 96, "resy": 96}),
         ("grid_astig", {"spp": 2, "resx": 96, "resy": 96}), ("grid_chromatic", {"spp": 2, "resx": 96, "resy": 96}),
         ("grid_distortion", {"spp": 2, "resx": 96, "resy": 96}), ("grid_exp-doublet", {"spp": 2, "resx": 96, "resy": 96})]

REF_SCRIPT = r'''
import sys, json
import mitsuba as mi
mi.set_variant("scalar_rgb")
import numpy as np
mi.Thread.thread().logger().set_log_level(mi.LogLevel.Error)
cases = json.loads(sys.argv[1])
out = {}
for name, kw in cases:
    scene = mi.load_file("%s/" + name + ".xml", **{k: str(v) for k, v in kw.items()})
    out[name] = np.array(mi.render(scene, seed=0), dtype=np.float32)
np.savez(sys.argv[2], **out)
''' % SCENES


def main(tag):
    mi = __import__("lensrt")
    from test_ref_parity import oracle_scalar_render, rel_rmse
    R = os.path.join(ROOT, "baseline", "_ref", "pinned", "python")
    with tempfile.TemporaryDirectory() as td:
        npz = os.path.join(td, "ref.npz")
        env = dict(os.environ, LD_LIBRARY_PATH=f"{R}/mitsuba:{R}/drjit", PYTHONPATH=R)
        subprocess.run([sys.executable, "-c", REF_SCRIPT, json.dumps(CASES), npz], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        ref = dict(np.load(npz))
    report = {"scenes_dir": SCENES, "note": "reference scalar_rgb render vs this repo's XML loader + pinned oracle (scalar sampling convention), same seed",
              "cases": []}
    for name, kw in CASES:
        d = mi.xml_to_dict(path=os.path.join(SCENES, name + ".xml"), **kw)
        fd = mi.scene.scene_data_from_dict(d)[1][0]["film"]
        # the reference picks its render block size from the film size and ITS thread count (integrator.cpp:171-183)
        bs = mi.scene.reference_block_size(fd["width"], fd["height"], os.cpu_count())
        img, miss = oracle_scalar_render(mi, d, int(kw["spp"]), analytic=True, block_size=bs)
        r = ref[name]
        err = rel_rmse(img, r)
        close = float((np.abs(img - r).max(-1) <= 1e-4 * np.maximum(r.max(-1), 1e-3)).mean())
        report["cases"].append({"scene": name + ".xml", "parameters": kw, "shape": list(r.shape), "block_size": bs, "rel_rmse": err, "pixels_within_1e-4": close,
                                "mean_reference": float(r.mean()), "mean_ours": float(img.mean())})
        print(f"{name:18s} rel-RMSE {err:.3e}  pixels within 1e-4: {close:.4f}  mean {r.mean():.5f} / {img.mean():.5f}", flush=True)
    # cbox_ptracer.xml: stale in the fork in the same way as cbox_path.xml (both loaders refuse `fov` / `fov_axis` on a realistic
    # lens with the same message). With those two lines removed — a scratch copy next to a link to the scene's meshes — it is
    # the fork's own particle-tracer scene (tessar lens, max_depth -1): reference ptracer vs the pinned oracle's ptracer.
    from test_ptracer_parity import oracle_ptracer
    with tempfile.TemporaryDirectory() as td:
        txt = open(os.path.join(SCENES, "cbox_ptracer.xml")).read()
        kept = [l for l in txt.splitlines() if 'name="fov_axis"' not in l and 'name="fov"' not in l]
        assert len(kept) == len(txt.splitlines()) - 2
        fn = os.path.join(td, "cbox_ptracer_nofov.xml")
        open(fn, "w").write("\n".join(kept))
        os.symlink(os.path.join(SCENES, "meshes"), os.path.join(td, "meshes"))
        kw = {"spp": 8, "res": 96}
        script = ("import sys, mitsuba as mi\nmi.set_variant('scalar_rgb')\nimport numpy as np\n"
                  "mi.Thread.thread().logger().set_log_level(mi.LogLevel.Error)\n"
                  "sc = mi.load_file(sys.argv[1], spp='8', res='96')\n"
                  "np.save(sys.argv[2], np.array(mi.render(sc, seed=0), dtype=np.float32))\n")
        env = dict(os.environ, LD_LIBRARY_PATH=f"{R}/mitsuba:{R}/drjit", PYTHONPATH=R)
        subprocess.run([sys.executable, "-c", script, fn, os.path.join(td, "ref.npy")], check=True, env=env, capture_output=True, text=True)
        n_threads = os.cpu_count()          # the scalar variants' thread pool (nanothread default)
        r = np.load(os.path.join(td, "ref.npy"))
        d = mi.xml_to_dict(path=fn, **kw)
        img = oracle_ptracer(mi, d, 0, n_threads)
        err = rel_rmse(img, r)
        close = float((np.abs(img - r).max(-1) <= 1e-4 * np.maximum(r.max(-1), 1e-3)).mean())
        report["cases"].append({"scene": "cbox_ptracer.xml with its two stale sensor properties (fov, fov_axis) removed", "parameters": kw,
                                "integrator": "ptracer (reference) vs oracle ptracer, scalar convention, %d threads" % n_threads, "shape": list(r.shape),
                                "rel_rmse": err, "pixels_within_1e-4": close, "mean_reference": float(r.mean()), "mean_ours": float(img.mean())})
        print(f"{'cbox_ptracer*':18s} rel-RMSE {err:.3e}  pixels within 1e-4: {close:.4f}  mean {r.mean():.5f} / {img.mean():.5f}", flush=True)
    path = os.path.join(ROOT, "profiles", f"{tag}_fork_scenes.json")
    json.dump(report, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "dev")
