"""Times the REFERENCE's own scalar_rgb build (baseline/_ref, built by tools/build_reference.sh) on bench.py's workload.
Run by bench.py in a subprocess with PYTHONPATH=baseline/_ref/<flavour>/python; imports nothing of the product package
(the scene generator mitsuba3-camera_b200/workloads.py and transform.py are loaded by path: pure numpy).

The height-field meshes are handed to the reference as binary PLY files (face_normals = true: flat shading, as the
product's in-memory `mesh` shapes) written to a scratch directory. Prints one JSON line."""
import argparse
import importlib.util
import json
import os
import sys
import tempfile
import time

import mitsuba as mi   # before numpy (import-order gotcha of this build)

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mitsuba3-camera_b200")


def load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec); sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def write_ply(path, verts, faces):
    head = ["ply", "format binary_little_endian 1.0", f"element vertex {len(verts)}", "property float x", "property float y",
            "property float z", f"element face {len(faces)}", "property list uchar int vertex_indices", "end_header"]
    with open(path, "wb") as fh:
        fh.write(("\n".join(head) + "\n").encode())
        fh.write(np.ascontiguousarray(verts, "<f4").tobytes())
        rec = np.zeros(len(faces), dtype=[("n", "u1"), ("i", "<i4", 3)])
        rec["n"] = 3; rec["i"] = faces
        fh.write(rec.tobytes())


def to_reference(d, tmp):
    """Our scene dictionary -> one the reference loads: Transform4f -> ScalarTransform4f, in-memory meshes -> PLY files."""
    out = {}
    for k, v in d.items():
        if isinstance(v, dict):
            if v.get("type") == "mesh":
                fn = os.path.join(tmp, f"{k}.ply")
                write_ply(fn, np.asarray(v["vertices"], np.float32), np.asarray(v["faces"], np.int32))
                v = {kk: vv for kk, vv in v.items() if kk not in ("vertices", "faces")}
                v.update(type="ply", filename=fn, face_normals=True)
            out[k] = to_reference(v, tmp)
        elif hasattr(v, "matrix"):
            out[k] = mi.ScalarTransform4f(np.asarray(v.matrix, np.float32))
        elif isinstance(v, np.ndarray) and k == "data":      # in-memory bitmap texture
            out[k] = mi.TensorXf(np.ascontiguousarray(v, np.float32))
        else:
            out[k] = v
    return out


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--workload", default="c5")
    p.add_argument("--width", type=int, default=3840); p.add_argument("--height", type=int, default=2160)
    p.add_argument("--n-side", type=int, default=708); p.add_argument("--lens", default="gauss")
    p.add_argument("--n-boxes", type=int, default=64); p.add_argument("--tex-size", type=int, default=2048)
    p.add_argument("--max-depth", type=int, default=6); p.add_argument("--spp", type=int, default=2)
    p.add_argument("--crop", type=int, nargs=4, default=None, help="crop_width crop_height offset_x offset_y")
    p.add_argument("--steps", type=int, default=1); p.add_argument("--warmup", type=int, default=0)
    p.add_argument("--threads", type=int, default=0)
    p.add_argument("--integrator", default="path", help="path | ptracer")
    a = p.parse_args()
    threads = a.threads or os.cpu_count()
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Warn)
    try:
        mi.Thread.set_thread_count(threads)      # scalar variants: nanothread pool size
    except Exception:
        pass
    load_by_path("transform", os.path.join(PKG, "transform.py"))
    wl = load_by_path("lrt_workloads", os.path.join(PKG, "workloads.py"))
    with tempfile.TemporaryDirectory() as tmp:
        if a.workload == "c5":
            d = wl.heightfield_scene(n_side=a.n_side, width=a.width, height=a.height, spp=a.spp, lens_design=a.lens, max_depth=a.max_depth)
        elif a.workload == "c4":
            d = wl.textured_boxes_scene(n_boxes=a.n_boxes, width=a.width, height=a.height, spp=a.spp, lens_design=a.lens, max_depth=a.max_depth,
                                        tex_size=a.tex_size)
        elif a.workload == "c1":
            d = mi.cornell_box(); d["integrator"]["max_depth"] = a.max_depth
            d["sensor"]["film"]["width"] = a.width; d["sensor"]["film"]["height"] = a.height
        elif a.workload == "pathb":      # the notebooks' pipeline in miniature (tests/ref_scenes.py path_b_scene): ptracer, perspective sensor
            rs = load_by_path("ref_scenes", os.path.join(ROOT, "tests", "ref_scenes.py"))
            d = rs.path_b_scene(lambda m: mi.ScalarTransform4f(np.asarray(m, np.float32)), spp=a.spp, res=a.width)
            a.height = a.width
        else:
            raise SystemExit("unknown workload " + a.workload)
        if a.integrator != "path" and a.workload != "pathb":
            d["integrator"]["type"] = a.integrator
        if a.crop and a.workload != "pathb":
            d["sensor"]["film"].update(crop_width=a.crop[0], crop_height=a.crop[1], crop_offset_x=a.crop[2], crop_offset_y=a.crop[3])
        t0 = time.perf_counter()
        scene = mi.load_dict(to_reference(d, tmp) if a.workload in ("c5", "c4") else d)
        load_s = time.perf_counter() - t0
    cw, ch = (a.crop[0], a.crop[1]) if a.crop else (a.width, a.height)
    times = []
    for i in range(a.warmup + a.steps):
        t0 = time.perf_counter()
        img = mi.render(scene, spp=a.spp, seed=i)
        if i >= a.warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    print(json.dumps({"seconds": sec, "paths": cw * ch * a.spp, "threads": threads, "scene_load_s": load_s,
                      "mean": float(np.array(img).mean()), "variant": "scalar_rgb"}))


if __name__ == "__main__":
    main()
