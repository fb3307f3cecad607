"""Generates tests/golden/ref_mesh_lens.npz with the REFERENCE's own scalar_rgb build (tools/build_reference.sh pinned): rays
through a MESHED singlet — the notebooks' lens representation (icosphere caps whose vertex z follows the sag function, smooth
vertex normals computed by the reference's Mesh, `dielectric` BSDF) — interaction by interaction with the reference's own
scene.ray_intersect and bsdf.sample, and central finite differences of the landing points with respect to five lens variables,
each perturbed scene loaded and traced by the reference again. Run in the build container only:

    R=baseline/_ref/pinned/python; LD_LIBRARY_PATH=$R/mitsuba:$R/drjit PYTHONPATH=$R python tests/golden/gen_ref_mesh_lens.py

Nothing of the reference is stored except its outputs. The cap meshes come from the product's host-side generator
(mesh_lens.surface_meshes: numpy only) and are stored with the outputs."""
import os
import sys
import tempfile

import mitsuba as mi   # must be imported before numpy

mi.set_variant("scalar_rgb")
import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
f32 = np.float32
RAY_EPSILON = f32(8.94069671630859375e-05)

# variables: (row, column) of the interface table (rear -> front rows16) and the finite-difference step
VARIABLES = [("c_front", 1, 3, 2.5e-3), ("c_rear", 0, 3, 2.0e-3), ("z_front", 1, 1, 2.0e-4), ("K_front", 1, 4, 0.2), ("a4_front", 1, 6, 5.0e-4)]


def write_ply(path, verts, faces):
    head = ["ply", "format binary_little_endian 1.0", f"element vertex {len(verts)}", "property float x", "property float y",
            "property float z", f"element face {len(faces)}", "property list uchar int vertex_indices", "end_header"]
    with open(path, "wb") as fh:
        fh.write(("\n".join(head) + "\n").encode())
        fh.write(np.ascontiguousarray(verts, "<f4").tobytes())
        rec = np.zeros(len(faces), dtype=[("n", "u1"), ("i", "<i4", 3)])
        rec["n"] = 3; rec["i"] = faces
        fh.write(rec.tobytes())


def sag64(row, x, y):
    """even asphere in the unitless radius (realisticlens.cpp:680-704 == lens.py:210-237), float64"""
    Re = row[2]
    q = (x * x + y * y) / (Re * Re)
    z = np.zeros_like(q)
    for i in range(6, -1, -1):
        z = z * q + row[6 + i]
    z = z * q * q
    z = z + row[3] * q / (1 + np.sqrt(np.clip(1 - (1 + row[4]) * row[3] * row[3] * q, 0, None)))
    return row[1] - Re * z


def build_scene(rows, xy, faces, tmp):
    """the reference's scene of the two caps: vertex z from the sag (float64, rounded once), no normals in the file — the
    reference's PLY loader computes them (Mesh::recompute_vertex_normals)"""
    d = {"type": "scene"}
    for i in range(len(rows)):
        V = np.concatenate([xy[i], sag64(rows[i].astype(np.float64), xy[i][:, 0], xy[i][:, 1])[:, None]], 1).astype(f32)
        fn = os.path.join(tmp, f"cap{i}.ply")
        write_ply(fn, V, faces[i])
        d[f"lens{i}"] = {"type": "ply", "filename": fn, "face_normals": False,
                         "bsdf": {"type": "dielectric", "int_ior": float(rows[i, 13]), "ext_ior": float(rows[i, 14])}}
    return mi.load_dict(d)


def chain(scene, od, max_interactions=4):
    """every ray through the scene with the reference's ray_intersect + the dielectric's transmission lobe. Returns the
    per-interaction records [n, K, 18] = in o3, in d3, t, p3, sh_n3, out d3, fresnel F, shape index; landing points; alive."""
    ctx = mi.BSDFContext()
    ids = [s.id() for s in scene.shapes()]
    n = len(od)
    rec = np.full((n, max_interactions, 18), np.nan, np.float64)
    land = np.zeros((n, 2), np.float64); alive = np.ones(n, bool)
    for i in range(n):
        o = od[i, :3].astype(f32); d = od[i, 3:].astype(f32)
        d = (d / np.linalg.norm(d.astype(np.float64))).astype(f32)
        for k in range(max_interactions):
            ray = mi.Ray3f([float(x) for x in o], [float(x) for x in d])
            si = scene.ray_intersect(ray)
            if not si.is_valid():
                break
            bsdf = si.bsdf()
            bs, w = bsdf.sample(ctx, si, 0.9999999, [0.5, 0.5])
            if mi.has_flag(bs.sampled_type, mi.BSDFFlags.DeltaReflection):
                alive[i] = False
                break
            d2 = np.array(si.to_world(bs.wo), f32)
            eta = float(bs.eta) if si.wi.z > 0 else 1.0 / float(bs.eta)
            F = float(mi.fresnel(float(si.wi.z), float(bs.eta if si.wi.z > 0 else 1.0 / bs.eta))[0])
            p = np.array(si.p, f32)
            rec[i, k] = [*o, *d, float(si.t), *p, *np.array(si.sh_frame.n, f32), *d2, F, ids.index(si.shape.id())]
            c = RAY_EPSILON * f32(0.01) * (f32(1) + np.abs(p).max())
            o = (p + d2 * c).astype(f32); d = d2
            del eta
        t = -float(o[2]) / float(d[2])
        land[i] = [float(o[0]) + t * float(d[0]), float(o[1]) + t * float(d[1])]
    return rec, land, alive


def main():
    mi.Thread.thread().logger().set_log_level(mi.LogLevel.Warn)
    # the table and the cap meshes come from the product's host-side generator, run in a process of its own (torch and the
    # reference build do not share a process)
    import subprocess
    stage = os.path.join(tempfile.gettempdir(), "mesh_lens_stage.npz")
    code = ("import sys, numpy as np; sys.path.insert(0, %r); import lensrt as L\n"
            "rows = L.lens_design.asphere_table([30.0, -45.0], [4.0, 52.0], [8, 8], [1.5168, 1.000277]).astype(np.float32)\n"
            "rows[1, 4] = -0.6; rows[1, 6] = 0.002; rows[0, 6] = -0.001\n"
            "caps, _ = L.mesh_lens.surface_meshes(rows, subdiv=3)\n"
            "np.savez(%r, rows=rows, **{f'xy{i}': c[0][:, :2] for i, c in enumerate(caps)}, **{f'faces{i}': c[1].astype(np.int32) for i, c in enumerate(caps)})\n"
            % (ROOT, stage))
    env = {k: v for k, v in os.environ.items() if k not in ("PYTHONPATH", "LD_LIBRARY_PATH")}
    subprocess.run([sys.executable, "-c", code], check=True, env=env)
    with np.load(stage) as st:
        rows = st["rows"]
        xy = [st[f"xy{i}"] for i in range(len(rows))]; faces = [st[f"faces{i}"] for i in range(len(rows))]
    os.remove(stage)
    # rays: from a point source 3 degrees off axis towards a disk of points on the front cap, started 2 cm before it
    rng = np.random.default_rng(5)
    n = 96
    r = 0.85 * rows[-1, 2] * np.sqrt(rng.random(n)); phi = 2 * np.pi * rng.random(n)
    ap = np.stack([r * np.cos(phi), r * np.sin(phi), np.full(n, rows[-1, 1])], 1)
    src = np.array([1.0 * np.tan(np.radians(3.0)), 0.3 * np.tan(np.radians(3.0)), rows[-1, 1] + 1.0])
    d = ap - src; d /= np.linalg.norm(d, axis=1, keepdims=True)
    od = np.concatenate([ap - 0.02 * d, d], 1).astype(f32)
    out = {"rows": rows, "rays": od, "variables": np.array([v[0] for v in VARIABLES]), "var_index": np.array([[v[1], v[2]] for v in VARIABLES]),
           "var_step": np.array([v[3] for v in VARIABLES])}
    for i in range(len(rows)):
        out[f"xy{i}"] = xy[i]; out[f"faces{i}"] = faces[i]
    with tempfile.TemporaryDirectory() as tmp:
        scene = build_scene(rows, xy, faces, tmp)
        rec, land, alive = chain(scene, od)
        out["records"] = rec; out["landing"] = land; out["alive"] = alive
        print("alive", alive.sum(), "of", n, "interactions per ray", np.isfinite(rec[:, :, 0]).sum(1).mean(), "landing rms (mm)",
              1e3 * np.sqrt(((land - land.mean(0)) ** 2).sum(1).mean()))
        # vertex normals as the reference computed them (first cap), for the oracle's own recompute_vertex_normals
        for i, s in enumerate(scene.shapes()):
            params = mi.traverse(s)
            key = [k for k in params.keys() if k.endswith("vertex_normals")][0]
            out[f"normals_{s.id()}"] = np.array(params[key], f32).reshape(-1, 3)
        fd = np.zeros((len(VARIABLES), n, 2))
        for j, (name, ri, ci, h) in enumerate(VARIABLES):
            lands = []
            for sgn in (+1, -1):
                rr = rows.astype(np.float64).copy(); rr[ri, ci] += sgn * h
                sc = build_scene(rr, xy, faces, tmp)
                _, l, a = chain(sc, od)
                assert (a == alive).all()
                lands.append(l)
            fd[j] = (lands[0] - lands[1]) / (2 * h)
            print(name, "d landing / d theta: rms", np.sqrt((fd[j] ** 2).sum(1).mean()), "shift at +h (mm)", 1e3 * np.abs(lands[0] - land).max())
        out["fd_landing"] = fd
    np.savez_compressed(os.path.join(HERE, "ref_mesh_lens.npz"), **out)
    print("wrote ref_mesh_lens.npz")


if __name__ == "__main__":
    main()
