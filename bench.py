#!/usr/bin/env python3
"""bench.py — Mpaths/s of the render hot path (realistic-lens sample_ray -> path integrator -> film splat).

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run, one rank per GPU)
  python bench.py --impl reference ...                      (the CPU arm: oracle port on all host cores)

Workload (config.workload): BASELINE.json configs[4] — the 1M-triangle diffuse/conductor scene through the
`gauss` realistic lens at 3840x2160 — the configuration the metric "Mpaths/s at 1/2/4/8 B200" is quoted on.
One step = one pass of --spp-per-step samples per pixel over the whole film (the 4096-spp job is 4096/spp such
steps; throughput in paths/s does not depend on the split). Pixel tiles are dealt round-robin over ranks, each
rank accumulates into its own film storage and one NCCL reduce to rank 0 merges them (weak scaling is NOT used: the
film is fixed, so `scaling` is "strong")."""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--spp-per-step", type=int, default=64,
                   help="samples per pixel of one step (one pass; the reference's own passes are bounded by W*H*spp < 2^32, i.e. <= 517 spp here)")
    p.add_argument("--cpu-spp", type=int, default=16, help="spp of the parity crop rendered by the oracle port (same film crop, same scene)")
    p.add_argument("--ref-spp", type=int, default=1, help="spp of the bounded sample the reference's own scalar_rgb build renders (FULL film: "
                                                            "the same vignetting mix as the GPU's step)")
    p.add_argument("--ref-flavour", default="native", choices=["native", "pinned"], help="which build under baseline/_ref is timed")
    p.add_argument("--workload", default="c5", choices=["c5", "c4"],
                   help="c5 (default, the configuration the metric is quoted on): 1M-triangle height field, `gauss` lens, 3840x2160; "
                        "c4: 64 textured boxes through the Nikon `exp2a` stack with exit-pupil sampling, 1920x1080 (1024 spp = 16 steps)")
    p.add_argument("--width", type=int, default=None)
    p.add_argument("--height", type=int, default=None)
    p.add_argument("--n-side", type=int, default=708)
    p.add_argument("--n-boxes", type=int, default=64)
    p.add_argument("--tex-size", type=int, default=2048)
    p.add_argument("--lens", default=None)
    p.add_argument("--max-depth", type=int, default=6)
    p.add_argument("--cpu-crop", type=int, nargs=2, default=[1280, 720], help="crop window of the CPU baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-lens-grad", action="store_true", help="skip the lens-gradient iterations/s section (C2 / C3 shapes)")
    p.add_argument("--lens-iters", type=int, default=10)
    a = p.parse_args()
    dflt = {"c5": (3840, 2160, "gauss"), "c4": (1920, 1080, "exp2a")}[a.workload]
    a.width, a.height, a.lens = a.width or dflt[0], a.height or dflt[1], a.lens or dflt[2]
    return a


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.reasons, self.stop_flag, self.max_mhz = gpu_index, [], set(), False, None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_throttle_reasons.hw_slowdown,clocks_throttle_reasons.hw_thermal_slowdown," \
            "clocks_throttle_reasons.sw_thermal_slowdown,clocks_throttle_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0])); self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if "Active" in v and "Not" not in v:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def workload_dict(mi, a):
    if a.workload == "c4":
        return mi.workloads.textured_boxes_scene(n_boxes=a.n_boxes, width=a.width, height=a.height, spp=a.spp_per_step, lens_design=a.lens,
                                                 max_depth=a.max_depth, tex_size=a.tex_size)
    return mi.workloads.heightfield_scene(n_side=a.n_side, width=a.width, height=a.height, spp=a.spp_per_step, lens_design=a.lens,
                                max_depth=a.max_depth)


def parity_crop(mi, torch, a, d, port):
    """GPU render of the crop the oracle port rendered (same seed / spp / sample streams): rel-RMSE of the developed images
    and the lens ray-miss counts, which must be equal."""
    import copy
    import numpy as np
    cw, ch = min(a.cpu_crop[0], a.width), min(a.cpu_crop[1], a.height)
    d2 = dict(d); d2["sensor"] = dict(d["sensor"]); d2["sensor"]["film"] = dict(d["sensor"]["film"])
    d2["sensor"]["film"].update(crop_width=cw, crop_height=ch, crop_offset_x=(a.width - cw) // 2, crop_offset_y=(a.height - ch) // 2)
    sensor = mi.load_dict(d2["sensor"])
    scene = parity_crop.scene
    spp = min(a.cpu_spp, a.spp_per_step)
    img = mi.render(scene, sensor=sensor, seed=0, spp=spp).cpu().numpy().astype(np.float64)
    miss_gpu = int(scene.stats()["lens_misses"])
    want = port["image"].astype(np.float64)
    rel = float(np.sqrt(np.mean((img - want) ** 2)) / max(np.sqrt(np.mean(want ** 2)), 1e-12))
    assert miss_gpu == port["lens_miss"], f"lens ray-miss count differs: GPU {miss_gpu}, oracle {port['lens_miss']}"
    return {"what": f"centred {cw}x{ch} crop, {spp} spp, seed 0: GPU vs oracle port (pinned to the reference, tests/test_ref_parity.py)",
            "rel_rmse": rel, "lens_miss_gpu": miss_gpu, "lens_miss_cpu": int(port["lens_miss"]), "image_mean_gpu": float(img.mean()),
            "image_mean_cpu": float(want.mean())}


def cpu_sample(a, scene_arrays, sensor_dict, steps=1, warmup=0, keep_image=False):
    """The CPU arm: the oracle port (OpenMP, all host cores) on a centred crop of the same film at --cpu-spp samples per pixel
    (a bounded sample: paths/s of the port does not depend on the pass size)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as ol
    ld = ol.load_lens_designs()
    cw, ch = a.cpu_crop
    cw, ch = min(cw, a.width), min(ch, a.height)
    film = ol.FilmSpec(a.width, a.height, [cw, ch, (a.width - cw) // 2, (a.height - ch) // 2], 35.0)
    h = ol.orc_scene(scene_arrays)
    tw = sensor_dict["to_world"].matrix
    cam = ol.orc_camera(ld.design(a.lens, {"object_distance": sensor_dict["object_distance"]}), film, to_world=tw,
                        near=sensor_dict["near_clip"], far=sensor_dict["far_clip"], object_distance=sensor_dict["object_distance"])
    if sensor_dict.get("sample_exit_pupil"):     # the 64 x 2^20-ray LUT on the host: ~20 s, not timed
        ol.cam_exit_pupil(ol.oracle(), "orc", cam, shutter_open_time=sensor_dict.get("shutter_close", 0.0) - sensor_dict.get("shutter_open", 0.0))
    cpu_spp = min(a.cpu_spp, a.spp_per_step)
    paths = cw * ch * cpu_spp
    times = []
    st = miss = None
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        st_i, miss_i = ol.orc_render(h, film, cpu_spp, seed=i, max_depth=a.max_depth, cam_h=cam)
        if i == 0:
            st, miss = st_i, miss_i
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    out = {"value": paths / sec / 1e6, "unit": "Mpaths/s", "cores": os.cpu_count(), "kind": "port",
           "sample": f"oracle port (OpenMP), centred {cw}x{ch} crop of the {a.width}x{a.height} film, {cpu_spp} spp ({paths} paths, {sec:.2f} s/step)"}
    if keep_image:
        out["image"], out["lens_miss"] = ol.develop(st), int(miss)
    return out, sec


def reference_sample(a, steps=1, warmup=0, extra=()):
    """The CPU arm proper: the REFERENCE's own scalar_rgb build (tools/build_reference.sh -> baseline/_ref, shipped to the GPU box)
    renders the same scene over the FULL film at --ref-spp samples per pixel with every host core, in a subprocess that never
    imports the product (tools/ref_bench.py). None if the build is not there."""
    py = os.path.join(ROOT, "baseline", "_ref", a.ref_flavour, "python")
    if not os.path.isdir(os.path.join(py, "mitsuba")):
        return None
    env = dict(os.environ); env["PYTHONPATH"] = py; env.pop("OMP_NUM_THREADS", None)
    # the build tree the extension modules' RUNPATH names does not exist on the GPU box: the libraries travel next to them
    env["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(py, "mitsuba"), os.path.join(py, "drjit"), env.get("LD_LIBRARY_PATH", "")])
    cmd = [sys.executable, os.path.join(ROOT, "tools", "ref_bench.py"), "--workload", a.workload, "--n-boxes", str(a.n_boxes),
           "--tex-size", str(a.tex_size), "--width", str(a.width), "--height", str(a.height),
           "--n-side", str(a.n_side), "--lens", a.lens, "--max-depth", str(a.max_depth), "--spp", str(a.ref_spp),
           "--steps", str(steps), "--warmup", str(warmup)] + list(extra)
    for k in range(0, len(extra), 2):        # later options override earlier ones in argparse; drop the duplicates for readability
        if extra[k] in cmd[:-len(extra)]:
            i = cmd.index(extra[k]); del cmd[i:i + 2]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=1500, env=env, cwd="/tmp")
        # (the particle tracer's progress bar shares the last line with the result: take the JSON object from its first brace)
        r = json.loads([ln[ln.index('{"seconds"'):] for ln in out.stdout.splitlines() if '{"seconds"' in ln][-1])
    except Exception as e:
        sys.stderr.write(f"reference build failed to run: {e}\n{out.stderr[-2000:] if 'out' in dir() else ''}\n")
        return None
    return {"value": r["paths"] / r["seconds"] / 1e6, "unit": "Mpaths/s", "cores": r["threads"], "kind": "reference",
            "sample": f"the reference's own scalar_rgb build ({a.ref_flavour}: "
                      + ("-march=x86-64-v3, -ffp-contract=fast" if a.ref_flavour == "native" else "plain x86-64, -ffp-contract=off")
                      + f"), full {a.width}x{a.height} film at {a.ref_spp} spp ({r['paths']} paths, {r['seconds']:.2f} s/step, native kd-tree, "
                        f"scene load {r['scene_load_s']:.1f} s not timed)",
            "image_mean": r["mean"]}, r["seconds"]


def doublet_table(mi):
    """C2: the notebooks' cemented doublet (doublet_full_optim.ipynb cell 6) as three even-asphere interfaces."""
    L = mi.lens_design
    n1, n2 = L.cauchy_ior(1.45608, 0.021087), L.cauchy_ior(1.56895, 0.026305)
    return L.asphere_table([24.0, -24.0, -168.0107], [3.0, 2.2529, 46.7471], [8, 8, 8], [n1, n2, 1.000277])


def triplet_table(mi):
    """C3: a Cooke triplet (six even-asphere surfaces; paraxial EFL 50.0 mm, BFL 46.3 mm, semi-diameters 7 / 5 / 7 mm),
    synthetic prescription."""
    L = mi.lens_design
    crown, flint = L.cauchy_ior(1.6030, 0.00600), L.cauchy_ior(1.6000, 0.01500)
    return L.asphere_table([17.223, -128.586, -16.413, 15.643, 114.483, -14.009], [3.242, 2.431, 1.621, 2.431, 3.242, 46.2], [7, 7, 5, 5, 7, 7],
                           [crown, 1.000277, flint, 1.000277, crown, 1.000277])


def widened_rows_section(mi, torch, a, scene):
    """Device-timed figures for two rows built after the headline path: the mesh-lens trace with its adjoint and
    the dispersion mode of the lens step (8f-4). CUDA events, 3 warm-up + 5 timed repetitions each."""
    def timed(fn, reps=5, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    out = {}
    # mesh lens: the notebooks' doublet, subdivision 6 (2 x 20 000 triangles + baffles), 2^20 rays from a field point
    rows = doublet_table(mi)
    lens = mi.mesh_lens.MeshLens(mi, rows, subdiv=6)
    src = mi.lens_design.field_source(2.0, 1.0e3, front_z=float(rows[-1, 1]))
    off = 2.0 * float(rows[-1, 2])
    n = 1 << 20
    ap, d, o = lens._spot_rays(src, n, 1, off)
    rt = torch.tensor(rows, dtype=torch.float64, device="cuda", requires_grad=True)

    def fwd():
        lens.trace(o, d)

    def fwd_bwd():
        rt.grad = None
        po, pd, alive, w = lens.trace_ad(rt, o, d)
        t = -po[:, 2] / pd[:, 2]
        pf = (po + t[:, None] * pd)[:, :2][alive]
        ((pf - pf.mean(0)) ** 2).sum(1).mean().backward()
    ms_f, ms_fb = timed(fwd), timed(fwd_bwd)
    out["mesh_lens"] = {"rays": n, "triangles": int(lens.scene.info()["triangles"]), "trace_ms": ms_f, "Mrays_per_s": n / ms_f / 1e3,
                        "design_iteration_ms": ms_fb, "iters_per_s": 1e3 / ms_fb,
                        "what": "meshed doublet, subdivision 6: lrt_scene_trace_refractive alone; and one design iteration = sag -> vertices -> "
                                "vertex normals (torch) -> BVH refit -> trace -> rms-spot loss -> lrt_scene_trace_refractive_backward -> torch "
                                "chain to the 3 x 16 table entries"}
    # dispersion: sample_ray with one wavelength per ray through the headline lens
    sens = mi.RealisticLensCamera(scene.sensors()[0].film, a.lens, spectral=True)
    g = torch.Generator(device="cuda").manual_seed(0)
    m = 1 << 24
    pos = torch.rand((m, 2), device="cuda", generator=g); apx = torch.rand((m, 2), device="cuda", generator=g)
    wl = 380.0 + 320.0 * torch.rand(m, device="cuda", generator=g)
    ms_wl = timed(lambda: sens.sample_ray_wavelength(pos, apx, wl))
    ms_rgb = timed(lambda: sens.sample_ray(0.0, None, pos, apx))
    out["dispersion"] = {"rays": m, "sample_ray_wl_ms": ms_wl, "Grays_per_s": m / ms_wl / 1e6, "sample_ray_rgb_ms": ms_rgb,
                         "Grays_per_s_rgb": m / ms_rgb / 1e6,
                         "what": "lrt_sensor_sample_ray_wl on the `%s` lens (per-ray Cauchy / Sellmeier IOR of both media of each of its "
                                 "interfaces) beside lrt_sensor_sample_ray (constant IORs)" % a.lens}
    # the notebooks' "Path B" (tests/ref_scenes.py path_b_scene): mesh lens + directionalarea source + film plane + perspective sensor
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_scenes as rs
    res, spp = 256, 64
    sc = mi.load_dict(rs.path_b_scene(lambda m: mi.Transform4f(m), spp=spp, res=res))
    it = sc.integrator()
    seeds = iter(range(1000))
    ms_pb = timed(lambda: it.render(sc, seed=next(seeds)))
    out["path_b"] = {"light_paths": res * res * spp, "ms_per_render": ms_pb, "M_light_paths_per_s": res * res * spp / ms_pb / 1e3,
                     "what": "ptracer on the notebooks' pipeline in miniature: meshed singlet (dielectric caps), directionalarea field source, "
                             "diffuse film plane, perspective sensor; %d x %d film, %d spp" % (res, res, spp)}
    if not a.no_cpu_baseline:
        pr = reference_sample(a, extra=("--workload", "pathb", "--width", str(res), "--spp", str(spp)))
        if pr is not None:
            out["path_b"]["cpu_baseline"] = {"value": pr[0]["value"], "unit": "M light paths/s", "cores": pr[0]["cores"], "kind": "reference",
                                             "sample": "the reference's scalar_rgb ptracer on the same scene dictionary, same film and spp"}
    return out


def lens_grad_section(mi, torch, dist, a, rank, world):
    """Lens-gradient iterations/s: one iteration = per field a primal spot render (spp), the RMS-spot loss, the adjoint
    render (spp_grad, fresh seed), an all-reduce of the table gradient (N > 1: samples are sharded by index) and an Adam
    step. Timed with CUDA events around whole iterations (host work included), max over ranks."""
    out = {}
    shapes = {"C2": dict(table=doublet_table, fields=[0.0, 2.5, 5.0], res=128, spp=256, spp_grad=128, efl=None),
              "C3": dict(table=triplet_table, fields=[0.0, 1.875, 3.75, 5.625, 7.5], res=512, spp=256, spp_grad=256, efl=0.05)}
    for name, cfg in shapes.items():
        rows = cfg["table"](mi)
        try:
            prob = mi.lens_design.LensDesignProblem(mi, rows, cfg["fields"], resolution=cfg["res"], full_res=4096, spp=cfg["spp"],
                                                    spp_grad=cfg["spp_grad"], lr=3e-5, efl_target=cfg["efl"],
                                                    shard=(rank, world))
        except Exception as e:   # a blocked chief ray in a synthetic prescription must not take the bench line down
            out[name] = {"error": str(e)[:200]}
            continue
        def it(record=False):
            # sharded runs: the spot accumulators are all-reduced before the (non-linear) loss and the table gradients after
            # the adjoint (LensDesignProblem.step); variables, table and optimiser state stay on the device
            prob.step(record_loss=record)
        it(record=True)
        for _ in range(3):
            it()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.lens_iters):
            it()
        e1.record(); torch.cuda.synchronize()
        it(record=True)
        ms = torch.tensor([e0.elapsed_time(e1) / a.lens_iters], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_eager = ms.item()
        graph_note = None
        try:    # the whole iteration (table, renders, loss, adjoints, all-reduces, Adam) replayed from one CUDA graph
            prob.capture()
            for _ in range(3):
                prob.step_graph()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0.record()
            for _ in range(a.lens_iters):
                prob.step_graph()
            e1.record(); torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / a.lens_iters], device="cuda", dtype=torch.float64)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            prob.losses.append(float(prob.last_loss))
        except Exception as e:
            graph_note = "CUDA-graph capture failed, eager timing reported: " + str(e)[:160]
        rays = len(cfg["fields"]) * cfg["res"] ** 2 * (cfg["spp"] + cfg["spp_grad"])
        out[name] = {"workload": f"{len(rows)}-surface even-asphere lens, {len(prob.free)} free variables, {len(cfg['fields'])} fields, "
                                 f"{cfg['res']}^2 px, spp {cfg['spp']} primal + {cfg['spp_grad']} adjoint" + (", EFL held at 50 mm" if cfg["efl"] else ""),
                     "iters_per_s": 1e3 / ms.item(), "ms_per_iter": ms.item(), "lens_rays_per_iter": rays,
                     "Grays_per_s": rays / (ms.item() * 1e-3) / 1e9, "loss_first": prob.losses[0], "loss_last": prob.losses[-1],
                     "ms_per_iter_eager": ms_eager, "mode": graph_note or "one CUDA graph per iteration (device-side seeds); eager figure beside it"}
    return out


def lens_grad_cpu(mi, a):
    """CPU arm of the lens-gradient metric: the oracle's float32 spot render + float64 dual-number gradient (the only CPU
    derivative of this path that exists: the reference cannot differentiate its lens table) on a bounded sample of C2:
    one field, 128^2 x 2 spp, 30 seeds; scaled to the full iteration by ray count. Single thread."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as ol
    from test_lens_adjoint import COLS
    rows = doublet_table(mi)
    from mitsuba3_camera_b200.lens_designs import InterfaceList
    film = ol.FilmSpec(4096, 4096, [128, 128, 1984, 1984], 35.0)
    h = ol.orc_camera(InterfaceList(rows, film_z=0.0, fine_focus=False), film)
    O = ol.oracle()
    f2s = np.zeros(16, np.float32); O.orc_cam_film_to_sample(h, ol.fp(f2s))
    s2f = np.zeros(16, np.float32); O.orc_cam_sample_to_film(h, ol.fp(s2f))
    src = mi.lens_design.field_source(0.0, 1.0e3, front_z=float(rows[-1, 1]))
    spp = 2
    t0 = time.perf_counter()
    st, valid = ol.orc_spot_render(h, film, src, [1, 1, 1], 0, spp, origin_offset=0.016)
    G = np.random.default_rng(0).standard_normal((128, 128, 4)).astype(np.float32)
    sel = [(i, c) for i in range(3) for c in [1, 3, 4, 6, 7, 8, 9, 10, 11, 12]]
    ol.orc_spot_grad(rows, f2s, 1.0 / (abs(float(s2f[0])) * abs(float(s2f[5]))), 1e-2, 1e4, film, src, [1, 1, 1], 1, spp, G, sel,
                     origin_offset=0.016)
    sec = time.perf_counter() - t0
    rays_sample = 128 * 128 * spp * 2
    rays_iter = 3 * 128 * 128 * (256 + 128)
    return {"value": 1.0 / (sec * rays_iter / rays_sample), "unit": "iters/s (C2)", "cores": 1, "kind": "port", "meaningful": False,
            "note": "NOT a like-for-like baseline: one thread, forward-mode float64 dual numbers (30 seeds) of the checker; the reference "
                    "cannot differentiate its analytic lens table at all and no llvm_ad_rgb variant can run on this box. Reported for "
                    "completeness only: do not quote the ratio.",
            "sample": f"1 field, 128^2 x {spp} spp primal + adjoint ({rays_sample} lens rays, {sec:.2f} s), scaled by ray count to the "
                      f"{rays_iter}-ray iteration; forward-mode float64 dual numbers (30 seeds)"}


def main():
    a = parse()
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    what = (f"C5: {2 * (a.n_side - 1) ** 2}-triangle diffuse/conductor height field, `{a.lens}` realistic lens" if a.workload == "c5" else
            f"C4: ground + {a.n_boxes} boxes (vertex normals + uvs), {a.tex_size}^2 RGB bitmap texture (bilinear), Cu conductors, "
            f"Nikon `{a.lens}` 22-interface lens with exit-pupil sampling")
    job = 4096 if a.workload == "c5" else 1024
    config = {"workload": f"{what}, {a.width}x{a.height}, path max_depth={a.max_depth}, {a.spp_per_step} spp per step "
                          f"(1/{job // max(a.spp_per_step, 1)} of the {job}-spp job)",
              "l2": "flushed between steps (256 MiB write); BVH+triangles ~60 MB are L2-resident by design",
              "parallelism": f"tile-sharded x{world}, one NCCL all-reduce of the film per step, overlapped with the next step's kernel" if world > 1 else "single GPU"}

    if a.impl == "reference":
        if rank != 0:
            return
        # torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core (libgomp reads this at load time)
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count())
        got = reference_sample(a, steps=a.steps, warmup=a.warmup)      # the reference's own build: nothing of ours is loaded
        if got is None:     # no reference build in the tree: the oracle port (needs the scene front end of the package)
            import lensrt as mi
            d = workload_dict(mi, a)
            arrays = mi.scene_data_from_dict(d)[0].arrays()
            got = cpu_sample(a, arrays, d["sensor"], steps=a.steps, warmup=a.warmup)
        cb, sec = got
        line = {"impl": "reference", "metric": "Mpaths/s", "value": cb["value"], "unit": "Mpaths/s", "n_gpus": a.gpus, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import lensrt as mi
    capi = mi.capi
    d = workload_dict(mi, a)
    t0 = time.perf_counter()
    scene = mi.load_dict(d)
    parity_crop.scene = scene
    build_s = time.perf_counter() - t0
    sensor = scene.sensors()[0]
    H, W = a.height, a.width
    # two film accumulators: the all-reduce of step i (communication stream) overlaps the render kernel of step i+1
    storages = [torch.zeros((H, W, 4), dtype=torch.float32, device="cuda") for _ in range(2 if world > 1 else 1)]
    comm = torch.cuda.Stream() if world > 1 else None
    reduced = [None, None]          # event: the all-reduce that last used storages[k] has finished
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    step_no = [0]
    host_img = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    host_lens = torch.from_numpy(sensor.interfaces().copy()).pin_memory()
    paths_per_step = H * W * a.spp_per_step

    def step(seed, e2e=False):
        if e2e:   # parameters_changed(): lens table from pinned host memory -> device
            sensor.set_interfaces(host_lens.numpy())
        k = step_no[0] % len(storages); step_no[0] += 1
        storage = storages[k]
        main = torch.cuda.current_stream()
        if reduced[k] is not None:
            main.wait_event(reduced[k])         # the collective that read this accumulator two steps ago is done
        storage.zero_()
        mi.render(scene, seed=seed, spp=a.spp_per_step, develop=False, shard=(rank, world), storage=storage)
        if world > 1:                           # sum of the film accumulators over ranks (all-reduce)
            done = torch.cuda.Event(); done.record(main)
            with torch.cuda.stream(comm):
                comm.wait_event(done)
                dist.all_reduce(storage)
                reduced[k] = torch.cuda.Event(); reduced[k].record(comm)
        if e2e and rank == 0:
            if world > 1:
                main.wait_event(reduced[k])     # the end-to-end step hands the merged image to the host
            img = mi.develop_rgb(storage)
            host_img.copy_(img, non_blocking=True)

    def drain():
        for ev in reduced:
            if ev is not None:
                torch.cuda.current_stream().wait_event(ev)

    def timed(n_steps, e2e, seed0):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        # one timed region around the K steps (the collective of step i overlaps the kernel of step i+1; the last one is
        # drained inside the region). The L2 flush between steps is a 256 MiB write on the same stream: ~0.1 ms of the step.
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))]
        ev[0][0].record()
        for i in range(n_steps):
            flush.fill_(i & 0xff)
            step(seed0 + i, e2e)
        drain()
        ev[0][1].record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([sum(s.elapsed_time(e) for s, e in ev)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() / n_steps

    for i in range(a.warmup):
        step(1000 + i)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = capi.lib.lrt_launch_count()
    ms = timed(a.steps, False, 0)
    launches = capi.lib.lrt_launch_count() - l0
    stats = scene.stats()
    if world > 1:    # sum the per-rank counters (every rank traced its own tiles)
        t = torch.tensor([float(stats[k]) for k in sorted(stats)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        stats = {k: int(v) for k, v in zip(sorted(stats), t.tolist())}
    for i in range(min(a.warmup, 2)):
        step(2000 + i, e2e=True)
    ms_e2e = timed(a.steps, True, 100)
    sampler.stop_flag = True
    # adjoint of the path loop (reflectance / radiance gradients): same workload, one pass of lrt_render_backward per step
    gstore = torch.randn((H, W, 4), dtype=torch.float32, device="cuda")
    drain(); torch.cuda.synchronize()

    def bstep(seed):
        gb, ge = mi.render_backward(scene, gstore, seed=seed, spp=a.spp_per_step, shard=(rank, world))
        if world > 1:
            dist.all_reduce(gb); dist.all_reduce(ge)
    bstep(0); torch.cuda.synchronize()
    eb = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    eb[0].record()
    for i in range(max(a.steps // 2, 1)):
        bstep(10 + i)
    eb[1].record(); torch.cuda.synchronize()
    ms_bwd = torch.tensor([eb[0].elapsed_time(eb[1]) / max(a.steps // 2, 1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_bwd, op=dist.ReduceOp.MAX)
    # particle tracer (8f-1) on the SAME scene through the same lens: light paths per second at a 1024 x 576 film, 4 spp
    ptr = None
    if not a.no_lens_grad:
        try:
            sens0 = scene.sensors()[0]
            pfilm = mi.Film(1024, 576, diagonal=sens0.film.diagonal, rfilter="tent")
            psens = mi.RealisticLensCamera(pfilm, sens0.lens_design, to_world=mi.Transform4f(sens0.to_world), near_clip=sens0.near_clip,
                                           far_clip=sens0.far_clip, **sens0.props)
            psens.sampler = mi.IndependentSampler(4, 0)
            pit = mi.PTracerIntegrator(a.max_depth if hasattr(a, "max_depth") else 6, 5, False)
            pit.render(scene, sensor=psens, seed=1, spp=4, shard=(rank, world)); torch.cuda.synchronize()
            ep = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            ep[0].record()
            n_pt = max(a.steps, 3)
            for i in range(n_pt):
                pit.render(scene, sensor=psens, seed=10 + i, spp=4, shard=(rank, world))
            ep[1].record(); torch.cuda.synchronize()
            ms_pt = torch.tensor([ep[0].elapsed_time(ep[1]) / n_pt], device="cuda", dtype=torch.float64)
            if world > 1:
                dist.all_reduce(ms_pt, op=dist.ReduceOp.MAX)
            pst = scene.stats()
            # its lens adjoint (lrt_ptrace_backward: every connection of the same light paths differentiated), rank 0's share
            gimg = torch.ones((576, 1024, 3), dtype=torch.float32, device="cuda")
            pit.render_backward_lens(scene, gimg, sensor=psens, seed=1, spp=4, shard=(rank, world)); torch.cuda.synchronize()
            eb2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            eb2[0].record()
            for i in range(3):
                pit.render_backward_lens(scene, gimg, sensor=psens, seed=10 + i, spp=4, shard=(rank, world))
            eb2[1].record(); torch.cuda.synchronize()
            ms_ptb = eb2[0].elapsed_time(eb2[1]) / 3
            ptr = {"value": 1024 * 576 * 4 / (ms_pt.item() * 1e-3) / 1e6, "unit": "M light paths/s", "ms_per_render": ms_pt.item(),
                   "lens_adjoint_ms": ms_ptb, "lens_adjoint_M_paths_per_s": 1024 * 576 * 4 / world / (ms_ptb * 1e-3) / 1e6,
                   "film": [1024, 576], "spp": 4, "connections_per_path": pst["lens_misses"] / max(pst["samples"], 1),
                   "rays_per_path": pst["rays"] / max(pst["samples"], 1),
                   "what": "lrt_ptrace (queue kernel) on the headline scene and lens: visible emitters + emitter rays + scene bounces, every "
                           "vertex connected to the lens (sample_direction) and splatted (normalised, atomics); per-rank counters"}
        except Exception as e:     # the extra measurement must never take the headline line down
            ptr = {"error": str(e)[:200]}
    lens_grad = None if a.no_lens_grad else lens_grad_section(mi, torch, dist, a, rank, world)
    pupil = None
    if rank == 0 and not a.no_lens_grad:   # exit-pupil LUT of the workload's lens: 64 x 2^20 lens traces in one kernel
        sens = scene.sensors()[0]
        sens.set_exit_pupil(True, 1.0); torch.cuda.synchronize()      # warm-up
        t0 = time.perf_counter()
        sens.set_exit_pupil(True, 1.0)                                # synchronous: kernel + 1.3 KB read-back
        sec = time.perf_counter() - t0
        sens.set_exit_pupil(False, 0.0)
        pupil = {"ms": sec * 1e3, "lens_rays": 64 << 20, "Grays_per_s": (64 << 20) / sec / 1e9,
                 "what": "lrt_sensor_set_exit_pupil: compute_exit_pupil_bounds of the `%s` lens (host wall clock, synchronous call)" % a.lens}

    extras = None
    if rank == 0 and not a.no_lens_grad:
        try:
            extras = widened_rows_section(mi, torch, a, scene)
        except Exception as e:     # the extra measurements must never take the headline line down
            extras = {"error": str(e)[:200]}

    if rank == 0:
        value = paths_per_step / (ms * 1e-3) / 1e6
        e2e_v = paths_per_step / (ms_e2e * 1e-3) / 1e6
        # algorithmic bytes of the dominant (only) kernel per launch: BVH nodes + triangles fetched, splat taps, film
        fp = 25 if a.spp_per_step >= 4 else 16
        alg_bytes = stats["bvh_nodes"] * 80 + stats["triangles"] * 48 + stats["samples"] * fp * 16 + H * W * 16 * world
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        peak = peaks.get("hbm_gbs", 6546.9)
        ach = alg_bytes / world / (ms * 1e-3) / 1e9
        traffic, traffic_src, ncu_facts = None, None, None
        tj = os.path.join(ROOT, "profiles", "render_traffic.json")
        if os.path.exists(tj):   # dram__bytes_read.sum + dram__bytes_write.sum of one k_render launch, from ncu --set full
            t = json.load(open(tj))
            if t.get("spp_per_step") == a.spp_per_step and t.get("width") == a.width and t.get("height") == a.height and world == 1:
                traffic, traffic_src = t["dram_bytes_per_launch"] / 1e9, t.get("source")
                ncu_facts = {k: t[k] for k in ("issue_slot_utilisation", "active_lanes_per_instruction", "l1_hit_rate", "l2_hit_rate",
                                               "top_stalls_per_issue") if k in t}
        # measured ceilings of this box (lrt_measure_peaks: FMA / issue / L1 / L2 / shared-atomic micro-kernels) and the stage
        # rooflines SURVEY.md 8(d) asks for: counters per launch from the committed ncu capture of this very workload
        # (profiles/render_traffic.json), time measured live
        pk = (C.c_double * 6)()
        capi.check(capi.lib.lrt_measure_peaks(pk, None))
        measured = {"fp32_tflops": pk[0], "issue_gwarp_inst_per_s": pk[1], "l1_hit_gbs": pk[2], "l2_hit_gbs": pk[3],
                    "smem_atomic_g_per_s": pk[4], "sm_mhz": pk[5], "hbm_gbs": peak,
                    "how": "lrt_measure_peaks (csrc/lrt_peaks.cu), best of 3, CUDA events; hbm_gbs from MEASURED_PEAKS.json"}
        stages = None
        if tj and os.path.exists(tj):
            t = json.load(open(tj))
            if t.get("counters") and traffic is not None:
                c, sec = t["counters"], ms * 1e-3
                inst, tinst = c["smsp__inst_executed.sum"], c["smsp__thread_inst_executed.sum"]
                stages = [
                    {"stage": "whole kernel (lens trace, traversal, shading)", "bound": "issue", "achieved": inst / sec / 1e9, "peak": pk[1],
                     "unit": "G warp-inst/s", "frac": inst / sec / 1e9 / pk[1], "lane_frac": tinst / inst / 32.0,
                     "note": "frac x lane_frac = share of the SMs' lane-issue capacity doing useful work"},
                    {"stage": "FP32 pipe (FFMA/FMUL/FADD)", "bound": "fp32", "achieved": 2 * c.get("sm__inst_executed_pipe_fma.sum", 0) * 32 / sec / 1e12,
                     "peak": pk[0], "unit": "TFLOP/s (warp-level FMA-pipe instructions x 32 x 2)",
                     "frac": 2 * c.get("sm__inst_executed_pipe_fma.sum", 0) * 32 / sec / 1e12 / pk[0]},
                    {"stage": "BVH traversal + stack (L1)", "bound": "l1", "achieved": c["l1tex__t_sectors.sum"] * 32 / sec / 1e9, "peak": pk[2],
                     "unit": "GB/s of L1 sectors", "frac": c["l1tex__t_sectors.sum"] * 32 / sec / 1e9 / pk[2]},
                    {"stage": "BVH traversal (L2)", "bound": "l2", "achieved": c["lts__t_sectors.sum"] * 32 / sec / 1e9, "peak": pk[3],
                     "unit": "GB/s of L2 sectors", "frac": c["lts__t_sectors.sum"] * 32 / sec / 1e9 / pk[3]},
                    {"stage": "DRAM", "bound": "hbm", "achieved": t["dram_bytes_per_launch"] / sec / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": t["dram_bytes_per_launch"] / sec / 1e9 / peak, "note": "the scene lives in L2: DRAM is not a bound"},
                ]
        traced = stats["samples"] - stats["lens_misses"]
        line = {"metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": config,
                "e2e": {"value": e2e_v, "unit": "Mpaths/s", "h2d_bytes_per_step": int(host_lens.numel() * 4),
                        "d2h_bytes_per_step": int(H * W * 3 * 4), "ms_per_step": ms_e2e},
                "gpu_launches": int(launches),
                "paths": {"samples_per_step": int(stats["samples"]), "lens_miss_fraction": stats["lens_misses"] / max(stats["samples"], 1),
                          "traced_Mpaths_per_s": traced / (ms * 1e-3) / 1e6,
                          "note": "value counts every film sample (W*H*spp, the metric's definition); vignetted samples trace no ray, "
                                  "traced_Mpaths_per_s counts only the paths that left the lens"},
                "measured_peaks": measured, "roofline_stages": stages,
                "clocks": sampler.result(),
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                             "traffic_unit": "GB per launch (ncu dram read+write)", "traffic_source": traffic_src,
                             "algorithmic_gb_per_launch": alg_bytes / world / 1e9,
                             "ncu": ncu_facts,   # what actually bounds the kernel (issue slots / latency), from the same capture
                             "kernel": "lrt::k_render", "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)",
                             "dram_achieved_gbs": (traffic / (ms * 1e-3)) if traffic is not None else None,
                             "note": "contract figure: algorithmic bytes (nodes*80 + triangles*48 + samples*taps*16 + film) / launch time vs the "
                                     "measured HBM copy rate. Those bytes are served by L1/L2 (DRAM traffic per launch = `traffic`, "
                                     "dram_achieved_gbs): the kernel is issue / latency bound, see roofline_stages and DESIGN.md 3"},
                "stats": {k: int(v) for k, v in stats.items()}, "scene_build_s": build_s,
                "bvh": scene.info()}
        line["render_backward"] = {"value": paths_per_step / (ms_bwd.item() * 1e-3) / 1e6, "unit": "Mpaths/s", "ms_per_step": ms_bwd.item(),
                                   "what": "lrt_render_backward on the same workload: gradients of diffuse reflectance / conductor "
                                           "specular reflectance / area-light radiance (one-pass PRB)"}
        if ptr is not None:
            if "value" in ptr and not a.no_cpu_baseline and world == 1:
                pr = reference_sample(a, extra=("--integrator", "ptracer", "--width", "1024", "--height", "576", "--spp", "4"))
                if pr is not None:
                    ptr["cpu_baseline"] = {"value": pr[0]["value"], "unit": "M light paths/s", "cores": pr[0]["cores"], "kind": "reference",
                                           "sample": "the reference's scalar_rgb ptracer on the same scene, 1024x576 film, 4 spp"}
            line["ptracer"] = ptr
        if lens_grad is not None:
            line["lens_grad"] = lens_grad
        if pupil is not None:
            line["exit_pupil_lut"] = pupil
        if extras is not None:
            line["widened_rows"] = extras
        if not a.no_cpu_baseline and world == 1:     # the contract wants the CPU baseline on rank 0 at N = 1 only
            port, _ = cpu_sample(a, scene.arrays, d["sensor"], keep_image=True)
            ref = reference_sample(a)
            line["cpu_baseline_port"] = {k: v for k, v in port.items() if k not in ("image", "lens_miss")}
            line["cpu_baseline"] = ref[0] if ref else line["cpu_baseline_port"]
            # parity on the headline workload itself: the same centred crop, same seed, same spp on the GPU vs the oracle port
            line["parity"] = parity_crop(mi, torch, a, d, port)
            if lens_grad is not None:
                line["lens_grad"]["cpu_baseline"] = lens_grad_cpu(mi, a)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
