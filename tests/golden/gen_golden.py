#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REFERENCE's own code (oracle/_ref/libreflens.so, compiled from
/root/reference by `make -C oracle ref`). Run in the build container only; the vectors are committed
because /root/reference does not exist on the GPU box.

  python tests/golden/gen_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle_lib import *  # noqa

ld = load_lens_designs()

R = ref()
assert R is not None, "needs oracle/_ref (build container)"
rng = np.random.default_rng(20261104)

# ---- RNG
pcg = {}
for (st, sq) in [(0x853c49e6748fea9b, 0xda3e39cb94b95bdb), (0, 0xda3e39cb94b95bdb), (42, 54), (0xffffffff, 0xfffffffe)]:
    u = np.zeros(64, np.uint32); f = np.zeros(64, f32)
    R.ref_pcg32_u32(C.c_uint64(st), C.c_uint64(sq), 64, u.ctypes.data_as(UP))
    R.ref_pcg32_f32(C.c_uint64(st), C.c_uint64(sq), 64, fp(f))
    pcg[f"{st}_{sq}"] = (u, f)
tea_in = rng.integers(0, 2**32, size=(256, 2), dtype=np.uint64).astype(np.uint32)
tea_out = np.zeros((256, 2), np.uint32)
for i in range(256):
    R.ref_tea32(C.c_uint32(int(tea_in[i, 0])), C.c_uint32(int(tea_in[i, 1])), 4, tea_out[i].ctypes.data_as(UP))
np.savez_compressed(os.path.join(HERE, "rng.npz"), tea_in=tea_in, tea_out=tea_out,
                    **{f"pcg_u_{k}": v[0] for k, v in pcg.items()}, **{f"pcg_f_{k}": v[1] for k, v in pcg.items()})

# ---- small maths
N = 4096
uv = rng.random((N, 2), dtype=f32)
uv[:8] = [[0.5, 0.5], [0, 0], [1 - 2**-24, 1 - 2**-24], [0.5, 0.25], [0.25, 0.5], [0, 0.5], [0.5, 0], [0.75, 0.75]]
disk = np.zeros((N, 2), f32); R.ref_concentric_disk(N, fp(uv), fp(disk))
hemi = np.zeros((N, 3), f32); R.ref_cosine_hemisphere(N, fp(uv), fp(hemi))
ci = (rng.random(N, dtype=f32) * 2 - 1).astype(f32); ci[:4] = [0, 1, -1, 1e-4]
eta = (rng.random(N, dtype=f32) * 1.4 + 0.6).astype(f32); eta[:4] = [1.5, 1.0, 1.5, 0.66]
fres = np.zeros((N, 4), f32); R.ref_fresnel(N, fp(ci), fp(eta), fp(fres))
abc = rng.standard_normal((N, 3)).astype(f32); abc[:3] = [[0, 1, 2], [0, 0, 1], [1, 0, -4]]
quad = np.zeros((N, 3), f32); R.ref_solve_quadratic(N, fp(abc), fp(quad))
np.savez_compressed(os.path.join(HERE, "maths.npz"), uv=uv, disk=disk, hemi=hemi, cos_i=ci, eta=eta, fresnel=fres,
                    abc=abc, quad=quad)

# ---- lens: per design the interface table after autofocus, sample_ray and world-side traces
out = {}
N = 512
FILMS = {"default": FilmSpec(256, 192, [200, 100, 20, 30], 35.0), "small": FilmSpec(128, 128, None, 4.0)}
for name in ld.all_designs():
    for fname, film in FILMS.items():
        h = ref_camera(name, film)
        key = f"{name}.{fname}"
        out[key + ".ifs"] = cam_interfaces(R, "ref", h)
        m = np.zeros(16, f32); R.ref_cam_sample_to_film(h, fp(m)); out[key + ".s2f"] = m
        pos = rng.random((N, 2), dtype=f32); ap = rng.random((N, 2), dtype=f32)
        pos[0] = ap[0] = 0.5   # the on-axis ray (killed in exp1* by a t<0 rounding)
        # concentrate half of the aperture samples near the axis so that long stacks transmit some rays
        ap[N // 2:] = 0.5 + (ap[N // 2:] - 0.5) * 0.05
        out[key + ".pos"] = pos; out[key + ".ap"] = ap
        out[key + ".ray"] = cam_sample_ray(R, "ref", h, pos, ap)
        front = out[key + ".ifs"][-1]
        od = np.zeros((N, 6), f32)
        od[:, 0:2] = (rng.random((N, 2), dtype=f32) - 0.5) * 2 * front[2]
        od[:, 2] = front[1] + 0.05
        d = np.zeros((N, 3), f32); d[:, 0:2] = (rng.random((N, 2), dtype=f32) - 0.5) * 0.3; d[:, 2] = -1
        od[:, 3:] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
        out[key + ".world_od"] = od
        out[key + ".world_out"] = cam_trace(R, "ref", h, od, True)
        R.ref_cam_destroy(h)
np.savez_compressed(os.path.join(HERE, "lens.npz"), **out)
for k in ("rng.npz", "maths.npz", "lens.npz"):
    print(k, os.path.getsize(os.path.join(HERE, k)) // 1024, "KiB")

# ---- Sensor::sample_direction (realisticlens.cpp:2301-2394): distant / near reference points in camera space
rng2 = np.random.default_rng(20261105)
out = {}
N = 256
for name in ld.all_designs():
    film = FILMS["default"]
    h = ref_camera(name, film)
    ifs = cam_interfaces(R, "ref", h)
    front_z, front_r = ifs[-1, 1], ifs[-1, 2]
    p = np.zeros((N, 3), f32)
    ang = np.radians(rng2.random(N) * 6.0); phi = rng2.random(N) * 2 * np.pi
    dist = np.where(np.arange(N) % 2 == 0, 1000.0, 0.5 + rng2.random(N) * 5)
    p[:, 0] = dist * np.tan(ang) * np.cos(phi); p[:, 1] = dist * np.tan(ang) * np.sin(phi); p[:, 2] = front_z + dist
    p[0] = [0, 0, front_z + 10.0]
    s2 = rng2.random((N, 2), dtype=f32); s2[0] = 0.5
    s2[N // 2:] = 0.5 + (s2[N // 2:] - 0.5) * 0.1
    o = np.zeros((N, 11), f32)
    R.ref_cam_sample_direction(h, N, fp(p), fp(s2), fp(o))
    out[name + ".p"] = p; out[name + ".s"] = s2; out[name + ".out"] = o
    R.ref_cam_destroy(h)
np.savez_compressed(os.path.join(HERE, "sample_direction.npz"), **out)
print("sample_direction.npz", os.path.getsize(os.path.join(HERE, "sample_direction.npz")) // 1024, "KiB",
      {k[:-4]: int((v[:, 9] > 0).sum()) for k, v in out.items() if k.endswith(".out")})
