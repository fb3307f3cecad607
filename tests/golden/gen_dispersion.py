#!/usr/bin/env python3
"""Generates tests/golden/dispersion.npz from the REFERENCE's own lens code instantiated as a SPECTRAL variant
(oracle/_ref/libreflens.so: RefLensCamera<float, mitsuba::Spectrum<float, 4>>, compiled from /root/reference by
`make -C oracle ref`): per lens design the interface positions after the spectral variants' autofocus and the nominal IORs,
Sensor::sample_ray and the world-side lens trace with one wavelength per ray — the fork's dispersion mode
(DispersiveMaterial::compute_ior(ray), realisticlens.cpp:169-183). Run in the build container only:

  python tests/golden/gen_dispersion.py
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
R.ref_scam_create.restype = C.c_void_p
R.ref_scam_film_z.restype = c_f
rng = np.random.default_rng(20261105)
film = FilmSpec(256, 192, [200, 100, 20, 30], 35.0)
N = 384
out = {}
for name in ld.all_designs():
    tw = np.ascontiguousarray(EYE, f32)
    h = C.c_void_p(R.ref_scam_create(name.encode(), C.c_int(film.width), C.c_int(film.height), ip(film.crop), c_f(film.diagonal_mm), fp(tw),
                                     c_f(1e-2), c_f(1e4), c_f(0.0), C.c_int(0), None, None))
    assert h.value, name
    n = R.ref_scam_num_interfaces(h)
    tab = np.zeros((n, 4), f32)
    for i in range(n):
        R.ref_scam_interface(h, C.c_int(i), fp(tab[i]))
    out[name + ".table"] = tab            # z, element radius, ior_left, ior_right at Wavelength(589.3f)
    pos = rng.random((N, 2), dtype=f32); ap = rng.random((N, 2), dtype=f32)
    ap[N // 2:] = 0.5 + (ap[N // 2:] - 0.5) * 0.05
    pos[3 * N // 4:] = 0.5 + (pos[3 * N // 4:] - 0.5) * 0.05      # long stacks only transmit near the axis
    wl = (380.0 + 320.0 * rng.random(N)).astype(f32)
    wl[:3] = [589.3, 486.1, 656.3]
    ray = np.zeros((N, 10), f32)
    R.ref_scam_sample_ray(h, C.c_int(N), fp(pos), fp(ap), fp(wl), fp(ray))
    out[name + ".pos"] = pos; out[name + ".ap"] = ap; out[name + ".wl"] = wl; out[name + ".ray"] = ray
    front_z, front_r = tab[-1, 0], tab[-1, 1]
    od = np.zeros((N, 6), f32)
    od[:, 0:2] = (rng.random((N, 2), dtype=f32) - 0.5) * 2 * front_r
    od[:, 2] = front_z + 0.05
    d = np.zeros((N, 3), f32); d[:, 0:2] = (rng.random((N, 2), dtype=f32) - 0.5) * 0.3; d[:, 2] = -1
    od[:, 3:] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    tr = np.zeros((N, 7), f32)
    R.ref_scam_trace(h, C.c_int(1), C.c_int(N), fp(od), fp(wl), fp(tr))
    out[name + ".world_od"] = od; out[name + ".world_out"] = tr
    # the same rays, all at one blue and one red wavelength: the chromatic shift the mode exists for
    for tag, w in (("blue", 450.0), ("red", 650.0)):
        t2 = np.zeros((N, 7), f32)
        R.ref_scam_trace(h, C.c_int(1), C.c_int(N), fp(od), fp(np.full(N, w, f32)), fp(t2))
        out[f"{name}.world_{tag}"] = t2
    if name in ("doublet", "gauss"):
        # the spectral variants' exit-pupil LUT (64 x 2^20 rays, each with its own Halton wavelength) and sample_ray through it
        lut = np.zeros((64, 4), f32)
        R.ref_scam_compute_exit_pupil(h, fp(lut))
        R.ref_scam_set_shutter_open_time(h, c_f(1.0))
        rayp = np.zeros((N, 10), f32)
        R.ref_scam_sample_ray(h, C.c_int(N), fp(pos), fp(ap), fp(wl), fp(rayp))
        out[name + ".pupil_lut"] = lut; out[name + ".pupil_ray"] = rayp
        print(f"   spectral exit-pupil LUT: sample_ray active {(rayp[:, 7] > 0).mean():.2f}")
    # Sensor::sample_direction of the spectral instantiation: distant / near reference points in camera space
    pd_ = np.zeros((N, 3), f32)
    ang = np.radians(rng.random(N) * 6.0); phi = rng.random(N) * 2 * np.pi
    dist = np.where(np.arange(N) % 2 == 0, 1000.0, 0.5 + rng.random(N) * 5)
    pd_[:, 0] = dist * np.tan(ang) * np.cos(phi); pd_[:, 1] = dist * np.tan(ang) * np.sin(phi); pd_[:, 2] = front_z + dist
    s2 = rng.random((N, 2), dtype=f32); s2[N // 2:] = 0.5 + (s2[N // 2:] - 0.5) * 0.1
    od11 = np.zeros((N, 11), f32)
    R.ref_scam_sample_direction(h, C.c_int(N), fp(pd_), fp(s2), fp(od11))
    out[name + ".dir_p"] = pd_; out[name + ".dir_s"] = s2; out[name + ".dir_out"] = od11
    act = ray[:, 7] > 0
    print(f"{name:14s} {n:2d} interfaces, sample_ray active {act.mean():.2f}, world trace active {(tr[:, 6] > 0).mean():.2f}")
    R.ref_scam_destroy(h)
np.savez_compressed(os.path.join(HERE, "dispersion.npz"), **out)
print("dispersion.npz", os.path.getsize(os.path.join(HERE, "dispersion.npz")) // 1024, "KiB")
