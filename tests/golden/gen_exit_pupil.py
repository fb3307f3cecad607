#!/usr/bin/env python3
"""Golden vectors for exit-pupil sampling, generated with the REFERENCE's own code (oracle/_ref/libreflens.so, built from
/root/reference by oracle/Makefile; build container only):
  * RadicalInverse::eval (include/mitsuba/core/qmc.h:68-91) for the bases 2, 3, 5;
  * compute_exit_pupil_bounds (src/sensors/realisticlens.cpp:1685-1809): the 64 x 4 LUT of a few lens designs;
  * sample_ray with sample_exit_pupil = true and shutter_open_time = 1 (:1812-1845, 2223-2291).
usage: python tests/golden/gen_exit_pupil.py    (about 20 s per lens: the reference traces 64 x 2^20 rays on one core)"""
import os
import sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import FilmSpec, cam_exit_pupil, cam_sample_ray, radical_inverse, ref, ref_camera  # noqa: E402

f32 = np.float32
R = ref()
assert R is not None, "needs oracle/_ref (make -C oracle ref)"
rng = np.random.default_rng(20261107)
out = {}
idx = np.concatenate([np.arange(0, 1024), rng.integers(0, 1 << 20, 3072), [(1 << 20) - 1, (1 << 32) + 12345]]).astype(np.uint64)
out["ri.index"] = idx
for b in range(3):
    out[f"ri.base{b}"] = radical_inverse(R, "ref", b, idx)
FILMS = {"doublet": FilmSpec(256, 256, None, 35.0), "gauss": FilmSpec(256, 192, [200, 100, 20, 30], 35.0),
         "asph": FilmSpec(128, 128, None, 20.0), "exp1a": FilmSpec(256, 256, None, 35.0)}
N = 2048
for name, film in FILMS.items():
    h = ref_camera(name, film, shutter_open_time=1.0)
    out[name + ".film"] = np.array([film.width, film.height, *film.crop, film.diagonal_mm], np.float64)
    out[name + ".bounds"] = cam_exit_pupil(R, "ref", h)
    pos = rng.random((N, 2), dtype=f32); ap = rng.random((N, 2), dtype=f32)
    pos[0] = 0.5; ap[0] = 0.5
    out[name + ".pos"] = pos; out[name + ".ap"] = ap
    out[name + ".ray"] = cam_sample_ray(R, "ref", h, pos, ap)
    print(name, "transmitted", float((out[name + ".ray"][:, 7] > 0).mean()), "bounds[0]", out[name + ".bounds"][0])
    R.ref_cam_destroy(h)
np.savez_compressed(os.path.join(HERE, "exit_pupil.npz"), **out)
print("exit_pupil.npz", os.path.getsize(os.path.join(HERE, "exit_pupil.npz")) // 1024, "KiB")
