"""CPU suite, part 2: liblensrt.so loads, exports every symbol include/lensrt.h declares, refuses compute
without a GPU (no CPU fallback), and its HOST-side lens construction matches the oracle bit for bit."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle_lib import *  # noqa

ld = load_lens_designs()


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "lensrt.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(lrt_[a-z0-9_]+)\s*\(", txt)))


def test_exports_every_declared_symbol(mi):
    syms = header_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(mi.capi.lib, s), f"liblensrt.so does not export {s}"


def test_no_cpu_fallback(mi):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    out = np.zeros(8, f32)
    rc = mi.capi.lib.lrt_sampler_draw_host(C.c_uint32(0), C.c_uint32(0), C.c_int64(4), C.c_int32(2), fp(out))
    assert rc == -2 and b"no CUDA device" in mi.capi.lib.lrt_last_error()
    s = mi.RealisticLensCamera(mi.Film(64, 64), "doublet")
    with pytest.raises(mi.capi.LensrtError):
        s.sample_ray(0, 0, np.full((4, 2), 0.5, f32), np.full((4, 2), 0.5, f32))


@pytest.mark.parametrize("name", ld.all_designs())
def test_host_lens_construction_matches_oracle(mi, golden, name):
    g = golden["lens"]
    film = mi.Film(256, 192, crop=(200, 100, 20, 30), diagonal=35.0)
    s = mi.RealisticLensCamera(film, name)
    assert (bits(s.interfaces()) == bits(g[f"{name}.default.ifs"])).all()
    assert (bits(s.info()["sample_to_film"].ravel()) == bits(g[f"{name}.default.s2f"])).all()
    od = g[f"{name}.default.world_od"]
    o, d, act = s.trace(od[:, :3], od[:, 3:], from_world=True)
    want = g[f"{name}.default.world_out"]
    assert (act == (want[:, 6] > 0)).all()
    assert (bits(np.concatenate([o, d], 1)[act]) == bits(want[act, :6])).all()


def test_set_interfaces_roundtrip(mi):
    s = mi.RealisticLensCamera(mi.Film(64, 64), "exp2a")
    rows = s.interfaces()
    s.set_interfaces(rows)
    assert (bits(s.interfaces()) == bits(rows)).all()


def test_blocked_focus_ray_is_an_error(mi):
    rows = ld.doublet(6.0, 0.025, 0.005).rows.copy()
    rows[:, 2] = 1e-6   # semi-diameters too small for the paraxial focus rays (realisticlens.cpp:2132-2134)
    with pytest.raises(mi.capi.LensrtError):
        mi.RealisticLensCamera(mi.Film(64, 64), "custom", prescription=ld.InterfaceList(rows))


def test_dispersive_bsdf_row_is_a_dielectric_at_589nm(mi):
    """dispersive_dielectric.cpp:425-431 at the RGB variants' fixed 589.3 nm: eta = n_d(int) / n_d(ext) up to float32 rounding."""
    r = mi.bsdf_row({"type": "dispersive", "int_ior": 1.7, "int_V_d": 30.0})
    assert r[0] == 2 and abs(float(r[14]) - 1.7) < 1e-6
    r2 = mi.bsdf_row({"type": "dispersive"})
    assert abs(float(r2[14]) - 1.5046) < 1e-6
