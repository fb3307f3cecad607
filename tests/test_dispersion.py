"""Dispersion mode (the fork's spectral variants): every ray carries a wavelength and each interface refracts
with the Cauchy / Sellmeier IORs of its media at that wavelength (DispersiveMaterial::compute_ior(ray),
realisticlens.cpp:169-183,226-240). Golden vectors: the reference's own lens classes and camera methods INSTANTIATED AS A
SPECTRAL VARIANT in oracle/_ref (tests/golden/gen_dispersion.py -> dispersion.npz), all 15 designs. CPU part: the table a
spectral variant builds (autofocus with Wavelength(589.3f) rays) and the host twin of the wavelength-aware trace, bit for bit;
the kernels are compared with the same vectors on the GPU (bottom)."""
import os

import numpy as np
import pytest

from oracle_lib import *  # noqa

ld = load_lens_designs()
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def G():
    return dict(np.load(os.path.join(HERE, "golden", "dispersion.npz")))


def film(mi):
    return mi.Film(256, 192, crop=(200, 100, 20, 30), diagonal=35.0)


@pytest.mark.parametrize("name", ld.all_designs())
def test_spectral_table_and_host_trace_match_the_reference(mi, G, name):
    s = mi.RealisticLensCamera(film(mi), name, spectral=True)
    rows = s.interfaces()
    tab = G[name + ".table"]
    assert (bits(rows[:, 1]) == bits(tab[:, 0])).all() and (bits(rows[:, 2]) == bits(tab[:, 1])).all()       # z after autofocus, radii
    assert (bits(rows[:, 13]) == bits(tab[:, 2])).all() and (bits(rows[:, 14]) == bits(tab[:, 3])).all()     # IORs at 0.001f * 589.3f um
    od, wl = G[name + ".world_od"], G[name + ".wl"]
    for key, w in ((".world_out", wl), (".world_blue", np.full(len(od), 450.0, f32)), (".world_red", np.full(len(od), 650.0, f32))):
        o, d, act = s.trace_wavelength(od[:, :3], od[:, 3:], w, from_world=True, device=False)
        want = G[name + key]
        assert (act == (want[:, 6] > 0)).all(), key
        assert (bits(np.concatenate([o, d], 1)[act]) == bits(want[act, :6])).all(), key


def test_the_rgb_table_differs_from_the_spectral_one_by_the_nominal_wavelength(mi, G):
    """0.5893f vs 0.001f * 589.3f: one ulp in the wavelength, visible in the IORs (and through the autofocus in the z's)"""
    a = mi.RealisticLensCamera(film(mi), "gauss").interfaces()
    b = mi.RealisticLensCamera(film(mi), "gauss", spectral=True).interfaces()
    assert np.abs(a[:, 13:15] - b[:, 13:15]).max() < 5e-7 and np.abs(a[:, 1] - b[:, 1]).max() < 1e-6
    assert (bits(b[:, 13]) == bits(G["gauss.table"][:, 2])).all()


def test_blue_focuses_shorter_than_red(mi, G):
    """the point of the mode: longitudinal chromatic aberration. A ray parallel to the axis through the biconvex singlet (one
    normal-dispersion Cauchy glass) crosses the axis closer to the lens in blue than in red."""
    s = mi.RealisticLensCamera(film(mi), "singlet", spectral=True)
    rows = s.interfaces()
    h = 0.3 * rows[-1, 2]
    o = np.array([[h, 0.0, rows[-1, 1] + 0.01]], f32); d = np.array([[0.0, 0.0, -1.0]], f32)
    z = {}
    for w in (450.0, 650.0):
        po, pd, act = s.trace_wavelength(o, d, w, from_world=True, device=False)
        assert act.all()
        z[w] = float(po[0, 2] - po[0, 0] * pd[0, 2] / pd[0, 0])          # axis crossing
    assert z[450.0] > z[650.0] + 1e-5                                     # blue crosses nearer to the lens (larger z: film at z = 0)


def test_materials_roundtrip_and_errors(mi):
    s = mi.RealisticLensCamera(film(mi), "tessar")
    m = s.materials()
    assert m.shape == (8, 2, 7)
    rows = s.interfaces()
    c = mi.RealisticLensCamera(film(mi), "custom", prescription=ld.InterfaceList(rows, film_z=0.0, fine_focus=False))
    od = np.array([[0.0, 0.0, rows[-1, 1] + 0.01, 0.0, 0.0, -1.0]], f32)
    with pytest.raises(mi.capi.LensrtError):
        c.trace_wavelength(od[:, :3], od[:, 3:], 550.0, from_world=True, device=False)      # no dispersion data yet
    c.set_materials(m)
    a = s.trace_wavelength(od[:, :3], od[:, 3:], 550.0, from_world=True, device=False)
    b = c.trace_wavelength(od[:, :3], od[:, 3:], 550.0, from_world=True, device=False)
    assert (bits(np.concatenate(a[:2], 1)) == bits(np.concatenate(b[:2], 1))).all()
    with pytest.raises(mi.capi.LensrtError):
        c.set_materials(m[:3])


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", ld.all_designs())
def test_gpu_dispersion_kernels_match_the_reference(mi, G, name):
    """lrt_sensor_sample_ray_wl and lrt_sensor_trace_wl against the reference's spectral instantiation: rays, weights and miss
    masks bit for bit, one wavelength per ray."""
    s = mi.RealisticLensCamera(film(mi), name, spectral=True)
    o, d, maxt, w = s.sample_ray_wavelength(G[name + ".pos"], G[name + ".ap"], G[name + ".wl"])
    got = np.concatenate([o.cpu().numpy(), d.cpu().numpy(), maxt.cpu().numpy()[:, None], w.cpu().numpy()[:, None]], 1)
    want = G[name + ".ray"]
    act = want[:, 7] > 0
    assert ((got[:, 7] > 0) == act).all()
    assert (bits(got[act]) == bits(want[act, :8])).all()
    od = G[name + ".world_od"]
    for key, wl in ((".world_out", G[name + ".wl"]), (".world_blue", 450.0), (".world_red", 650.0)):
        po, pd, a = s.trace_wavelength(od[:, :3], od[:, 3:], wl, from_world=True)
        want = G[name + key]
        a = a.cpu().numpy()
        assert (a == (want[:, 6] > 0)).all(), key
        assert (bits(np.concatenate([po.cpu().numpy(), pd.cpu().numpy()], 1)[a]) == bits(want[a, :6])).all(), key


@pytest.mark.gpu
def test_gpu_dispersion_limits(mi, G):
    """no rays; a sensor without dispersion data; exit-pupil sampling (whose spectral LUT is not built) is refused."""
    import ctypes as C
    import torch
    capi, lib = mi.capi, mi.capi.lib
    s = mi.RealisticLensCamera(film(mi), "gauss", spectral=True)
    assert lib.lrt_sensor_trace_wl(s.handle, C.c_int32(1), C.c_int64(0), None, None, None, None) == 0
    assert lib.lrt_sensor_sample_ray_wl(s.handle, C.c_int64(0), None, None, None, None, None) == 0
    rows = s.interfaces()
    c = mi.RealisticLensCamera(film(mi), "custom", prescription=ld.InterfaceList(rows, film_z=0.0, fine_focus=False))
    with pytest.raises(mi.capi.LensrtError):
        c.sample_ray_wavelength(np.full((4, 2), 0.5, f32), np.full((4, 2), 0.5, f32), 550.0)
    # a LUT built the rgb way (constant IORs) is not the spectral variants' LUT: refused for per-ray wavelengths
    p = mi.RealisticLensCamera(film(mi), "gauss", spectral=True)
    capi.check(lib.lrt_sensor_set_exit_pupil(p.handle, C.c_int32(1), C.c_float(1.0), None))
    with pytest.raises(mi.capi.LensrtError):
        p.sample_ray_wavelength(np.full((4, 2), 0.5, f32), np.full((4, 2), 0.5, f32), 550.0)
    # a scalar wavelength broadcasts; 589.3 nm through the spectral table = the table's own nominal IORs
    o, d, maxt, w = s.sample_ray_wavelength(G["gauss.pos"][:64], G["gauss.ap"][:64], 589.3)
    o2, d2, maxt2, w2 = s.sample_ray(0.0, None, torch.tensor(G["gauss.pos"][:64], device="cuda"), torch.tensor(G["gauss.ap"][:64], device="cuda"))
    assert (bits(o.cpu().numpy()) == bits(o2.cpu().numpy())).all() and (bits(d.cpu().numpy()) == bits(d2.cpu().numpy())).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["doublet", "gauss"])
def test_gpu_spectral_exit_pupil_lut_and_sampling(mi, G, name):
    """compute_exit_pupil_bounds as a spectral variant runs it — every ray of the 64 x 2^20-ray build carries the wavelength
    lerp(380, 700, RadicalInverse(2, i)) nm (realisticlens.cpp:1718-1727) — and sample_ray through that LUT with per-ray
    wavelengths: LUT, rays, weights and miss masks bit for bit against the reference's spectral instantiation."""
    s = mi.RealisticLensCamera(film(mi), name, spectral=True, sample_exit_pupil=True, shutter_open=0.0, shutter_close=1.0)
    lut = s.exit_pupil_bounds()
    assert (bits(lut) == bits(G[name + ".pupil_lut"])).all()
    rgb = mi.RealisticLensCamera(film(mi), name, sample_exit_pupil=True).exit_pupil_bounds()
    assert np.abs(rgb - lut).max() > 0                                  # (not the constant-IOR LUT)
    o, d, maxt, w = s.sample_ray_wavelength(G[name + ".pos"], G[name + ".ap"], G[name + ".wl"])
    got = np.concatenate([o.cpu().numpy(), d.cpu().numpy(), maxt.cpu().numpy()[:, None], w.cpu().numpy()[:, None]], 1)
    want = G[name + ".pupil_ray"]
    act = want[:, 7] > 0
    assert act.mean() > 0.5 and ((got[:, 7] > 0) == act).all()
    assert (bits(got[act]) == bits(want[act, :8])).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ld.all_designs())
def test_gpu_spectral_sample_direction(mi, G, name):
    """Sensor::sample_direction of a spectral variant: the reference traces the connection at Wavelength(589.3f) whatever the
    ray carries (realisticlens.cpp:2339-2344, its own TODO), i.e. with the nominal IORs of the spectral table — what
    lrt_sensor_sample_direction does on a sensor built with spectral=True. Bit for bit, all designs."""
    import ctypes as C
    s = mi.RealisticLensCamera(film(mi), name, spectral=True)
    p, s2, want = G[name + ".dir_p"], G[name + ".dir_s"], G[name + ".dir_out"]
    got = np.zeros_like(want)
    mi.capi.check(mi.capi.lib.lrt_sensor_sample_direction_host(s.handle, C.c_int64(len(p)), fp(np.ascontiguousarray(p)),
                                                               fp(np.ascontiguousarray(s2)), fp(got)))
    assert ((got[:, 9] > 0) == (want[:, 9] > 0)).all(), "connection mask differs"
    assert (bits(got) == bits(want)).all()
