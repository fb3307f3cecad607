"""GPU parity, part 1: sampler streams and the lens step through the C ABI vs the oracle (bit-exact)."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import *  # noqa

ld = load_lens_designs()
pytestmark = pytest.mark.gpu


def test_sampler_streams_bit_exact(mi):
    O = oracle()
    for seed, idx0, n, nd in [(0, 0, 4096, 8), (7, 123456, 1000, 46), (0xffffffff, 0xfffffff0, 16, 3)]:
        want = np.zeros((n, nd), f32)
        O.orc_sampler_jit_f32(C.c_uint32(seed), C.c_uint32(idx0), n, nd, fp(want))
        got = np.zeros((n, nd), f32)
        mi.capi.check(mi.capi.lib.lrt_sampler_draw_host(C.c_uint32(seed), C.c_uint32(idx0), C.c_int64(n), C.c_int32(nd), fp(got)))
        assert (bits(got) == bits(want)).all()


FILMS = {"default": (FilmSpec(256, 192, [200, 100, 20, 30], 35.0), dict(width=256, height=192, crop=(200, 100, 20, 30), diagonal=35.0)),
         "small": (FilmSpec(128, 128, None, 4.0), dict(width=128, height=128, diagonal=4.0))}


@pytest.mark.parametrize("name", ld.all_designs())
@pytest.mark.parametrize("fname", list(FILMS))
def test_sample_ray_golden(mi, golden, name, fname):
    """CUDA sample_ray == the reference's own code on the committed golden vectors (rays, weights, miss mask)."""
    g = golden["lens"]
    key = f"{name}.{fname}"
    s = mi.RealisticLensCamera(mi.Film(**FILMS[fname][1]), name)
    o, d, maxt, w = s.sample_ray(0.0, 0.0, g[key + ".pos"], g[key + ".ap"])
    got = np.concatenate([o.numpy(), d.numpy(), maxt.numpy()[:, None], w.numpy()], 1)
    want = g[key + ".ray"]
    assert ((got[:, 7] > 0) == (want[:, 7] > 0)).all(), "lens ray-miss mask differs"
    act = want[:, 7] > 0
    assert (bits(got[act]) == bits(want[act])).all()


@pytest.mark.parametrize("name", ["doublet", "gauss", "fisheye", "asph", "exp2a", "exp2c", "hypercentric"])
def test_sample_ray_vs_oracle_1m(mi, name):
    """2^20 PCG32-drawn (u,v,a,b) per design: masks and rays bit-equal to the oracle; with a to_world pose."""
    import torch
    O = oracle()
    n = 1 << 20
    draws = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    mi.capi.check(mi.capi.lib.lrt_sampler_draw(C.c_uint32(11), C.c_uint32(0), C.c_int64(n), C.c_int32(4),
                                               mi.capi.dev_ptr(draws), mi.capi.stream_ptr()))
    small = name in ("asph", "exp2a", "exp2c")
    if small:   # concentrate aperture samples so that a useful fraction is transmitted
        draws[:, 2:4] = 0.5 + (draws[:, 2:4] - 0.5) * 0.2
    th = 0.3
    tw = np.array([[np.cos(th), 0, np.sin(th), 1.5], [0, 1, 0, -2], [-np.sin(th), 0, np.cos(th), 0.25], [0, 0, 0, 1]], f32)
    kw = dict(width=320, height=200, diagonal=8.0 if small else 35.0)
    s = mi.RealisticLensCamera(mi.Film(**kw), name, to_world=tw)
    o, d, maxt, w = s.sample_ray(0.0, 0.0, draws[:, 0:2].contiguous(), draws[:, 2:4].contiguous())
    got = torch.cat([o, d, maxt[:, None], w], 1).cpu().numpy()
    h = orc_camera(ld.design(name, {}), FilmSpec(320, 200, None, kw["diagonal"]), to_world=tw)
    dn = draws.cpu().numpy()
    want = cam_sample_ray(O, "orc", h, dn[:, 0:2], dn[:, 2:4])
    assert ((got[:, 7] > 0) == (want[:, 7] > 0)).all(), "lens ray-miss mask differs"
    act = want[:, 7] > 0
    assert act.sum() > 100
    assert (bits(got[act]) == bits(want[act])).all()


# ------------------------------------------------------------------------------------------------ exit-pupil sampling (L7)
PUPIL_FILMS = {"doublet": dict(width=256, height=256, diagonal=35.0), "gauss": dict(width=256, height=192, crop=(200, 100, 20, 30), diagonal=35.0),
               "asph": dict(width=128, height=128, diagonal=20.0), "exp1a": dict(width=256, height=256, diagonal=35.0)}


def test_radical_inverse_host_entry(mi):
    g = np.load(os.path.join(ROOT, "tests", "golden", "exit_pupil.npz"))
    for b in range(3):
        got = np.array([mi.capi.lib.lrt_radical_inverse(C.c_int32(b), C.c_uint64(int(i))) for i in g["ri.index"][:1500]], f32)
        assert (bits(got) == bits(g[f"ri.base{b}"][:1500])).all()
    assert mi.capi.lib.lrt_radical_inverse(C.c_int32(3), C.c_uint64(1)) == -1.0


@pytest.mark.parametrize("name", list(PUPIL_FILMS))
def test_exit_pupil_lut_and_sample_ray_golden(mi, name):
    """The LUT kernel (64 x 2^20 Halton rays) and sample_ray in pupil mode == the reference's own code, bit for bit."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "exit_pupil.npz"))
    s = mi.RealisticLensCamera(mi.Film(**PUPIL_FILMS[name]), name, sample_exit_pupil=True, shutter_open=0.0, shutter_close=1.0)
    assert s.shutter_open_time() == 1.0
    assert (bits(s.exit_pupil_bounds()) == bits(g[name + ".bounds"])).all()
    o, d, maxt, w = s.sample_ray(0.0, 0.0, g[name + ".pos"], g[name + ".ap"])
    got = np.concatenate([o.numpy(), d.numpy(), maxt.numpy()[:, None], w.numpy()], 1)
    want = g[name + ".ray"]
    assert ((got[:, 7] > 0) == (want[:, 7] > 0)).all(), "lens ray-miss mask differs"
    act = want[:, 7] > 0
    assert (bits(got[act]) == bits(want[act])).all()
    # switching the mode off restores rear-element sampling; a new interface table rebuilds the LUT
    rows = s.interfaces()
    s.set_interfaces(rows)
    assert (bits(s.exit_pupil_bounds()) == bits(g[name + ".bounds"])).all()
    s.set_exit_pupil(False)
    with pytest.raises(mi.capi.LensrtError):
        s.exit_pupil_bounds()


def test_exit_pupil_default_shutter_is_black(mi):
    """shutter_open_time() is 0 unless the scene sets shutter_open/close: the reference's pupil-mode weight is then 0."""
    s = mi.RealisticLensCamera(mi.Film(width=64, height=64), "doublet", sample_exit_pupil=True)
    rng = np.random.default_rng(3)
    o, d, maxt, w = s.sample_ray(0.0, 0.0, rng.random((256, 2), dtype=f32), rng.random((256, 2), dtype=f32))
    assert float(w.abs().max()) == 0.0
