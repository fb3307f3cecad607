"""Film step. CPU: the oracle's ImageBlock::put against the NumPy brute-force model of the reference's own
test (src/render/tests/test_imageblock.py:37-121, restated here), filters against closed forms, host filter
code of the product against the oracle. GPU: lrt_film_put / develop against the oracle."""
import ctypes as C

import os

import numpy as np
import pytest

from oracle_lib import *  # noqa

KIND = {"box": 0, "tent": 1, "gaussian": 2}


def orc_eval(kind, x, param=0.0):
    O = oracle()
    x = np.ascontiguousarray(x, f32)
    out = np.zeros_like(x)
    O.orc_filter_eval(C.c_int(kind), c_f(param), C.c_int(x.size), fp(x), fp(out))
    return out


def orc_put(kind, size, offset, border, channels, coalesce, normalize, pos, values, param=0.0):
    O = oracle()
    O.orc_filter_radius.restype = c_f
    t = np.zeros((size[1] + 2 * border, size[0] + 2 * border, channels), np.float64)
    pos = np.ascontiguousarray(pos, f32).reshape(-1, 2); values = np.ascontiguousarray(values, f32).reshape(-1, channels)
    O.orc_block_put(C.c_int(kind), c_f(param), C.c_int(size[0]), C.c_int(size[1]), C.c_int(offset[0]), C.c_int(offset[1]),
                    C.c_int(border), C.c_int(channels), C.c_int(int(coalesce)), C.c_int(int(normalize)),
                    C.c_int64(len(pos)), fp(pos), fp(values), t.ctypes.data_as(C.POINTER(C.c_double)))
    return t


def border_size(kind):
    r = {0: 0.5, 1: 1.0, 2: 2.0}[kind]
    return int(np.ceil(r - 0.5 - 2 * 8.94069671630859375e-05))


def test_filter_closed_forms():
    x = np.linspace(-2.5, 2.5, 1001).astype(f32)
    assert np.allclose(orc_eval(1, x), np.maximum(0, 1 - np.abs(x)), atol=1e-7)
    g = orc_eval(2, x)
    want = np.maximum(np.exp(-x.astype(np.float64) ** 2 / (2 * 0.25)) - np.exp(-4.0 / (2 * 0.25)), 0)
    assert np.abs(g - want).max() < 1e-3 and g[np.abs(x) >= 2].max() < 1e-5
    assert orc_eval(2, np.array([2.0, -2.0], f32)).max() < 4e-6   # shifted so that f(radius) ~ 0 (float32 cancellation)
    assert (orc_eval(0, np.array([-0.5, 0.4999, 0.5], f32)) == [1, 1, 0]).all()


@pytest.mark.parametrize("filter_name", ["gaussian", "tent", "box"])
@pytest.mark.parametrize("border", [False, True])
@pytest.mark.parametrize("offset", [[0, 0], [1, 2]])
@pytest.mark.parametrize("normalize", [False, True])
@pytest.mark.parametrize("coalesce", [False, True])
def test_put_vs_bruteforce_model(filter_name, border, offset, normalize, coalesce):
    """test_imageblock.py::test02_put, JIT flavour: block.put(pos, [1]) == eval(-p_x) * eval(-p_y) on the grid."""
    kind = KIND[filter_name]
    b = border_size(kind) if border else 0
    for j in range(5):
        for i in range(5):
            pos = np.array([3.3 + 0.25 * i, 3 + 0.25 * j], f32)
            got = orc_put(kind, (6, 6), offset, b, 1, coalesce, normalize, pos, [1.0])[..., 0]
            n = 6 + 2 * b
            shift = 0.5 - pos.astype(np.float64) - b
            gx, gy = np.arange(n) + shift[0], np.arange(n) + shift[1]
            ref_norm = np.outer(orc_eval(kind, (-gy).astype(f32)), orc_eval(kind, (-gx).astype(f32)))
            shift = shift + np.array(offset)
            gx, gy = np.arange(n) + shift[0], np.arange(n) + shift[1]
            ref = np.outer(orc_eval(kind, (-gy).astype(f32)), orc_eval(kind, (-gx).astype(f32))).astype(np.float64)
            if normalize and kind != 0:
                # the model of the reference's test normalises by the un-offset, un-clipped filter sum
                wx = orc_eval(kind, (np.arange(-8, 9) + 0.5 - (pos[0] - np.floor(pos[0]))).astype(f32)).sum()
                wy = orc_eval(kind, (np.arange(-8, 9) + 0.5 - (pos[1] - np.floor(pos[1]))).astype(f32)).sum()
                if wx * wy > 0:
                    ref = ref / (wx * wy)
            # the float32 Remez polynomial carries ~1e-5 of cancellation noise near the filter edge
            assert np.allclose(got, ref, atol=3e-5), (pos, got, ref)


def test_put_boundary_box():
    # test_imageblock.py::test03_put_boundary: samples outside the block are dropped
    t = orc_put(0, (3, 3), (0, 0), 0, 1, True, False, [[-0.1, 1], [3.0, 1], [2.99, 2.99], [1, -0.01]], [1, 1, 1, 1])
    assert t.sum() == 1 and t[2, 2, 0] == 1


def test_product_filter_host_matches_oracle(mi):
    x = np.linspace(-2.2, 2.2, 4001).astype(f32)
    for name, kind in KIND.items():
        f = mi.ReconstructionFilter(name)
        assert np.abs(f.eval(x) - orc_eval(kind, x)).max() < 2e-7
        assert f.border_size() == border_size(kind)


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("filter_name", ["gaussian", "tent", "box"])
@pytest.mark.parametrize("border", [False, True])
@pytest.mark.parametrize("normalize", [False, True])
@pytest.mark.parametrize("coalesce", [False, True])
def test_gpu_put_vs_oracle(mi, filter_name, border, normalize, coalesce):
    import torch
    rng = np.random.default_rng(5)
    n, size, offset, ch = 20000, (37, 23), (3, 5), 4
    pos = (rng.random((n, 2)) * [size[0] + 6, size[1] + 6] + [offset[0] - 3, offset[1] - 3]).astype(f32)
    pos[:4] = [[offset[0], offset[1]], [offset[0] + size[0], offset[1] + size[1]], [offset[0] + 0.5, offset[1] + 0.5], [offset[0] - 2, offset[1] + 1]]
    vals = rng.random((n, ch)).astype(f32)
    blk = mi.ImageBlock(size, offset, ch, mi.ReconstructionFilter(filter_name), border=border, normalize=normalize, coalesce=coalesce)
    blk.put(pos, vals)
    got = blk.tensor().cpu().numpy().astype(np.float64)
    want = orc_put(KIND[filter_name], size, offset, blk.border_size(), ch, coalesce, normalize, pos, vals)
    assert got.shape == want.shape
    err = np.abs(got - want).max() / max(1.0, np.abs(want).max())
    assert err < 2e-6, err


@pytest.mark.gpu
def test_gpu_put_reference_test_points(mi):
    """The 25 points of test_imageblock.py::test02_put through the CUDA kernel (gaussian, all modes)."""
    for coalesce in (False, True):
        for j in range(5):
            for i in range(5):
                pos = np.array([[3.3 + 0.25 * i, 3 + 0.25 * j]], f32)
                blk = mi.ImageBlock((6, 6), (1, 2), 1, mi.ReconstructionFilter("gaussian"), border=True, coalesce=coalesce)
                blk.put(pos, [1.0])
                want = orc_put(2, (6, 6), (1, 2), blk.border_size(), 1, coalesce, False, pos, [1.0])
                assert np.allclose(blk.tensor().cpu().numpy(), want, atol=1e-6)


@pytest.mark.gpu
def test_gpu_develop(mi):
    import torch
    rng = np.random.default_rng(6)
    st = rng.random((19, 31, 4)).astype(f32); st[3, 4, 3] = 0; st[0, 0] = 0
    got = mi.develop_rgb(torch.from_numpy(st).cuda()).cpu().numpy()
    want = np.zeros((19 * 31, 3), f32)
    oracle().orc_develop_rgb(C.c_int64(19 * 31), st.astype(np.float64).ctypes.data_as(C.POINTER(C.c_double)), fp(want))
    assert (bits(got.reshape(-1, 3)) == bits(want)).all()


def test_exr_and_pfm_round_trip(mi, tmp_path):
    """HDRFilm::write / mi.util.write_bitmap stand-in: uncompressed scanline OpenEXR and PFM, read back bit for bit."""
    rng = np.random.default_rng(5)
    for c in (1, 3, 4):
        img = (rng.random((7, 13, c), dtype=np.float32) * 100 - 10).astype(np.float32)
        p = tmp_path / f"img{c}.exr"
        mi.util.write_bitmap(p, img)
        back = mi.imageio.read_exr(p)
        assert back.shape == img.shape and (back.view(np.uint32) == img.view(np.uint32)).all()
        # structure: magic, version 2, scanline offsets point at blocks that start with their own y
        raw = open(p, "rb").read()
        assert raw[:8] == bytes.fromhex("762f3101") + (2).to_bytes(4, "little")
    half = tmp_path / "half.exr"
    img = rng.random((5, 6, 3), dtype=np.float32)
    mi.imageio.write_exr(half, img, half=True)
    assert np.abs(mi.imageio.read_exr(half) - img).max() < 1e-3
    pf = tmp_path / "img.pfm"
    mi.util.write_bitmap(pf, img)
    assert (mi.imageio.read_pfm(pf) == img).all()
    with pytest.raises(ValueError):
        mi.util.write_bitmap(tmp_path / "img.png", img)
    # an independent OpenEXR implementation (OpenCV's bundled one), when present, reads the same pixels
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    try:
        import cv2
    except ImportError:
        return
    p3 = tmp_path / "cv.exr"
    mi.util.write_bitmap(p3, img)
    a = cv2.imread(str(p3), cv2.IMREAD_UNCHANGED)
    if a is not None:   # builds without the EXR codec return None
        assert a.dtype == np.float32 and (a[..., ::-1] == img).all()
