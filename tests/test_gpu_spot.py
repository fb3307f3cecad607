"""GPU parity, lens-design path: sample_direction (bit-exact vs the reference's vectors), light-traced spot images
vs the float32 oracle, and the CUDA adjoint (hand-derived reverse sweep) vs the oracle's float64 dual-number
gradient — tolerances: image rel-RMSE <= 1e-4 (float32 atomics vs float64 sums), valid-sample counts exact,
gradient cosine similarity >= 0.9999 and per-entry error <= 2% of the largest entry."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import *  # noqa
from test_lens_adjoint import spot_case, COLS

ld = load_lens_designs()
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ld.all_designs())
def test_sample_direction_golden(mi, name):
    g = np.load(os.path.join(ROOT, "tests", "golden", "sample_direction.npz"))
    s = mi.RealisticLensCamera(mi.Film(256, 192, crop=(200, 100, 20, 30), diagonal=35.0), name)
    p, s2, want = g[name + ".p"], g[name + ".s"], g[name + ".out"]
    got = np.zeros_like(want)
    mi.capi.check(mi.capi.lib.lrt_sensor_sample_direction_host(s.handle, C.c_int64(len(p)), fp(np.ascontiguousarray(p)),
                                                               fp(np.ascontiguousarray(s2)), fp(got)))
    assert ((got[:, 9] > 0) == (want[:, 9] > 0)).all(), "connection mask differs"
    assert (bits(got) == bits(want)).all()


def make_sensor(mi, c, rfilter="tent"):
    cw, ch, cx, cy = (int(v) for v in c["film"].crop)
    film = mi.Film(c["film"].width, c["film"].height, crop=(cw, ch, cx, cy), diagonal=35.0, rfilter=rfilter)
    s = mi.RealisticLensCamera(film, c["name"], **c["props"])
    assert (bits(s.interfaces()) == bits(c["rows"])).all()
    return s


# (design, field angle in degrees, filter, full-film resolution that sets the pixel pitch, source distance in metres)
# Sources sit at the focus distance (6 m), not at "infinity": sample_direction traces from the reference point in
# float32 (as the reference does), and from 1 km away the cancellation in o + t d leaves ~0.3 px of position noise,
# which would make a per-sample comparison of float32 and float64 tap sets meaningless.
CASES = [("doublet", 0.0, "tent", 2048, 6.0), ("doublet", 2.0, "tent", 2048, 6.0), ("doublet", 1.0, "gaussian", 2048, 6.0),
         ("gauss", 3.0, "tent", 512, 6.0), ("tessar", 2.0, "tent", 1024, 6.0), ("asph", 2.0, "tent", 8192, 0.1),
         ("exp2a", 0.3, "gaussian", 4096, 6.0)]


@pytest.mark.parametrize("name,angle,rfilter,full,dist", CASES)
def test_spot_image_matches_oracle(mi, name, angle, rfilter, full, dist):
    import torch
    c = spot_case(name, angle, n_pix=32, full=full, spp=16, distance=dist)
    s = make_sensor(mi, c, rfilter)
    inten = [1.0, 0.7, 0.4]
    st = torch.zeros((32, 32, 4), device="cuda")
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    mi.spot_render(s, c["src"], inten, seed=11, spp=16, storage=st, stats=stats)
    want, valid = orc_spot_render(c["h"], c["film"], c["src"], inten, 11, 16, filter_kind={"tent": 1, "gaussian": 2}[rfilter])
    stats = stats.cpu().numpy()
    assert stats[0] == 32 * 32 * 16 and stats[1] == valid and valid > 100
    assert rel_rmse(st.cpu().numpy(), want) <= 1e-4
    # sample-index sharding replays the same streams: 3 shards sum to the unsharded image
    parts = sum(mi.spot_render(s, c["src"], inten, seed=11, spp=16, shard=(k, 3)) for k in range(3))
    assert rel_rmse(parts.cpu().numpy(), want) <= 1e-4


@pytest.mark.parametrize("name,angle,rfilter,full,dist", CASES)
def test_spot_adjoint_matches_dual_number_gradient(mi, name, angle, rfilter, full, dist):
    import torch
    c = spot_case(name, angle, n_pix=32, full=full, spp=4, distance=dist)
    s = make_sensor(mi, c, rfilter)
    rng = np.random.default_rng(5)
    G = rng.standard_normal((32, 32, 4)).astype(f32)
    rows = c["rows"]
    sel = [(i, col) for i in range(len(rows)) for col in COLS[int(rows[i, 0])]][:60]
    inten = [1.0, 0.7, 0.4]
    fk = {"tent": 1, "gaussian": 2}[rfilter]
    phi, want = orc_spot_grad(rows, c["f2s"], c["norm"], 1e-2, 1e4, c["film"], c["src"], inten, 9, 4, G, sel, filter_kind=fk)
    gr = mi.spot_render_backward(s, c["src"], torch.from_numpy(G).cuda(), inten, seed=9, spp=4).cpu().numpy()
    got = np.array([gr[i, col] for i, col in sel], np.float64)
    cos = float(got @ want / (np.linalg.norm(got) * np.linalg.norm(want)))
    assert cos >= 0.9999, (cos, got, want)
    # per-entry bound: a handful of boundary samples are vignetted in float32 but not in float64 (or vice versa)
    assert np.abs(got - want).max() <= 2e-2 * np.abs(want).max(), (got, want)
    # entries outside `sel` that the table does not use stay zero (kind, semi-diameter, r_scale, pad)
    assert np.all(gr[:, [0, 2, 5, 15]] == 0)
    # sharded adjoint = unsharded adjoint
    parts = sum(mi.spot_render_backward(s, c["src"], torch.from_numpy(G).cuda(), inten, seed=9, spp=4, shard=(k, 2)) for k in range(2))
    assert np.abs(parts.cpu().numpy() - gr).max() <= 1e-3 * np.abs(gr).max()


def test_rms_loss_gradient_and_descent(mi):
    """End to end: rows -> spot image -> losses.py rms_loss; autograd through lrt_spot_render_backward agrees with the
    checker on G = d(loss)/d(image), and sign-gradient descent on one curvature radius steadily shrinks the spot (measured: 0.786 -> 0.678 px^2 in 25 steps)."""
    import torch
    c = spot_case("doublet", 1.0, n_pix=64, full=512, spp=8, distance=6.0)
    s = make_sensor(mi, c, "tent")
    rows0 = c["rows"].copy()
    k = int(np.argmax(rows0[:, 0] == 0))   # first spherical interface from the rear
    rows0[k, 3] *= 1.03    # perturb its radius: defocus
    rows = torch.tensor(rows0, requires_grad=True)
    img = mi.spot_image(rows, s, c["src"], seed=3, spp=8)
    loss = mi.rms_loss(img)
    loss.backward()
    # checker with the same G
    img2 = img.detach().clone().requires_grad_(True)
    mi.rms_loss(img2).backward()
    G = np.zeros((64, 64, 4), f32); G[..., :3] = img2.grad.cpu().numpy()
    sel = [(i, col) for i in range(len(rows0)) for col in COLS[int(rows0[i, 0])]]
    _, want = orc_spot_grad(rows0, c["f2s"], c["norm"], 1e-2, 1e4, c["film"], c["src"], [1, 1, 1], 3, 8, G, sel)
    got = np.array([rows.grad[i, col].item() for i, col in sel])
    cos = float(got @ want / (np.linalg.norm(got) * np.linalg.norm(want)))
    assert cos >= 0.999, (cos, got, want)
    # descent on one scalar (rear radius) with a normalised step
    R = torch.tensor(float(rows0[k, 3]), requires_grad=True)
    losses = []
    for it in range(25):
        r = torch.tensor(rows0).clone()
        r[k, 3] = R
        l = mi.rms_loss(mi.spot_image(r, s, c["src"], seed=100 + it, spp=8))
        losses.append(float(l.detach()))
        g, = torch.autograd.grad(l, R)
        with torch.no_grad():
            R -= 2e-3 * abs(float(rows0[k, 3])) * torch.sign(g)
    assert np.mean(losses[-3:]) < 0.9 * np.mean(losses[:3]), losses


def test_design_problem_iterations_reduce_the_spot(mi):
    """C2-shaped loop (cemented doublet as three even aspheres, 30 free variables, three fields): Adam iterations driven by
    the CUDA adjoint reduce the summed RMS spot size; with an EFL target the paraxial focal length is held."""
    import torch
    L = mi.lens_design
    n1, n2 = L.cauchy_ior(1.45608, 0.021087), L.cauchy_ior(1.56895, 0.026305)
    rows = L.asphere_table([24.0, -24.0, -168.0107], [3.0, 2.2529, 46.7471], [8, 8, 8], [n1, n2, 1.000277])
    prob = L.LensDesignProblem(mi, rows, [0.0, 2.5, 5.0], resolution=64, full_res=2048, spp=32, spp_grad=16, lr=2e-4)
    assert len(prob.free) == 30
    losses = [prob.step() for _ in range(30)]
    assert np.isfinite(losses).all()
    assert np.mean(losses[-5:]) < 0.8 * np.mean(losses[:5]), losses
    # EFL-constrained variant: rear vertex / curvature follow from the other 28 variables
    prob = L.LensDesignProblem(mi, rows, [0.0, 2.5], resolution=64, full_res=2048, spp=32, spp_grad=16, lr=2e-4,
                               efl_target=0.05)
    assert len(prob.free) == 28
    for _ in range(10):
        prob.step()
    r = prob.rows().detach()
    efl, bfl = L.efl_bfl(r[:, 1], r[:, 3] / r[:, 2], r[:, 13], r[:, 14])
    assert abs(float(efl) - 0.05) < 1e-6 and abs(float(bfl) - float(r[0, 1])) < 1e-6
    assert np.isfinite(prob.losses).all()


@pytest.mark.parametrize("name,angle,rfilter,full,dist", [("doublet", 2.0, "tent", 2048, 1e3), ("asph", 2.0, "tent", 8192, 1e3),
                                                            ("exp2a", 0.3, "gaussian", 4096, 1e3)])
def test_spot_with_near_origin_from_a_distant_source(mi, name, angle, rfilter, full, dist):
    """origin_offset > 0 (the design loop's setting): rays of a source 1 km away start near the lens, so float32 stays
    well-conditioned — image and adjoint agree with the checkers that apply the same offset, even for aspheres whose
    absolute Newton tolerance cannot be met when tracing from 1 km."""
    import torch
    c = spot_case(name, angle, n_pix=32, full=full, spp=4, distance=6.0 if name != "asph" else 0.1)   # window from a near chief ray
    s = make_sensor(mi, c, rfilter)
    src = c["src"] * np.float32(dist / (6.0 if name != "asph" else 0.1))
    src[2] = c["rows"][-1, 1] + dist
    off = 2.0 * float(c["rows"][-1, 2])
    fk = {"tent": 1, "gaussian": 2}[rfilter]
    inten = [1.0, 0.7, 0.4]
    st = mi.spot_render(s, src, inten, seed=4, spp=4, origin_offset=off)
    want, valid = orc_spot_render(c["h"], c["film"], src, inten, 4, 4, filter_kind=fk, origin_offset=off)
    assert valid > 100 and rel_rmse(st.cpu().numpy(), want) <= 1e-4
    rows = c["rows"]
    sel = [(i, col) for i in range(len(rows)) for col in COLS[int(rows[i, 0])]][:60]
    # a smooth image gradient (the shape an RMS-spot loss produces). With white-noise G the per-sample terms (~1e6) cancel
    # to ~1e3 and the comparison would only measure float32's 2.4e-4 px resolution of splat positions near pixel 2048.
    yy, xx = np.mgrid[0:32, 0:32].astype(f32)
    G = np.repeat((((xx - 16) ** 2 + (yy - 16) ** 2) / 64)[..., None], 4, axis=2).astype(f32)
    _, wantg = orc_spot_grad(rows, c["f2s"], c["norm"], 1e-2, 1e4, c["film"], src, inten, 9, 4, G, sel, filter_kind=fk, origin_offset=off)
    gr = mi.spot_render_backward(s, src, torch.from_numpy(G).cuda(), inten, seed=9, spp=4, origin_offset=off).cpu().numpy()
    got = np.array([gr[i, col] for i, col in sel], np.float64)
    cos = float(got @ wantg / (np.linalg.norm(got) * np.linalg.norm(wantg)))
    assert cos >= 0.9999 and np.abs(got - wantg).max() <= 2e-2 * np.abs(wantg).max(), (cos, got, wantg)


def _sharded_design_worker(rank, world, port, q):
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "tests"))
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    import lensrt as mi
    dist.init_process_group("gloo")              # both ranks share the one GPU of the box: gloo moves the CUDA tensors
    L = mi.lens_design
    rows = _doublet_rows(mi)
    prob = L.LensDesignProblem(mi, rows, [0.0, 2.5, 5.0], resolution=64, full_res=2048, spp=32, spp_grad=16, lr=2e-4, shard=(rank, world))
    prob.opt.zero_grad(set_to_none=True)
    l = prob.loss(prob.rows(), 0); l.backward()
    from mitsuba3_camera_b200.parallel import all_reduce_sum
    all_reduce_sum(prob.theta.grad)
    g0 = prob.theta.grad.detach().cpu().numpy().copy()
    for _ in range(3):
        prob.step()
    q.put((rank, float(l), g0, prob.theta.detach().cpu().numpy()))
    dist.destroy_process_group()


def _doublet_rows(mi):
    L = mi.lens_design
    n1, n2 = L.cauchy_ior(1.45608, 0.021087), L.cauchy_ior(1.56895, 0.026305)
    return L.asphere_table([24.0, -24.0, -168.0107], [3.0, 2.2529, 46.7471], [8, 8, 8], [n1, n2, 1.000277])


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_gpu_sharded_design_equals_single_process(mi):
    """C2-shaped design problem on 2 ranks (sample-index sharding, spot accumulators all-reduced before the loss, table
    gradients after the adjoint) = the single-process loss, gradient and Adam trajectory to float summation order."""
    import socket
    import torch
    import torch.multiprocessing as mp
    L = mi.lens_design
    prob = L.LensDesignProblem(mi, _doublet_rows(mi), [0.0, 2.5, 5.0], resolution=64, full_res=2048, spp=32, spp_grad=16, lr=2e-4)
    assert prob.theta.is_cuda and prob.rows().is_cuda
    prob.opt.zero_grad(set_to_none=True)
    l = prob.loss(prob.rows(), 0); l.backward()
    g1 = prob.theta.grad.detach().cpu().numpy().copy()
    for _ in range(3):
        prob.step()
    th1 = prob.theta.detach().cpu().numpy()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_sharded_design_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted((q.get(timeout=400) for _ in ps), key=lambda t: t[0])
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    for rank, loss, g, th in res:
        assert abs(loss - float(l)) <= 1e-4 * abs(float(l)), (loss, float(l))
        assert np.abs(g - g1).max() <= 1e-4 * np.abs(g1).max(), (np.abs(g - g1).max(), np.abs(g1).max())
        assert np.abs(th - th1).max() <= 1e-5 * max(np.abs(th1).max(), 1.0)


@pytest.mark.gpu
def test_gpu_fast_flavour_agrees_with_strict(mi):
    """lrt_spot_params.fast: fused multiply-adds + hardware rcp / sqrt approximations. Same samples, same algorithm: the spot
    image agrees to 2e-3 rel-RMSE (measured 6e-4: film positions move by ~1e-3 pixel, which the tent weights see), the
    transmitted-sample count to 1e-4, the table gradient to cosine 0.9999."""
    import torch
    import bench
    L = mi.lens_design
    prob = L.LensDesignProblem(mi, bench.triplet_table(mi), [3.75], resolution=128, full_res=4096, spp=64, spp_grad=64)
    s, src = prob.sensors[0], prob.sources[0]
    out = {}
    G = torch.randn((128, 128, 4), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    for fast in (False, True):
        stats = torch.zeros(2, dtype=torch.int64, device="cuda")
        st = L.spot_render(s, src, seed=3, spp=64, origin_offset=prob.origin_offset, stats=stats, fast=fast)
        gr = L.spot_render_backward(s, src, G, seed=3, spp=64, origin_offset=prob.origin_offset, fast=fast)
        out[fast] = (st.cpu().numpy(), gr.cpu().numpy().ravel().astype(np.float64), stats.cpu().numpy())
    a, b = out[False], out[True]
    assert rel_rmse(b[0][..., :3], a[0][..., :3]) < 2e-3
    assert abs(int(a[2][1]) - int(b[2][1])) <= 1e-4 * a[2][1] and a[2][1] > 1e5
    cos = float(a[1] @ b[1] / (np.linalg.norm(a[1]) * np.linalg.norm(b[1])))
    assert cos >= 0.9999, cos


@pytest.mark.gpu
def test_gpu_design_iteration_as_cuda_graph(mi):
    """LensDesignProblem.capture(): the iteration replayed from one CUDA graph draws fresh samples every replay (device-side
    seed counter) and follows the eager trajectory (same seeds, same arithmetic: equal to float atomics' summation order)."""
    import torch
    L = mi.lens_design
    rows = _doublet_rows(mi)
    eager = L.LensDesignProblem(mi, rows, [0.0, 2.5, 5.0], resolution=64, full_res=2048, spp=32, spp_grad=16, lr=2e-4)
    for _ in range(3 + 5):
        eager.step()
    graph = L.LensDesignProblem(mi, rows, [0.0, 2.5, 5.0], resolution=64, full_res=2048, spp=32, spp_grad=16, lr=2e-4)
    graph.capture(warmup=3)               # three warm-up iterations run eagerly and count as iterations 0..2
    losses = []
    for _ in range(5):
        losses.append(float(graph.step_graph()))
    assert len(set(losses)) == 5                                   # fresh samples every replay
    a, b = eager.theta.detach().cpu().numpy(), graph.theta.detach().cpu().numpy()
    assert np.abs(a - b).max() <= 1e-4 * max(np.abs(a).max(), 1.0), np.abs(a - b).max()
    assert np.allclose(losses, eager.losses[3:8], rtol=1e-3)
