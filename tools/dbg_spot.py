import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, torch
import lensrt as mi
from oracle_lib import *
from test_lens_adjoint import spot_case, COLS
from test_gpu_spot import make_sensor
name, angle, rfilter, full = "exp2a", 0.3, "gaussian", 4096
c = spot_case(name, angle, n_pix=32, full=full, spp=4, distance=6.0)
s = make_sensor(mi, c, rfilter)
rows = c["rows"]
sel = [(i, col) for i in range(len(rows)) for col in COLS[int(rows[i, 0])]][:60]
dist = 1000.0; off = 2.0 * float(rows[-1, 2])
src = c["src"] * np.float32(dist / 6.0); src[2] = rows[-1, 1] + dist
for Gname in ["rand", "w_only", "rgb_only", "lin_x", "quad"]:
    G = np.random.default_rng(8).standard_normal((32, 32, 4)).astype(f32)
    if Gname == "w_only": G[..., :3] = 0
    if Gname == "rgb_only": G[..., 3] = 0
    if Gname == "lin_x": G[:] = 0; G[..., 3] = np.arange(32)[None, :]
    if Gname == "quad": G[:] = 0; G[..., 3] = (np.arange(32)[None, :] - 16.0) ** 2 + (np.arange(32)[:, None] - 16.0) ** 2
    _, want = orc_spot_grad(rows, c["f2s"], c["norm"], 1e-2, 1e4, c["film"], src, [1, .7, .4], 9, 4, G, sel, filter_kind=2, origin_offset=off)
    gr = mi.spot_render_backward(s, src, torch.from_numpy(G).cuda(), [1, .7, .4], seed=9, spp=4, origin_offset=off).cpu().numpy()
    got = np.array([gr[i, col] for i, col in sel], np.float64)
    cos = float(got @ want / (np.linalg.norm(got) * np.linalg.norm(want) + 1e-300))
    print(Gname, "cos", cos, "maxrel", np.abs(got - want).max() / np.abs(want).max())
    if Gname in ("rand", "quad"):
        for (i, col), a, b in zip(sel, got, want): print("   ", i, int(rows[i,0]), col, "%.5g %.5g" % (a, b))
st = mi.spot_render(s, src, [1,.7,.4], seed=9, spp=4, origin_offset=off).cpu().numpy()
print("nonzero px", (st[...,3] > 0).sum(), "W sum", st[...,3].sum())
