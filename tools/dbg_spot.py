import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, torch
import lensrt as mi
from oracle_lib import *
from test_lens_adjoint import spot_case, COLS
from test_gpu_spot import make_sensor, CASES
for name, angle, rfilter, full, dist in CASES:
    c = spot_case(name, angle, n_pix=32, full=full, spp=4, distance=dist)
    s = make_sensor(mi, c, rfilter)
    rows = c["rows"]
    sel = [(i, col) for i in range(len(rows)) for col in COLS[int(rows[i, 0])]][:60]
    fk = {"tent": 1, "gaussian": 2}[rfilter]
    for Gname in ["rand", "w_only", "rgb_const_x"]:
        rng = np.random.default_rng(5)
        G = rng.standard_normal((32, 32, 4)).astype(f32)
        if Gname == "w_only": G[..., :3] = 0
        if Gname == "rgb_const_x":
            G[:] = 0; G[..., 0] = np.arange(32)[None, :]
        phi, want = orc_spot_grad(rows, c["f2s"], c["norm"], 1e-2, 1e4, c["film"], c["src"], [1, .7, .4], 9, 4, G, sel, filter_kind=fk)
        gr = mi.spot_render_backward(s, c["src"], torch.from_numpy(G).cuda(), [1, .7, .4], seed=9, spp=4).cpu().numpy()
        got = np.array([gr[i, col] for i, col in sel], np.float64)
        cos = float(got @ want / (np.linalg.norm(got) * np.linalg.norm(want) + 1e-300))
        print(name, angle, rfilter, Gname, "cos", cos, "maxrel", np.abs(got - want).max() / np.abs(want).max())
        if cos < 0.9999 and len(sel) < 20:
            for (i, col), a, b in zip(sel, got, want): print("   ", i, col, a, b)
