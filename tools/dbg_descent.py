import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, torch
import lensrt as mi
from oracle_lib import *
from test_lens_adjoint import spot_case, COLS
from test_gpu_spot import make_sensor
for full, npix, pert, step, iters in [(512, 64, 1.03, 2e-3, 25), (1024, 64, 1.02, 1e-3, 30), (2048, 48, 1.0, 5e-4, 20)]:
    c = spot_case("doublet", 1.0, n_pix=npix, full=full, spp=8, distance=6.0)
    s = make_sensor(mi, c, "tent")
    rows0 = c["rows"].copy()
    k = int(np.argmax(rows0[:, 0] == 0))
    base = float(mi.rms_loss(mi.spot_image(torch.tensor(rows0), s, c["src"], seed=1, spp=8)))
    rows0[k, 3] *= pert
    R = torch.tensor(float(rows0[k, 3]), requires_grad=True)
    losses = []
    for it in range(iters):
        r = torch.tensor(rows0).clone(); r[k, 3] = R
        l = mi.rms_loss(mi.spot_image(r, s, c["src"], seed=100 + it, spp=8))
        losses.append(round(float(l.detach()), 3))
        g, = torch.autograd.grad(l, R)
        with torch.no_grad():
            R -= step * abs(float(rows0[k, 3])) * torch.sign(g)
    print(full, npix, pert, "unperturbed loss", base, "losses", losses)
