"""mi.ad: the gradient-descent optimizers of src/python/python/ad/optimizers.py (SGD :109-196, Adam :198-330) over torch
tensors. Usage as in the reference:

    opt = mi.ad.Adam(lr=0.05); opt[key] = params[key]; params.update(opt)
    img = mi.render(scene, params, spp=4); loss = ...; loss.backward(); opt.step(); params.update(opt)
"""
from __future__ import annotations

import torch


class Optimizer:
    def __init__(self, lr, params=None):
        self.lr, self.variables, self.state, self.lr_v = float(lr), {}, {}, {}
        if params is not None:
            for k in params:
                self[k] = params[k]

    def __contains__(self, key):
        return key in self.variables

    def __getitem__(self, key):
        return self.variables[key]

    def __setitem__(self, key, value):
        """Register / overwrite a variable: a fresh leaf tensor with gradient tracking (optimizers.py:41-64)."""
        v = torch.as_tensor(value).detach().clone().to(torch.float32)
        v.requires_grad_(True)
        self.variables[key] = v
        self.reset(key)

    def __iter__(self):
        return iter(self.variables)

    def keys(self):
        return self.variables.keys()

    def items(self):
        return self.variables.items()

    def set_learning_rate(self, lr):
        """A float (all variables) or {key: rate} (optimizers.py:79-100)."""
        if isinstance(lr, dict):
            self.lr_v.update({k: float(v) for k, v in lr.items()})
        else:
            self.lr = float(lr)

    def _lr(self, key):
        return self.lr_v.get(key, self.lr)

    def reset(self, key):
        self.state[key] = None


class SGD(Optimizer):
    """Stochastic gradient descent with optional momentum (optimizers.py:109-196)."""

    def __init__(self, lr, momentum=0.0, mask_updates=False, params=None):
        assert 0 <= momentum < 1 and lr > 0
        self.momentum, self.mask_updates = float(momentum), mask_updates
        super().__init__(lr, params)

    def step(self):
        with torch.no_grad():
            for k, p in self.variables.items():
                if p.grad is None:
                    continue
                g = p.grad
                if self.momentum == 0:
                    p -= self._lr(k) * g
                else:
                    v = self.state[k] if self.state[k] is not None else torch.zeros_like(p)
                    v = self.momentum * v + g
                    self.state[k] = v
                    p -= self._lr(k) * v
                p.grad = None


class Adam(Optimizer):
    """Adam (optimizers.py:198-330) incl. the bias-corrected step size lr * sqrt(1 - b2^t) / (1 - b1^t)."""

    def __init__(self, lr, beta_1=0.9, beta_2=0.999, epsilon=1e-8, mask_updates=False, uniform=False, params=None):
        assert 0 <= beta_1 < 1 and 0 <= beta_2 < 1 and lr > 0 and epsilon > 0
        self.beta_1, self.beta_2, self.epsilon, self.mask_updates, self.uniform = beta_1, beta_2, epsilon, mask_updates, uniform
        super().__init__(lr, params)

    def step(self):
        with torch.no_grad():
            for k, p in self.variables.items():
                if p.grad is None:
                    continue
                g = p.grad
                t, m, v = self.state[k] if self.state[k] is not None else (0, torch.zeros_like(p), torch.zeros_like(p))
                t += 1
                m = self.beta_1 * m + (1 - self.beta_1) * g
                v = self.beta_2 * v + (1 - self.beta_2) * g * g
                self.state[k] = (t, m, v)
                lr_t = self._lr(k) * (1 - self.beta_2 ** t) ** 0.5 / (1 - self.beta_1 ** t)
                vt = v.max() if self.uniform else v
                p -= lr_t * m / (vt.sqrt() + self.epsilon)
                p.grad = None
