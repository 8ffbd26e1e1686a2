"""Linear regression + Adam (model_export.py:13-51): the checkpoint/export fixture.

``y_pred = x w + b`` with ``w`` = ``test/weights`` (1,1) ~ N(0,1) and ``b`` = ``test/bias`` (1,)
= 0; ``loss = sum((y - y_pred)^2 / n_samples)``; Adam(0.01); 500 steps of 100 random points of
``y = x + 20 sin(x/10)``, x in [0,100).
"""
from __future__ import annotations

import numpy as np
import torch

from ..optim.adam import Adam


def make_data(n_samples: int = 1000):
    x = np.arange(100, step=0.1)[:n_samples]
    y = x + 20 * np.sin(x / 10)
    return x.reshape(-1, 1).astype(np.float32), y.reshape(-1, 1).astype(np.float32)


class LinearRegression:
    def __init__(self, device="cpu", seed: int = 0, scope: str = "test", n_samples: int = 1000):
        g = torch.Generator().manual_seed(seed)
        self.scope = scope
        self.n_samples = n_samples
        self.device = torch.device(device)
        self.w = torch.randn(1, 1, generator=g).to(self.device)
        self.b = torch.zeros(1, device=self.device)
        self.opt = Adam(learning_rate=0.01)
        self._gw = torch.zeros(1, device=self.device)
        self._gb = torch.zeros(1, device=self.device)
        self._loss = torch.zeros(1, device=self.device)

    def train_step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self.device.type == "cuda":
            from .. import ops
            ops.linreg_fwd_bwd(x, y.reshape(-1), self.w, self.b, loss=self._loss, gw=self._gw, gb=self._gb,
                               inv_n=1.0 / self.n_samples)
            gw, gb = self._gw.view(1, 1), self._gb
        else:
            err = y - (x @ self.w + self.b)
            self._loss = (err ** 2 / self.n_samples).sum().reshape(1)
            gw = (-2.0 / self.n_samples) * (x.t() @ err)
            gb = (-2.0 / self.n_samples) * err.sum().reshape(1)
        self.opt.apply({f"{self.scope}/weights": self.w, f"{self.scope}/bias": self.b},
                       {f"{self.scope}/weights": gw, f"{self.scope}/bias": gb})
        return self._loss

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        return x @ self.w + self.b

    def variables(self):
        """All variables a ``tf.train.Saver()`` would see, incl. Adam slots (TF <= 1.0 prefixes the
        slots with the enclosing variable scope again: ``test/test/weights/Adam``)."""
        out = {f"{self.scope}/weights": self.w, f"{self.scope}/bias": self.b}
        out.update(self.opt.slot_variables(scope_prefix=f"{self.scope}/"))
        return out
