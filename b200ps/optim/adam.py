"""Adam (R9, K11) with TF's ``ApplyAdam`` update and slot layout (model_export.py:38).

Slots per variable ``m`` (``<var>/Adam``) and ``v`` (``<var>/Adam_1``) plus the shared
``beta1_power`` / ``beta2_power`` scalars, so a Saver dump has the reference's variable set.
"""
from __future__ import annotations

from typing import Dict

import torch


class Adam:
    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8):
        self.lr, self.beta1, self.beta2, self.eps = learning_rate, beta1, beta2, epsilon
        self.m: Dict[str, torch.Tensor] = {}
        self.v: Dict[str, torch.Tensor] = {}
        self.beta_pow = None
        self.t = 0

    def _slots(self, name, p):
        if name not in self.m:
            self.m[name] = torch.zeros_like(p)
            self.v[name] = torch.zeros_like(p)
        if self.beta_pow is None:
            self.beta_pow = torch.tensor([self.beta1, self.beta2], dtype=torch.float32, device=p.device)
        return self.m[name], self.v[name]

    def apply(self, params: Dict[str, torch.Tensor], grads: Dict[str, torch.Tensor]) -> None:
        from .. import ops
        self.t += 1
        for name, p in params.items():
            m, v = self._slots(name, p)
            g = grads[name]
            if p.is_cuda:
                ops.adam_apply(p.view(-1), g.view(-1), m.view(-1), v.view(-1), self.beta_pow, self.lr,
                               self.beta1, self.beta2, self.eps)
            else:
                b1p, b2p = float(self.beta_pow[0]), float(self.beta_pow[1])
                lr_t = self.lr * (1 - b2p) ** 0.5 / (1 - b1p)
                m.mul_(self.beta1).add_(g, alpha=1 - self.beta1)
                v.mul_(self.beta2).addcmul_(g, g, value=1 - self.beta2)
                p.sub_(lr_t * m / (v.sqrt() + self.eps))
        if self.beta_pow.is_cuda:
            ops.adam_advance(self.beta_pow, self.beta1, self.beta2)
        else:
            self.beta_pow[0] *= self.beta1
            self.beta_pow[1] *= self.beta2

    def slot_variables(self, scope_prefix: str = "") -> Dict[str, torch.Tensor]:
        """Checkpoint view: ``<prefix><var>/Adam``, ``<prefix><var>/Adam_1``, beta powers."""
        out = {}
        for name in self.m:
            out[f"{scope_prefix}{name}/Adam"] = self.m[name]
            out[f"{scope_prefix}{name}/Adam_1"] = self.v[name]
        if self.beta_pow is not None:
            out[f"{scope_prefix}beta1_power"] = self.beta_pow[0].reshape(())
            out[f"{scope_prefix}beta2_power"] = self.beta_pow[1].reshape(())
        return out
