"""Classification accuracy (P12) — ``reduce_mean(cast(equal(argmax(y,1), argmax(y_,1))))`` of
example.py:125-128, evaluated on a training batch every step and on the whole 10 000-image test set
in one call at the end (example.py:187).

On the GPU path the count of correct predictions is produced by the step kernels themselves (the fused
head, ``mlp_tail`` / ``mlp_fused_step``, write ``correct`` next to the loss), so a running accuracy costs
nothing extra; this module is the host-side accumulator around those counters plus the plain-tensor
form used by the CPU (gloo) path and the tests.
"""
from __future__ import annotations

import torch


def accuracy(logits: torch.Tensor, labels: torch.Tensor) -> float:
    """Fraction of rows whose arg-max class equals the label (labels: class ids, or one-hot rows like the
    reference's ``y_`` placeholder).  Ties resolve to the lowest class index, as in the kernels."""
    if labels.dim() == 2:
        labels = labels.argmax(1)
    return float((logits.argmax(1) == labels.to(logits.device).long()).float().mean().item())


class StreamingAccuracy:
    """Running accuracy over many batches: ``update(correct, n)`` with the kernels' counters (device
    int32 tensors are read only in ``result()``), or ``update_from_logits``."""

    def __init__(self):
        self._correct, self._seen, self._pending = 0, 0, []

    def update(self, correct, n: int) -> None:
        if isinstance(correct, torch.Tensor):
            self._pending.append(correct.detach().clone())       # no synchronisation here
        else:
            self._correct += int(correct)
        self._seen += int(n)

    def update_from_logits(self, logits: torch.Tensor, labels: torch.Tensor) -> None:
        if labels.dim() == 2:
            labels = labels.argmax(1)
        self.update((logits.argmax(1) == labels.to(logits.device).long()).sum(), labels.numel())

    def result(self) -> float:
        if self._pending:
            self._correct += int(torch.stack([c.reshape(()).long() for c in self._pending]).sum().item())
            self._pending = []
        return self._correct / max(1, self._seen)

    def reset(self) -> None:
        self._correct, self._seen, self._pending = 0, 0, []
