"""Saver (R12, P25): ``tf.train.Saver`` semantics over the tensor-bundle codec.

``save(variables, prefix, global_step)`` writes ``<prefix>-<step>.index`` /
``.data-00000-of-00001`` and maintains the text ``checkpoint`` state file
(``model_checkpoint_path`` + ``all_model_checkpoint_paths``), pruning to ``max_to_keep`` like
TF.  Variable names/shapes follow SURVEY §2.7; ``global_step`` is stored as an fp32 scalar
because the reference declares it without a dtype (example.py:70-72).

The trainers in the reference never actually checkpoint (``--checkpoint`` is declared and
unused, lr2.py:29; Supervisor gets no logdir) — here ``Supervisor(logdir=...)`` does, on the
chief only, atomically (tmp + rename), as TF's Supervisor would.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List, Optional

import numpy as np

from . import tf_bundle

_STATE_FILE = "checkpoint"


def _quote(p: str) -> str:
    return '"' + p.replace("\\", "\\\\").replace('"', '\\"') + '"'


def write_checkpoint_state(directory: str, latest: str, all_paths: List[str]) -> None:
    lines = [f"model_checkpoint_path: {_quote(latest)}"]
    lines += [f"all_model_checkpoint_paths: {_quote(p)}" for p in all_paths]
    tmp = os.path.join(directory, _STATE_FILE + ".tmp")
    with open(tmp, "w") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, os.path.join(directory, _STATE_FILE))


def read_checkpoint_state(directory: str):
    path = os.path.join(directory, _STATE_FILE)
    if not os.path.exists(path):
        return None, []
    latest, allp = None, []
    for line in open(path):
        m = re.match(r'\s*(model_checkpoint_path|all_model_checkpoint_paths):\s*"(.*)"\s*$', line)
        if not m:
            continue
        val = m.group(2).replace('\\"', '"').replace("\\\\", "\\")
        if m.group(1) == "model_checkpoint_path":
            latest = val
        else:
            allp.append(val)
    return latest, allp


def latest_checkpoint(directory: str) -> Optional[str]:
    """Prefix of the newest checkpoint in ``directory`` (``tf.train.latest_checkpoint``)."""
    latest, _ = read_checkpoint_state(directory)
    if latest is None:
        return None
    prefix = latest if os.path.isabs(latest) else os.path.join(directory, latest)
    return prefix if os.path.exists(prefix + ".index") else None


class Saver:
    def __init__(self, max_to_keep: int = 5):
        self.max_to_keep = max_to_keep
        self._kept: List[str] = []

    def save(self, variables: Dict[str, object], save_path: str, global_step: Optional[int] = None) -> str:
        prefix = save_path if global_step is None else f"{save_path}-{int(global_step)}"
        directory = os.path.dirname(os.path.abspath(prefix))
        os.makedirs(directory, exist_ok=True)
        tf_bundle.write_bundle(prefix, variables)
        base = os.path.basename(prefix)
        _, existing = read_checkpoint_state(directory)
        kept = [p for p in (self._kept or existing) if p != base] + [base]
        while self.max_to_keep and len(kept) > self.max_to_keep:
            old = kept.pop(0)
            for suffix in (".index", ".data-00000-of-00001"):
                try:
                    os.remove(os.path.join(directory, old + suffix))
                except OSError:
                    pass
        self._kept = kept
        write_checkpoint_state(directory, base, kept)
        return prefix

    @staticmethod
    def restore(save_path: str) -> Dict[str, np.ndarray]:
        """name -> ndarray for every variable stored under ``save_path`` (checksums verified)."""
        return tf_bundle.read_bundle(save_path)


def store_variables(store) -> Dict[str, np.ndarray]:
    """Checkpoint view of a ParameterStore: every variable gathered from its shard owners."""
    return {k: v.detach().cpu().numpy() for k, v in store.state_dict().items()}
