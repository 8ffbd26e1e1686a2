"""Supervisor (R7, P10): ``tf.train.Supervisor`` — chief initialises (or restores), the other
workers wait until the parameter servers hold initialised variables; optional periodic
checkpoints; ``stop()``.

Reference use: ``sv = Supervisor(is_chief=(task_index == 0), global_step=..., init_op=...)`` then
``with sv.prepare_or_wait_for_session(server.target) as sess`` (example.py:139-145, lr2.py:411-418).
[TF-lib] With a ``logdir`` the chief restores the newest checkpoint instead of running
``init_op`` and saves ``model.ckpt-<global_step>`` every ``save_model_secs`` (600 s); the
reference passes no logdir, so both default to off here too.  Unlike lr2.py:419 the
non-chief workers do NOT re-run the initialiser (that quirk resets a trained model).
"""
from __future__ import annotations

import os
import threading
import time
from typing import Callable, Optional

from ..io import log
from ..io.saver import Saver, latest_checkpoint, store_variables


class Session:
    """What ``prepare_or_wait_for_session`` hands back: a handle on the store (context manager)."""

    def __init__(self, store, supervisor):
        self.store, self.supervisor = store, supervisor

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def close(self):
        pass


class Supervisor:
    def __init__(self, is_chief: bool, store, ctx=None, logdir: Optional[str] = None, saver: Optional[Saver] = None,
                 save_model_secs: float = 600.0, recovery_wait_secs: float = 0.2, ready_timeout_secs: float = 120.0,
                 init_fn: Optional[Callable] = None, checkpoint_basename: str = "model.ckpt"):
        self.is_chief, self.store, self.ctx = is_chief, store, ctx
        self.logdir = logdir or None
        self.saver = saver or (Saver() if self.logdir else None)
        self.save_model_secs = save_model_secs
        self.recovery_wait_secs, self.ready_timeout_secs = recovery_wait_secs, ready_timeout_secs
        self.init_fn = init_fn
        self.checkpoint_basename = checkpoint_basename
        self._stop = threading.Event()
        self._saver_thread: Optional[threading.Thread] = None
        self.restored_from: Optional[str] = None

    # ------------------------------------------------------------------
    def prepare_or_wait_for_session(self, master: str = "", config=None) -> Session:
        gpu_store = hasattr(self.store, "pool")
        if self.is_chief:
            ckpt = latest_checkpoint(self.logdir) if self.logdir else None
            if ckpt:
                log.info(f"chief: restoring from {ckpt}")
                self.store.load_state_dict(Saver.restore(ckpt)) if not gpu_store else None
                self.restored_from = ckpt
            elif not gpu_store:
                log.info("chief: running init_op")
                self.store.initialize()
            if self.init_fn:
                self.init_fn(self.store)
        if gpu_store:
            # every rank fills its own shard deterministically (or loads the checkpoint the chief
            # found); the control-plane barrier is the "variables are initialised" signal
            path = self._broadcast_path(self.restored_from)
            if path:
                self.store.load_state_dict(Saver.restore(path))
                self.restored_from = path
            else:
                self.store.initialize()
            if self.ctx is not None:
                self.ctx.control_barrier(self.ready_timeout_secs)
        elif not self.is_chief:
            t0 = time.time()
            while not self.store.is_ready():
                if time.time() - t0 > self.ready_timeout_secs:
                    raise TimeoutError(f"worker waited {self.ready_timeout_secs}s for the chief to initialise "
                                       f"the parameter servers")
                time.sleep(self.recovery_wait_secs)
            log.info("variables initialised by the chief; session ready")
        if self.is_chief and self.logdir and self.save_model_secs > 0:
            self._saver_thread = threading.Thread(target=self._periodic_save, daemon=True, name="sv-saver")
            self._saver_thread.start()
        if not gpu_store:
            self.store.pull()
        return Session(self.store, self)

    managed_session = prepare_or_wait_for_session

    def _broadcast_path(self, path: Optional[str]) -> Optional[str]:
        if self.ctx is None or self.ctx.group is None:
            return path
        import torch.distributed as dist
        box = [path]
        dist.broadcast_object_list(box, src=0, group=self.ctx.group)
        return box[0]

    # ------------------------------------------------------------------
    def save(self) -> Optional[str]:
        """Chief only: gather every shard and write ``<logdir>/model.ckpt-<global_step>`` atomically."""
        if not (self.is_chief and self.logdir and self.saver):
            return None
        variables = store_variables(self.store)
        step = int(round(float(variables["global_step"])))
        path = self.saver.save(variables, os.path.join(self.logdir, self.checkpoint_basename), step)
        log.info(f"saved checkpoint {path}")
        return path

    def _periodic_save(self):
        while not self._stop.wait(self.save_model_secs):
            try:
                self.save()
            except Exception as e:  # never kill training from the saver thread
                log.error(f"periodic checkpoint failed: {e}")

    def should_stop(self) -> bool:
        return self._stop.is_set()

    def request_stop(self) -> None:
        self._stop.set()

    def stop(self) -> None:
        """``sv.stop()`` (example.py:191): stop the service threads.  (The worker then posts its
        done-token with ``server.post_done()`` so the ps can leave ``join()``.)"""
        self._stop.set()
        if self._saver_thread is not None:
            self._saver_thread.join(timeout=10)
