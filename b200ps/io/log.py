"""Levelled logging with the reference's ``[time] [job:task] msg`` prefix (lr2.py:38-48).

``tf.logging.set_verbosity(INFO)`` (lr2.py:15) maps to ``set_verbosity``.
"""
from __future__ import annotations

import logging
import sys
import time

DEBUG, INFO, WARN, ERROR = logging.DEBUG, logging.INFO, logging.WARNING, logging.ERROR

_logger = logging.getLogger("b200ps")
if not _logger.handlers:
    _h = logging.StreamHandler(sys.stderr)
    _h.setFormatter(logging.Formatter("%(levelname)s:b200ps:%(message)s"))
    _logger.addHandler(_h)
    _logger.setLevel(logging.INFO)
    _logger.propagate = False

_role = {"job_name": "worker", "task_index": 0}


def set_role(job_name: str, task_index: int) -> None:
    _role["job_name"], _role["task_index"] = job_name, int(task_index)


def set_verbosity(level: int) -> None:
    _logger.setLevel(level)


def format_message(msg, now=None) -> str:
    tm = time.strftime("%Y-%m-%d %H:%M:%S", time.localtime(time.time() if now is None else now))
    return " [%s] [%s:%d] %s" % (tm, _role["job_name"], _role["task_index"], msg)


def debug(msg) -> None:
    _logger.debug(format_message(msg))


def info(msg) -> None:
    _logger.info(format_message(msg))


def warn(msg) -> None:
    _logger.warning(format_message(msg))


def error(msg) -> None:
    _logger.error(format_message(msg))
