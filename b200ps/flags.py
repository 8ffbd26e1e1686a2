"""Flag system (N9): a small gflags-style registry over argparse.

The reference mixes three mechanisms — ``tf.app.flags`` (example.py:33-35, lr2.py:17-35),
bash ``shflags`` (run_lr2.sh:9-20) and ``cluster_conf.json`` (lr2.py:325).  Here one
registry serves the Python entry points and the shell launcher passes flags through
unchanged.  Names and defaults of every reference flag are preserved; the
framework adds ``--cluster_conf``, ``--sync``, ``--placement``, ``--backend`` and friends.

    from b200ps import flags
    flags.DEFINE_integer("task_index", 0, "Index of task within the job")
    FLAGS = flags.FLAGS
    flags.parse()          # or flags.run(main) like tf.app.run

Like ``tf.app.flags``, attributes may also be *set* on FLAGS without a definition
(model_export.py:10-11 does ``FLAGS.work_dir = './model/'``).
"""
from __future__ import annotations

import argparse
import sys
from typing import Any, Callable, Dict, List, Optional


class _FlagValues:
    def __init__(self):
        object.__setattr__(self, "_defs", {})      # name -> (type, default, help)
        object.__setattr__(self, "_values", {})
        object.__setattr__(self, "_parsed", False)
        object.__setattr__(self, "_extra", [])

    # -- definition ------------------------------------------------------------
    def _define(self, name: str, typ, default, helptext: str, short: Optional[str] = None):
        if name in self._defs:
            # re-definition with identical type/default is tolerated (modules imported twice)
            old = self._defs[name]
            if old[0] is not typ or old[1] != default:
                raise ValueError(f"flag --{name} already defined with a different type/default")
            return
        self._defs[name] = (typ, default, helptext, short)
        self._values[name] = default

    # -- parsing ---------------------------------------------------------------
    def _parser(self, prog=None) -> argparse.ArgumentParser:
        p = argparse.ArgumentParser(prog=prog, allow_abbrev=False,
                                    formatter_class=argparse.ArgumentDefaultsHelpFormatter)
        for name, (typ, default, helptext, short) in sorted(self._defs.items()):
            names = [f"--{name}"] + ([f"-{short}"] if short else [])
            if typ is bool:
                p.add_argument(*names, dest=name, nargs="?", const=True, default=default,
                               type=_parse_bool, help=helptext)
                p.add_argument(f"--no{name}", dest=name, action="store_false", help=argparse.SUPPRESS)
            else:
                p.add_argument(*names, dest=name, type=typ, default=default, help=helptext)
        return p

    def _parse(self, argv: Optional[List[str]] = None, known_only: bool = True) -> List[str]:
        argv = list(sys.argv[1:] if argv is None else argv)
        ns, extra = self._parser().parse_known_args(argv)
        for k, v in vars(ns).items():
            self._values[k] = v
        object.__setattr__(self, "_parsed", True)
        object.__setattr__(self, "_extra", extra)
        if extra and not known_only:
            raise SystemExit(f"unknown flags: {extra}")
        return extra

    def _reset(self):
        for name, (typ, default, _h, _s) in self._defs.items():
            self._values[name] = default
        object.__setattr__(self, "_parsed", False)

    # -- access ----------------------------------------------------------------
    def __getattr__(self, name: str) -> Any:
        vals = object.__getattribute__(self, "_values")
        if name in vals:
            return vals[name]
        raise AttributeError(f"unknown flag {name!r}")

    def __setattr__(self, name: str, value: Any):
        self._values[name] = value

    def __contains__(self, name):
        return name in self._values

    def as_dict(self) -> Dict[str, Any]:
        return dict(self._values)


def _parse_bool(s):
    if isinstance(s, bool):
        return s
    if s.lower() in ("1", "true", "t", "yes", "y"):
        return True
    if s.lower() in ("0", "false", "f", "no", "n"):
        return False
    raise argparse.ArgumentTypeError(f"not a boolean: {s!r}")


FLAGS = _FlagValues()


def DEFINE_string(name, default, helptext="", short=None):
    FLAGS._define(name, str, default, helptext, short)


def DEFINE_integer(name, default, helptext="", short=None):
    FLAGS._define(name, int, default, helptext, short)


def DEFINE_float(name, default, helptext="", short=None):
    FLAGS._define(name, float, default, helptext, short)


def DEFINE_boolean(name, default, helptext="", short=None):
    FLAGS._define(name, bool, default, helptext, short)


DEFINE_bool = DEFINE_boolean


def parse(argv: Optional[List[str]] = None, known_only: bool = True) -> List[str]:
    """Parse ``argv`` (default ``sys.argv[1:]``); returns the unparsed remainder."""
    return FLAGS._parse(argv, known_only)


def run(main: Callable, argv: Optional[List[str]] = None):
    """``tf.app.run`` equivalent (lr2.py:479-480): parse flags, then ``main(remaining_argv)``."""
    extra = parse(argv)
    sys.exit(main([sys.argv[0]] + extra))


def define_role_flags(job_name_default: str = "worker"):
    """--job_name / --task_index (example.py:33-34, lr2.py:20-21) + framework-wide additions."""
    DEFINE_string("job_name", job_name_default, "Either 'ps' or 'worker'", short="j")
    DEFINE_integer("task_index", 0, "Index of task within the job", short="i")
    DEFINE_string("cluster_conf", "cluster_conf.json",
                  "JSON cluster spec {\"ps\": [host:port...], \"worker\": [...]} (lr2.py:325)")
    DEFINE_boolean("sync", False, "synchronous replicas (chief + token barrier) instead of async Hogwild")
    DEFINE_integer("replicas_to_aggregate", 0, "gradients averaged per sync step (0 = all workers)")
    DEFINE_string("placement", "flat_shard",
                  "variable placement over ps shards: round_robin_variable (reference parity) | flat_shard")
    DEFINE_string("backend", "auto", "transport: auto | nvlink (GPU peer memory) | gloo (CPU processes)")
    DEFINE_string("logdir", "", "Supervisor logdir: periodic checkpoints + restore on start ('' = off)")
    DEFINE_integer("save_secs", 600, "Supervisor checkpoint period in seconds (with --logdir)")
    DEFINE_integer("seed", 1, "graph-level random seed (example.py:82 uses 1)")
    DEFINE_integer("fault_inject_step", -1, "test hook: the worker named by --fault_rank exits at this step")
    DEFINE_integer("fault_rank", -1, "test hook: task_index of the worker to kill")
    DEFINE_integer("heartbeat_steps", 0, "NVLink backend: check every N local steps that all ranks still answer "
                   "(0 = off); a silent rank raises PeerLost on the survivors before the next launch")
    DEFINE_float("barrier_timeout_secs", 120.0, "give up waiting for peers after this many seconds")
