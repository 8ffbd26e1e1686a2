"""Native sm_100a kernels (libb200ps.so) and their torch front ends."""
from . import native  # noqa: F401
from .kernels import *  # noqa: F401,F403
