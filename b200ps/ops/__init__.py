"""Native sm_100a kernels (libb200ps.so) and their torch front ends."""
from . import native  # noqa: F401
from .native import (ACT_NONE, ACT_SIGMOID, DT_BF16, DT_F32, EPI_ATOMIC_ADD, EPI_SGD,  # noqa: F401
                     EPI_STORE, NativeError)
from .kernels import *  # noqa: F401,F403
