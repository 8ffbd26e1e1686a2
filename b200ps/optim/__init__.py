from .adam import Adam  # noqa: F401
from .sgd import AsyncSGD, SyncReplicasSGD  # noqa: F401
