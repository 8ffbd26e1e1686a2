"""b200ps — a Blackwell-native (sm_100a) parameter-server training framework.

Capabilities follow Amano-Ginji/distributed-tensorflow-example (ps/worker roles, async and
synchronous replica updates, MNIST MLP, sparse logistic regression, checkpoints, export),
re-designed for one 8xB200 NVLink/NVSwitch box: every GPU is worker k and parameter-server
shard k, and the hot paths are hand-written CUDA kernels in libb200ps.so.
"""
__version__ = "0.1.0"
