"""rayne_b200 - B200-native scene hot path for Rayne (transform propagation, frustum culling, particles).

The product is ``librn_cuda.so`` (hand-written sm_100a CUDA behind the C ABI declared in ``include/rn_cuda.h``);
this package holds its sources (``csrc/``), the ctypes binding used by the tests and the bench (``capi``), the
synthetic scene generators (``scenegen``) and the one-process-per-GPU plumbing ``bench.py --gpus N`` runs on
(``shard``).  The Rayne-side binding (a ``RN::Scene`` subclass) lives in ``integration/``.
"""
__all__ = ["scenegen"]
