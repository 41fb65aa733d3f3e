"""metaopt_b200: B200-native candidate-scoring hot path for Orion's Bayesian-optimisation suggest().

Only what the path needs: `csrc/` (CUDA kernels + C ABI, built into libob200.so), the ctypes
binding (`_lib`), the device engine (`engine`), the host-side mirror of the reference's
Space/Dimension/BaseAlgorithm interface (`space`, `algo`) and the multi-GPU sharding helpers
(`dist`).  Nothing here imports `oracle/`; a missing CUDA library or device raises.
"""
__version__ = "0.1.0"
