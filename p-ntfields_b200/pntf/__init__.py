"""pntf -- B200-native hot path of P-NTFields (travel-time network evaluation, Eikonal loss and its
backward, AdamW, batched path integration) behind the reference's NN / Model Python API.

The compute lives in ``libpntf_b200.so`` (hand-written sm_100a CUDA, C-ABI in include/pntf_b200.h);
this package is the thin host side: flat parameter buffers, ctypes calls on the current torch CUDA
stream, the reference-shaped classes, the data-parallel trainer and the synthetic workload generator.
There is no CPU fallback: every op raises if the library or a CUDA device is missing.
"""
from . import capi  # noqa: F401
from .params import FlatParams, TENSOR_NAMES  # noqa: F401
from .nn import NN_single, NN_multi  # noqa: F401
from .model import Model_single, Model_multi  # noqa: F401

__all__ = ["capi", "FlatParams", "TENSOR_NAMES", "NN_single", "NN_multi", "Model_single", "Model_multi"]
