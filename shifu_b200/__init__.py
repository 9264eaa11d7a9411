"""shifu_b200 -- B200-native (sm_100a) implementation of Shifu's GBT / RF training hot path.

Package contents: csrc/ (CUDA kernels + the C ABI of include/shifu_b200.h), capi.py (ctypes binding of that ABI),
messages.py (builders of the flattened DTMasterParams message).  There is no CPU fallback.
"""
__version__ = "0.1"
