"""dist-gpu-accelerated-tree-search_b200 - B200-native (sm_100a) bound -> prune -> branch engine.

Python here is a thin ctypes binding over the C ABI of libbbgpu.so (include/bb_gpu.h) for tests and bench.py;
the product is the shared library and the C host executables under host/. The directory name is not a valid
Python identifier: import it through `__graft_entry__.load_package()` (registered as module `dgats_b200`).

There is no CPU fallback: importing works without a GPU (symbols only), every compute call needs a CUDA device
and raises BBError otherwise; a missing libbbgpu.so raises at import time.
"""
from .capi import *  # noqa: F401,F403
