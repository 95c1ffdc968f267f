"""Operators.

``reference``  plain PyTorch/numpy restatements (fp32/fp64) of every op — the
               numerical oracle for the CUDA kernels and the restated baseline;
``native``     ctypes bindings to ``libdtf_b200.so`` (hand-written sm_100a kernels,
               C ABI declared in ``include/dtf_b200.h``).
"""
