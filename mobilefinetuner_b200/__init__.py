"""mobilefinetuner_b200 — B200-native (sm_100a) training-step hot path behind MobileFineTuner's operator API.

The product is the C-ABI shared library ``libmft_b200.so`` (hand-written CUDA kernels, declared in
``include/mft_b200.h``) and the C++ host mirror of the reference operator API built on top of it
(``csrc/host``). This Python package only loads the library through ctypes; PyTorch is used by the
tests and the benchmark for device memory, streams and torch.distributed — plumbing, not compute.

There is no CPU fallback: importing :mod:`mobilefinetuner_b200.capi` raises if the library is missing.
"""
__version__ = "0.1.0"
