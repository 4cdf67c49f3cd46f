"""cp2_b200 — B200-native (sm_100a) implementation of kimathikaai/CP2's pre-training loss path.

`cp2_b200.builder.MODEL` mirrors the reference's `builder.MODEL`; `cp2_b200.ops` wraps the
C ABI of `cp2_b200/csrc/libcp2_b200.so` (include/cp2_b200.h).  There is no CPU fallback: the
ops raise if the shared library is missing or the tensors are not on a CUDA device.
"""
__version__ = "0.1.0"
