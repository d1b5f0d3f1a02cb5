"""move2kube_b200: B200-native source detection for move2kube (one path of the reference).

The product is the CUDA shared library `libm2kdetect.so` next to this file (C ABI in
include/m2kdetect.h).  The Python modules are the host-side mirror of the reference's interfaces for
that path (source.Translator, containerizer.Containerizer, the three common helpers) and a ctypes
binding; nothing here computes detection results on the CPU.
"""
from . import _abi  # noqa: F401

__all__ = ["_abi", "lib", "plan", "detection", "containerizer", "source", "common"]
