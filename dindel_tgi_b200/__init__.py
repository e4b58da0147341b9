"""dindel_tgi_b200 — B200-native read x candidate-haplotype likelihood engine (Dindel hot path).

Only what the path needs: `csrc/` (CUDA kernels + the C ABI, built into libdindel_b200.so),
`engine` (ctypes mirror of the reference interface over that ABI) and `synth` (seeded inputs).
Importing the package does not load the CUDA library; `engine.load_library()` does and fails
loudly when it is missing (there is no CPU fallback).
"""
from . import engine  # noqa: F401

__all__ = ["engine"]
