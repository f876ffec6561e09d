"""sparrowhawk_b200 — B200-native (sm_100a) front end of the sparrowhawk-asm
assembler: k-mer extraction + counting + min-count filter + de Bruijn graph +
unitig compaction as hand-written CUDA kernels behind a C ABI (include/shk.h).

The Python layer is a thin host-side mirror of the reference's `AssemblyHelper`
interface (www/src/workers/Assembler.ts:15-39); the compute lives in
libshk_b200.so and there is no CPU fallback.
"""
from ._ffi import ShkError, lib, LIB_PATH  # noqa: F401
from .assembly_helper import AssemblyHelper  # noqa: F401
from . import synth  # noqa: F401

__all__ = ["AssemblyHelper", "ShkError", "synth", "lib", "LIB_PATH"]
