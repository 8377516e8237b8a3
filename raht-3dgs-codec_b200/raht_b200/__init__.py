"""raht_b200 -- B200 (sm_100a) implementation of the RAHT-3DGS encode/decode hot path.

Host side of libraht_b200.so.  Importing this package loads the native library and fails loudly
when it is missing; there is no CPU fallback.
"""
from . import _lib
from . import ops
from .ops import RahtTree, LevelSequence, RlgrStreams

__all__ = ["_lib", "ops", "RahtTree", "LevelSequence", "RlgrStreams"]
