"""deep_vision_b200 — B200-native (sm_100a) evidence indexing for DeepVision.

Only the report-evidence hot path: CUDA kernels behind the C-ABI
`libdvb200.so` (include/dvb200.h) and the host-side mirror of the reference's Python
signatures.  No CPU fallback: compute entry points raise without the library + a GPU.
"""
from . import _lib  # noqa: F401
from ._lib import Context, DvbError  # noqa: F401

__version__ = "0.1.0"
