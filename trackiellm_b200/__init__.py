"""tk-b200-vision: B200-native (sm_100a) implementation of TrackieLLM's vision hot path.

The product is the C-ABI shared library `libtk_b200.so` (trackiellm_b200/csrc,
include/tk_b200_vision.h).  This package only loads it (`_capi`) and mirrors the
reference's host interface for the path (`vision`, `shard`).
"""
__version__ = "0.1.0"
