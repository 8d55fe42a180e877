"""verkko_b200 -- B200-native front end of Verkko's buildGraph stage (MBG): homopolymer / microsatellite
compression, syncmer selection over large k, 128-bit k-mer hashing and hashed k-mer counting.

The product is the C-ABI library ``libmbgfe.so`` (include/mbgfe.h; CUDA kernels in verkko_b200/csrc).  This package
is the thin host-side mirror used by the tests and bench.py: ``MbgFrontEnd`` wraps the C ABI with ctypes.
"""
from .frontend import MbgFrontEnd, MbgfeError, MODES, library_path  # noqa: F401
