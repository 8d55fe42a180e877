"""Synthetic HiFi-shaped workloads (ctypes wrapper of csrc/synth.cpp; host only, not on the hot path)."""
import ctypes as C
import os

import numpy as np

from .build import build_synth

_lib = None


def _load():
    global _lib
    if _lib is None:
        L = C.CDLL(build_synth())
        L.mbgsynth_genome.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32]
        L.mbgsynth_read_lengths.argtypes = [C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                            C.c_uint32, C.c_void_p, C.c_uint32]
        L.mbgsynth_reads.argtypes = [C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                     C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]
        _lib = L
    return _lib


def genome(seed, length, locus_spacing=250, threads=None):
    threads = threads or min(32, os.cpu_count() or 1)
    g = np.empty(length, dtype=np.uint8)
    _load().mbgsynth_genome(seed, length, locus_spacing, g.ctypes.data, threads)
    return g


def read_offsets(seed, g, n_reads, len_min, len_max, error_ppm=1000, first_read=0, threads=None):
    threads = threads or min(32, os.cpu_count() or 1)
    lengths = np.zeros(n_reads, dtype=np.uint64)
    _load().mbgsynth_read_lengths(seed, g.ctypes.data, len(g), first_read, n_reads, len_min, len_max, error_ppm,
                                  lengths.ctypes.data, threads)
    offsets = np.zeros(n_reads + 1, dtype=np.uint64)
    np.cumsum(lengths, out=offsets[1:])
    return offsets


def fill_reads(seed, g, n_reads, len_min, len_max, offsets, out, error_ppm=1000, first_read=0, threads=None):
    """writes the reads into `out` (uint8 array of at least offsets[-1] bytes; may be pinned memory)"""
    threads = threads or min(32, os.cpu_count() or 1)
    _load().mbgsynth_reads(seed, g.ctypes.data, len(g), first_read, n_reads, len_min, len_max, error_ppm,
                           offsets.ctypes.data, out.ctypes.data, threads)


def reads(seed, g, n_reads, len_min, len_max, error_ppm=1000, first_read=0, threads=None):
    offsets = read_offsets(seed, g, n_reads, len_min, len_max, error_ppm, first_read, threads)
    bases = np.empty(int(offsets[-1]), dtype=np.uint8)
    fill_reads(seed, g, n_reads, len_min, len_max, offsets, bases, error_ppm, first_read, threads)
    return bases, offsets
