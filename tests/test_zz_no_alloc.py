"""mbgfe_kat_kmer_hash_noalloc (include/mbgfe.h): the k-mer hash through the C ABI with its device scratch in a static
__device__ array instead of cudaMalloc'ed staging.  No counterpart in the reference; the checker is the existing entry
point mbgfe_kat_kmer_hash (itself pinned on MBG's golden hash and the oracle in tests/test_gpu_parity.py).  Integer work:
the two must agree bit for bit.  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NOALLOC_MAX_K = 16384      # MBGFE_NOALLOC_MAX_K


def test_noalloc_kmer_hash_equals_the_allocating_entry_point(oracle):
    from verkko_b200 import MbgFrontEnd
    from verkko_b200.frontend import MbgfeError
    fe = MbgFrontEnd(1001, 100)
    try:
        rng = np.random.default_rng(17)
        # two calls in a row on the shared scratch, the second shorter than the first: what the first left must not leak in
        first = rng.integers(0, 30948, 2001).astype(np.uint16)
        second = rng.integers(0, 30948, 31).astype(np.uint16)
        got1 = fe.kat_kmer_hash_noalloc(first)
        got2 = fe.kat_kmer_hash_noalloc(second)
        assert got1 == fe.kat_kmer_hash(first) == oracle.kmer_hash(first)
        assert got2 == fe.kat_kmer_hash(second) == oracle.kmer_hash(second)
        # MBG's own known answer (tests/test_gpu_parity.py::test_device_tables_known_answers)
        codes = np.array([(i * 7) % 4 for i in range(31)], dtype=np.uint16)
        assert fe.kat_kmer_hash_noalloc(codes) == (0xc8ec520a381594d1, 0x6515f3d84acb3010)
        # a sweep of sizes up to the limit, long then short then long again; the same call twice gives the same answer
        for k in (NOALLOC_MAX_K, 33, 1001, 35, NOALLOC_MAX_K - 1, 37, 64, 63, 65):
            codes = rng.integers(0, 30948, k).astype(np.uint16)
            want = fe.kat_kmer_hash(codes)
            assert fe.kat_kmer_hash_noalloc(codes) == want, k
            assert fe.kat_kmer_hash_noalloc(codes) == want, k
        # beyond the static array: a clear error, and the handle stays usable
        too_long = rng.integers(0, 30948, NOALLOC_MAX_K + 1).astype(np.uint16)
        with pytest.raises(MbgfeError, match="MBGFE_NOALLOC_MAX_K"):
            fe.kat_kmer_hash_noalloc(too_long)
        assert fe.kat_kmer_hash_noalloc(second) == got2
    finally:
        fe.close()


def test_noalloc_kmer_hash_leaves_free_memory_alone():
    import torch
    from verkko_b200 import MbgFrontEnd
    fe = MbgFrontEnd(1001, 100)
    try:
        codes = np.random.default_rng(5).integers(0, 30948, 1001).astype(np.uint16)
        want = fe.kat_kmer_hash_noalloc(codes)       # first use on this device: resolves the symbol's address
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(0)
        for _ in range(3):
            assert fe.kat_kmer_hash_noalloc(codes) == want
        free_after, _ = torch.cuda.mem_get_info(0)
        assert free_after == free_before
    finally:
        fe.close()
