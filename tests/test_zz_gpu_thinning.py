"""GPU parity of mbgfe_thin_unitigs (the newest entry point; kept in a file of its own that sorts last)."""
import numpy as np
import pytest

import workloads
from oracle_lib import lcg_bases


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 1001, 100, 1, 2), ("collapse-msat", 301, 30, 1, 2), ("hpc", 63, 33, 1, 3), ("collapse-msat", 101, 20, 2, 2), ("no", 63, 5, 1, 2)])
def test_unitig_kmer_thinning_against_oracle(oracle, mode, k, w, a, u):
    """the last piece of row f1 under MBG's defaults: filterKmersToUnitigKmers with filterWithinUnitig = true
    (MBG/MBG.cpp:1065-1185) -- kept k-mers and the overlaps of the new neighbours -- on the device against the oracle
    (pinned on the reference's whole pass-3 sequence in tests/test_oracle.py)"""
    from fe_helpers import check_thinning
    from verkko_b200.frontend import MbgFrontEnd
    big = workloads.coverage_reads(genome_len=120000, n_reads=200, len_min=9000, len_max=14000, error_ppm=3000) if k < 1001 else \
        workloads.coverage_reads(genome_len=200000, n_reads=300, len_min=12000, len_max=20000, error_ppm=2000)
    checked = 0
    for reads in (big, workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        fe = MbgFrontEnd(k, w, mode)
        try:
            fe.add_reads(*reads) if isinstance(reads, tuple) else fe.add_read_list(reads)
            fe.run()
            ug = fe.build_unitigs(a, u, one_cov_heuristic=True)
            thin, off = fe.thin_unitigs()
            thin, off = thin.copy(), off.copy()
        finally:
            fe.close()
        _, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=True, thin=True)
        checked += check_thinning(ug, thin, off, want, k, f"thinning {mode} k={k}")
    assert checked > 10


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 1001, 100, 1, 2), ("collapse-msat", 301, 30, 1, 2), ("hpc", 63, 33, 1, 3), ("collapse-msat", 101, 20, 2, 2), ("no", 63, 5, 1, 2)])
def test_thinning_applied_and_read_paths_on_the_thinned_list(oracle, mode, k, w, a, u):
    """row f2 as every default MBG run (and Verkko) takes it: filterKmersToUnitigKmers applied (second HashList::filter,
    renumbered unitigs, new sequence overlaps, MBG/MBG.cpp:1166-1184) and getReadPaths on the thinned list (MBG.cpp:1586;
    KmerMatcher.h:55-141 with the new overlaps), on the device against the oracle"""
    from fe_helpers import run_thinned_pass
    big = workloads.coverage_reads(genome_len=120000, n_reads=200, len_min=9000, len_max=14000, error_ppm=3000) if k < 1001 else \
        workloads.coverage_reads(genome_len=200000, n_reads=300, len_min=12000, len_max=20000, error_ppm=2000)
    n_paths = n_new = 0
    for reads in (big, workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)],
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        p, o = run_thinned_pass(mode, k, w, reads, a, u, oracle, what=f"thinned pass {mode} k={k}")
        n_paths += p
        n_new += o
    assert n_paths > 20 and n_new > 20
