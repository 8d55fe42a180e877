"""T4 of SURVEY.md 8(c): the device's tables handed to the reference's OWN back end give the GFA and GAF of `MBG -t 1` byte for
byte.  The reference side (oracle/_ref/MBG and oracle/_ref/libmbgref.so with oracle/ref_shim.cpp:refshim_backend) is test
infrastructure compiled from the reference's sources; the product never loads it."""
import hashlib

import numpy as np
import pytest

from oracle_lib import have_reference, reference_backend, reference_cli, write_fasta
from verkko_b200.frontend import MbgFrontEnd

pytestmark = pytest.mark.gpu


def device_tables(fa, mode, k, w, a, u, one_cov):
    """file -> table -> getUnitigGraph -> thinning applied, all on the device (rows a1 .. a14, f1)"""
    fe = MbgFrontEnd(k, w, mode)
    try:
        fe.add_file(str(fa))
        fe.run()
        table = dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
        ug = fe.build_unitigs(a, u, one_cov_heuristic=one_cov)
        unitig_edges = ug["unitig_edges"].copy()
        fe.thin_unitigs()
        th = fe.thinned()
        th["unitig_edges"] = unitig_edges
        return table, th
    finally:
        fe.close()


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built")
@pytest.mark.parametrize("genome,n_reads,k,w,r,big_r,guesswork", [(300_000, 480, 1001, 100, 0, 0, False), (1_000_000, 1600, 1001, 100, 15000, 4000, True),
                                                                  (100_000, 300, 301, 30, 4000, 1000, True)])
def test_t4_device_tables_through_the_reference_back_end(tmp_path, genome, n_reads, k, w, r, big_r, guesswork):
    """config #1-shaped reads (24x, 11-19 kb, 0.1 % errors); first without resolution (-r 0), then with Verkko's resolution flags
    (-r 15000 -R 4000 --do-unsafe-guesswork-resolutions, -a 1 -u 2, collapse-msat; Snakefiles/1-buildGraph.sm:88-100) but without
    --hpc-variant-onecopy-coverage, whose variant table comes from row f3 (not built)"""
    from verkko_b200 import synth
    g = synth.genome(7, genome)
    bases, offsets = synth.reads(11, g, n_reads, 11000, 19000, 1000)
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    mode, a, u = "collapse-msat", 1, 2
    want_gfa, want_gaf = reference_cli(fa, tmp_path / "ref.gfa", tmp_path / "ref.gaf", mode, k, w, a, u, r, big_r, guesswork)
    assert want_gfa.count(b"\nS\t") > 0 and want_gaf.count(b"\n") > n_reads // 2
    table, th = device_tables(fa, mode, k, w, a, u, one_cov=(u >= 2 and r > 0 and guesswork))
    gfa0, gaf0 = reference_backend(0, fa, tmp_path / "s0.gfa", tmp_path / "s0.gaf", mode, k, w, a, u, r, big_r, guesswork, **table)
    assert hashlib.md5(gfa0).hexdigest() == hashlib.md5(want_gfa).hexdigest(), "stage 0 (table) GFA differs from MBG -t 1"
    assert hashlib.md5(gaf0).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 0 (table) GAF differs from MBG -t 1"
    gfa1, gaf1 = reference_backend(1, fa, tmp_path / "s1.gfa", tmp_path / "s1.gaf", mode, k, w, a, u, r, big_r, guesswork,
                                   node_hash=th["node_hash"], node_cov=th["node_cov"], node_tip=th["node_tip"], edges=th["edges"], new_overlaps=th["new_overlaps"],
                                   unitig_begin=th["unitig_begin"], unitig_nodes=th["unitig_nodes"], unitig_edges=th["unitig_edges"])
    assert hashlib.md5(gfa1).hexdigest() == hashlib.md5(want_gfa).hexdigest(), "stage 1 (thinned list + unitigs) GFA differs from MBG -t 1"
    assert hashlib.md5(gaf1).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 1 (thinned list + unitigs) GAF differs from MBG -t 1"


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built")
@pytest.mark.parametrize("k,w,r,big_r", [(301, 30, 4000, 1000), (1001, 100, 15000, 4000)])
def test_t4_verkko_flag_set_with_hpc_variants(tmp_path, k, w, r, big_r):
    """Verkko's flag set (Snakefiles/1-buildGraph.sm:88-100): --hpc-variant-onecopy-coverage on, resolution on, guesswork, -a 1
    -u 2, collapse-msat.  The variant table is the reference's own (passes 1-2; row f3 is not built) and goes in through
    mbgfe_set_variants; pass 3 on the device (rows a1-a14, f1) through the reference's back end = MBG -t 1, byte for byte."""
    import workloads
    from oracle_lib import reference_hpc_variants
    bases, offsets = workloads.diploid_msat_reads(len_min=6000 if k < 1001 else 11000, len_max=10000 if k < 1001 else 19000)
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    mode, a, u, onecopy = "collapse-msat", 1, 2, 8
    want_gfa, want_gaf = reference_cli(fa, tmp_path / "ref.gfa", tmp_path / "ref.gaf", mode, k, w, a, u, r, big_r, True, extra=("--hpc-variant-onecopy-coverage", str(onecopy)))
    variants = reference_hpc_variants(fa, mode, k, w, a, u, onecopy)
    assert len(variants.offsets) > 10
    fe = MbgFrontEnd(k, w, mode)
    try:
        fe.add_file(str(fa))
        fe.set_variants(variants.hashes, variants.offsets, variants.len_begin, variants.lengths)
        fe.run()
        table = dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
        ug = fe.build_unitigs(a, u, one_cov_heuristic=True)
        unitig_edges = ug["unitig_edges"].copy()
        fe.thin_unitigs()
        th = fe.thinned()
    finally:
        fe.close()
    gfa0, gaf0 = reference_backend(0, fa, tmp_path / "s0.gfa", tmp_path / "s0.gaf", mode, k, w, a, u, r, big_r, True, variants=variants, **table)
    assert hashlib.md5(gfa0).hexdigest() == hashlib.md5(want_gfa).hexdigest() and hashlib.md5(gaf0).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 0"
    gfa1, gaf1 = reference_backend(1, fa, tmp_path / "s1.gfa", tmp_path / "s1.gaf", mode, k, w, a, u, r, big_r, True, variants=variants,
                                   node_hash=th["node_hash"], node_cov=th["node_cov"], node_tip=th["node_tip"], edges=th["edges"], new_overlaps=th["new_overlaps"],
                                   unitig_begin=th["unitig_begin"], unitig_nodes=th["unitig_nodes"], unitig_edges=unitig_edges)
    assert hashlib.md5(gfa1).hexdigest() == hashlib.md5(want_gfa).hexdigest() and hashlib.md5(gaf1).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 1"
