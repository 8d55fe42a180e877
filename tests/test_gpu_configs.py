"""Parity at the size and shape of BASELINE.json's configurations: config #1 at full size against the
oracle (T1 + T2 + T3 of SURVEY.md 8c), HERRO-shaped reads through the abundance filter at -a 2 / -a 3, and the order-free
table digest the bench uses for its `parity_sample` check."""
import numpy as np
import pytest

import workloads
from fe_helpers import run_frontend, run_frontend_unitigs
from oracle_lib import compare_runs, sorted_edges, table_digest

pytestmark = pytest.mark.gpu


def test_config1_ecoli_full_size(oracle):
    """BASELINE.json configs[0]: E. coli-shaped synthetic HiFi 24x -- 5 Mbp genome, 8 000 reads of 11-19 kb (~120 Mbase),
    k = 1001, w = 100, collapse-msat; every per-part array (T1), the node table (T2) and the edge table (T3), bit for bit"""
    from verkko_b200 import synth
    g = synth.genome(7, 5_000_000)
    reads = synth.reads(11, g, 8000, 11000, 19000, 1000)
    got = run_frontend("collapse-msat", 1001, 100, reads)
    want = oracle.run("collapse-msat", 1001, 100, reads)
    assert got["stats"]["raw_bases"] > 110_000_000 and len(want.node_cov) > 100_000
    compare_runs(got, want, what="config #1 (E. coli-shaped, full size)")
    assert got["stats"]["selected"] == int(want.node_cov.sum()) == len(want.positions)
    assert table_digest(got.node_hash, got.node_cov, got.node_tip, got.edges) == table_digest(want.node_hash, want.node_cov, want.node_tip, want.edges)


@pytest.mark.parametrize("ppm,a,u", [(2500, 2, 2), (2500, 3, 3), (2500, 2, 3.5), (7500, 2, 2), (7500, 3, 3)])
def test_herro_shaped_abundance_filter(oracle, ppm, a, u):
    """BASELINE.json configs[3]: HERRO / ONT-duplex-shaped reads (30-50 kb, 0.25 % and 0.75 % residual errors) -- 96 % to
    99.8 % of the distinct k-mers carry an error and are seen once, so the -a filter does real work: table,
    mbgfe_build_unitigs_ex (with and without the oneCovHeuristic) against the oracle at -a 2 and -a 3"""
    from test_gpu_parity import check_unitigs
    from verkko_b200 import synth
    g = synth.genome(23, 400_000)
    reads = synth.reads(5, g, 300, 30000, 50000, ppm)
    for one_cov in (False, True):
        table, got = run_frontend_unitigs("collapse-msat", 1001, 100, reads, a, u, one_cov=one_cov)
        otable, want = oracle.run_unitigs("collapse-msat", 1001, 100, reads, a, u, one_cov=one_cov)
        assert np.array_equal(table.node_hash, otable.node_hash) and np.array_equal(table.node_cov, otable.node_cov) and np.array_equal(table.node_tip, otable.node_tip)
        assert np.array_equal(sorted_edges(table.edges), sorted_edges(otable.edges))
        check_unitigs(got, want, f"herro {ppm} ppm -a {a} -u {u} one_cov={one_cov}")
        assert 0 < got.kept.sum() < len(got.kept), "the abundance filter must remove some k-mers and keep some"
    # the filter did real work: most distinct k-mers of such reads are error k-mers seen once
    assert (otable.node_cov < a).mean() > 0.9
