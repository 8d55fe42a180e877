"""Input larger than one resident batch: mbgfe_run_append batch by batch, then ONE table over the selected
k-mers of all batches (mbgfe_build_table, or mbgfe_exchange for the sharded table) == the reference's single HashList over the
whole input (the oracle on all reads)."""
import numpy as np
import pytest

import workloads
from oracle_lib import sorted_edges
from verkko_b200.frontend import NO_TABLE, MbgFrontEnd


def batches_of(bases, offsets, n_batches):
    n = len(offsets) - 1
    out = []
    for b in range(n_batches):
        lo, hi = b * n // n_batches, (b + 1) * n // n_batches
        out.append((bases[int(offsets[lo]):int(offsets[hi])].copy(), (offsets[lo:hi + 1] - offsets[lo]).copy()))
    return out


def run_in_batches(mode, k, w, reads, n_batches, library=None, files=None, **kw):
    fe = MbgFrontEnd(k, w, mode, library=library, **kw)
    try:
        for i, (b, o) in enumerate(batches_of(*reads, n_batches)):
            if files is not None:            # every other batch comes from a FASTA file (the streaming loader appends too)
                from oracle_lib import write_fasta
                if i % 2:
                    p = files / f"batch{i}.fa"
                    write_fasta(p, b, o)
                    fe.add_file(str(p))
                else:
                    fe.add_reads(b, o)
            else:
                fe.add_reads(b, o)
            fe.run_append()
        fe.build_table()
        return dict(positions=fe.positions(), hashes=fe.hashes(), node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges(),
                    stats=fe.stats())
    finally:
        fe.close()


def check(got, want):
    assert np.array_equal(got["positions"], want.positions) and np.array_equal(got["hashes"], want.hashes), "instances of all batches, in input order"
    assert np.array_equal(got["node_hash"], want.node_hash) and np.array_equal(got["node_cov"], want.node_cov) and np.array_equal(got["node_tip"], want.node_tip)
    assert np.array_equal(sorted_edges(got["edges"]), sorted_edges(want.edges))
    assert got["stats"]["selected"] == len(want.positions)


def test_emulated_append_runs(oracle, emu_library, tmp_path):
    reads = workloads.as_pair(workloads.coverage_reads(n_reads=12) if True else None)
    want = oracle.run("collapse-msat", 301, 30, reads)
    check(run_in_batches("collapse-msat", 301, 30, reads, 3, library=emu_library, files=tmp_path, piece_bytes=20000), want)
    check(run_in_batches("collapse-msat", 301, 30, reads, 1, library=emu_library), want)


def test_emulated_append_runs_sharded(oracle, emu_library):
    """two ranks, two batches each, one exchange at the end"""
    from test_distributed_cpu import check_partitions, shard
    reads = workloads.as_pair(workloads.coverage_reads(n_reads=12))
    world = 2
    fes = [MbgFrontEnd(301, 30, "collapse-msat", flags=NO_TABLE, rank=r, world=world, library=emu_library) for r in range(world)]
    try:
        for r, fe in enumerate(fes):
            for b, o in batches_of(*shard(*reads, r, world), 2):
                fe.add_reads(b, o)
                fe.run_append()
        MbgFrontEnd.exchange_group(fes)
        parts = [dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges()) for fe in fes]
    finally:
        for fe in fes:
            fe.close()
    check_partitions(parts, oracle.run("collapse-msat", 301, 30, reads), world)


@pytest.mark.gpu
@pytest.mark.parametrize("n_batches", [2, 5])
def test_append_runs_on_gpu(oracle, tmp_path, n_batches):
    from verkko_b200 import synth
    g = synth.genome(7, 300000)
    reads = synth.reads(11, g, 480, 11000, 19000, 1000)
    want = oracle.run("collapse-msat", 1001, 100, reads)
    check(run_in_batches("collapse-msat", 1001, 100, reads, n_batches, files=tmp_path), want)
    check(run_in_batches("collapse-msat", 1001, 100, reads, n_batches, piece_bytes=1 << 20), want)
