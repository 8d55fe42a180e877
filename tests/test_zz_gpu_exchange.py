"""The exchange kernels and the in-library exchange (mbgfe_exchange_group: every rank a handle of this process) on ONE GPU, so
that the driver's GPU test run covers the multi-GPU data path: k_owner_count2 / k_owner_pack / k_insert_records2 / k_slot_ids /
k_edge_owner_pack2 / k_insert_edge_records and the sharded export, against the oracle's single table."""
import numpy as np
import pytest

import workloads
from test_distributed_cpu import check_partitions, shard
from verkko_b200.frontend import NO_TABLE, MbgFrontEnd

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,mode,k,w", [(2, "collapse-msat", 1001, 100), (4, "collapse-msat", 1001, 100), (8, "hpc", 501, 50), (3, "no", 63, 33)])
def test_sharded_table_on_one_gpu(oracle, world, mode, k, w):
    from verkko_b200 import synth
    g = synth.genome(7, 200000)
    bases, offsets = synth.reads(11, g, 320, 11000, 19000, 1000)
    fes = [MbgFrontEnd(k, w, mode, flags=NO_TABLE, rank=r, world=world) for r in range(world)]
    try:
        for r, fe in enumerate(fes):
            fe.add_reads(*shard(bases, offsets, r, world))
            fe.run()
        MbgFrontEnd.exchange_group(fes)
        parts = [dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges()) for fe in fes]
        sent = sum(fe.stats()["exchange_bytes"] for fe in fes)
        # a second exchange over the same resident instances gives the same partitions
        MbgFrontEnd.exchange_group(fes)
        again = [dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges()) for fe in fes]
    finally:
        for fe in fes:
            fe.close()
    want = oracle.run(mode, k, w, (bases, offsets))
    check_partitions(parts, want, world)
    check_partitions(again, want, world)
    assert sent >= 32 * len(want.positions)


def test_sketch_sized_tables_on_gpu(oracle, monkeypatch):
    """tables sized from the distinct-count sketch (by default only for inputs above 2^26 instances): the exact tail of one handle,
    with an estimate forced 16x too low (overflow -> doubling), and the sharded tables of an exchange"""
    from verkko_b200 import synth
    g = synth.genome(7, 120000)
    bases, offsets = synth.reads(11, g, 400, 9000, 15000, 1000)          # ~40x coverage
    want = oracle.run("collapse-msat", 501, 50, (bases, offsets))
    monkeypatch.setenv("MBGFE_SKETCH_MIN", "0")

    def one():
        fe = MbgFrontEnd(501, 50, "collapse-msat")
        try:
            fe.add_reads(bases, offsets)
            fe.upload()
            fe.process()
            return dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges(), stats=fe.stats())
        finally:
            fe.close()

    got = one()
    assert all(np.array_equal(got[f], want[f]) for f in ("node_hash", "node_cov", "node_tip", "edges"))
    assert 2 * len(want.node_hash) <= got["stats"]["table_slots"] < len(want.positions)      # two per distinct k-mer, far below two per instance
    monkeypatch.setenv("MBGFE_SKETCH_DIV", "16")
    low = one()
    assert all(np.array_equal(low[f], want[f]) for f in ("node_hash", "node_cov", "node_tip", "edges")) and low["stats"]["table_grows"] >= 1
    monkeypatch.delenv("MBGFE_SKETCH_DIV")
    world = 3
    fes = [MbgFrontEnd(501, 50, "collapse-msat", flags=NO_TABLE, rank=r, world=world) for r in range(world)]
    try:
        for r, fe in enumerate(fes):
            fe.add_reads(*shard(bases, offsets, r, world))
            fe.run()
        MbgFrontEnd.exchange_group(fes)
        parts = [dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges()) for fe in fes]
        assert all(fe.stats()["table_slots"] < 2 * len(want.positions) // world for fe in fes)      # sized from the instances it would be >= 2 per record
    finally:
        for fe in fes:
            fe.close()
    check_partitions(parts, want, world)

