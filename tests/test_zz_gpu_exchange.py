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
