"""The sharded k-mer counting path (mbgfe_exchange) on CPU: the GPU-less emulator build as the per-rank engine.
* world_size 2 and 3 over torch.distributed / gloo: one process per rank, the library's exchange driven through the
  callback transport (the same exchange_run the NCCL transport drives on GPUs);
* the same through mbgfe_exchange_group (all ranks in one process, copies between their buffers).
Checks that the union of the rank partitions equals the single-rank table of the oracle and that every k-mer and edge
landed on its owner."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def shard(bases, offsets, rank, world):
    n = len(offsets) - 1
    lo, hi = rank * n // world, (rank + 1) * n // world
    return bases[int(offsets[lo]):int(offsets[hi])], offsets[lo:hi + 1] - offsets[lo]


def check_partitions(parts, want, world):
    from dist_helpers import owner_of
    from oracle_lib import sorted_edges
    for r, p in enumerate(parts):
        assert all(int(o) == r for o in owner_of(p["node_hash"][:, 1], world)), "k-mer on the wrong owner"
    # owner ranges are ordered, so concatenating the sorted partitions is the global sortByHash order
    nh = np.concatenate([p["node_hash"] for p in parts])
    cov = np.concatenate([p["node_cov"] for p in parts])
    tip = np.concatenate([p["node_tip"] for p in parts])
    assert np.array_equal(nh, want.node_hash) and np.array_equal(cov, want.node_cov) and np.array_equal(tip, want.node_tip)
    # edges: global ids; a rank holds the edges of which it owns `from` when the low bits of the two canonical hashes agree, else those
    # of which it owns `to` (balanced over the ranks: always `from`, the smaller hash, would load the low ranks); together they are
    # the oracle's edge table
    begin = np.concatenate([[0], np.cumsum([len(p["node_cov"]) for p in parts])])
    n_edges = [len(p["edges"]) for p in parts]
    for r, p in enumerate(parts):
        e = p["edges"]
        take_from = ((nh[e[:, 0], 0] ^ nh[e[:, 2], 0]) & np.uint64(1)) == 0      # columns: from, from_fw, to, to_fw, coverage, overlap
        holder = np.where(take_from, np.minimum(e[:, 0], e[:, 2]), np.maximum(e[:, 0], e[:, 2]))
        assert np.all((holder >= begin[r]) & (holder < begin[r + 1])), "edge on the wrong owner"
    if sum(n_edges) > 2000:
        assert max(n_edges) < 1.25 * sum(n_edges) / world + 50, f"edges not balanced over the ranks: {n_edges}"
    edges = np.concatenate([p["edges"] for p in parts])
    assert np.array_equal(sorted_edges(edges), sorted_edges(want.edges))


def _worker(rank, world, port, emu, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import workloads
    from dist_helpers import gloo_callbacks
    from verkko_b200.frontend import NO_TABLE, MbgFrontEnd
    bases, offsets = workloads.coverage_reads(n_reads=8)
    fe = MbgFrontEnd(501, 50, "collapse-msat", flags=NO_TABLE, rank=rank, world=world, library=emu)
    fe.add_reads(*shard(bases, offsets, rank, world))
    fe.run()
    fe.comm_init_callbacks(*gloo_callbacks(world))
    fe.exchange()
    fe.exchange()          # a second exchange over the same instances gives the same partition
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
    fe.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_counting_over_gloo(oracle, emu_library, tmp_path, world):
    import socket
    import torch.multiprocessing as mp
    import workloads
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(world, port, emu_library, str(tmp_path)), nprocs=world, join=True)
    want = oracle.run("collapse-msat", 501, 50, workloads.coverage_reads(n_reads=8))
    check_partitions([np.load(tmp_path / f"rank{r}.npz") for r in range(world)], want, world)


@pytest.mark.parametrize("world", [2, 4])
def test_partitioned_counting_in_one_process(oracle, emu_library, world):
    """mbgfe_exchange_group: all ranks are handles of this process"""
    import workloads
    from verkko_b200.frontend import NO_TABLE, MbgFrontEnd
    bases, offsets = workloads.coverage_reads(n_reads=8)
    fes = [MbgFrontEnd(501, 50, "collapse-msat", flags=NO_TABLE, rank=r, world=world, library=emu_library) for r in range(world)]
    try:
        for r, fe in enumerate(fes):
            fe.add_reads(*shard(bases, offsets, r, world))
            fe.run()
        MbgFrontEnd.exchange_group(fes)
        parts = [dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges()) for fe in fes]
    finally:
        for fe in fes:
            fe.close()
    check_partitions(parts, oracle.run("collapse-msat", 501, 50, (bases, offsets)), world)


@pytest.mark.parametrize("seed", range(8))
def test_partitioned_counting_random(oracle, emu_library, seed):
    """seeded random sharded runs in one process: world 2 .. 8 (also more ranks than reads), any mode, reads appended in one or two
    batches per rank, an empty rank, the sketch-sized tables on odd seeds -- the partitions together are the oracle's single table,
    and the edges are balanced over the ranks"""
    import workloads
    from verkko_b200 import synth
    from verkko_b200.frontend import NO_TABLE, MbgFrontEnd
    rng = np.random.default_rng(300 + seed)
    world = [2, 3, 5, 8, 4, 7, 6, 2][seed]
    mode = ["collapse-msat", "hpc", "no", "msat", "collapse-dinuc", "dinuc", "collapse", "collapse-msat"][seed]
    k = 31 + 2 * int(rng.integers(10, 120))
    w = int(rng.integers(3, min(50, k - 30) + 1))
    g = synth.genome(int(rng.integers(1, 1 << 20)), int(rng.integers(8000, 16000)))
    lmin = max(2 * k + 100, 1200)
    bases, offsets = synth.reads(int(rng.integers(1, 1 << 20)), g, int(rng.integers(6, 40)), lmin, lmin + 1500, int(rng.integers(500, 6000)))
    if seed & 1:
        os.environ["MBGFE_SKETCH_MIN"] = "0"
    try:
        fes = [MbgFrontEnd(k, w, mode, flags=NO_TABLE, rank=r, world=world, library=emu_library) for r in range(world)]
        try:
            for r, fe in enumerate(fes):
                b, o = shard(bases, offsets, r, world)
                n = len(o) - 1
                if seed % 3 == 0 and n >= 2:          # the rank's reads in two appended batches
                    cut = n // 2
                    fe.add_reads(b[:int(o[cut])], o[:cut + 1])
                    fe.run_append()
                    fe.add_reads(b[int(o[cut]):], o[cut:] - o[cut])
                    fe.run_append()
                else:
                    fe.add_reads(b, o)
                    fe.run()
            MbgFrontEnd.exchange_group(fes)
            parts = [dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges()) for fe in fes]
        finally:
            for fe in fes:
                fe.close()
    finally:
        os.environ.pop("MBGFE_SKETCH_MIN", None)
    check_partitions(parts, oracle.run(mode, k, w, (bases, offsets)), world)

