"""world_size-2 test of the sharded k-mer counting path on CPU: gloo backend, the GPU-less emulator build as the
per-rank engine ("device" pointers are host pointers there).  Checks that the union of the two rank partitions equals
the single-rank table of the oracle and that every k-mer landed on its owner."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, emu, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import workloads
    from verkko_b200.distributed import DistributedCounter
    from verkko_b200.frontend import NO_TABLE, MbgFrontEnd
    bases, offsets = workloads.coverage_reads(n_reads=8)
    n = len(offsets) - 1
    lo, hi = rank * n // world, (rank + 1) * n // world
    fe = MbgFrontEnd(501, 50, "collapse-msat", flags=NO_TABLE, rank=rank, world=world, library=emu)
    fe.add_reads(bases[int(offsets[lo]):int(offsets[hi])], offsets[lo:hi + 1] - offsets[lo])
    fe.run()
    counter = DistributedCounter(fe, world, rank, on_gpu=False)
    counter.exchange_and_count()
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges(),
             id_begin=counter.id_begin)
    fe.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_counting(oracle, emu_library, tmp_path, world):
    import socket
    import workloads
    from verkko_b200.distributed import owner_of
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(world, port, emu_library, str(tmp_path)), nprocs=world, join=True)
    want = oracle.run("collapse-msat", 501, 50, workloads.coverage_reads(n_reads=8))
    parts = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    for r, p in enumerate(parts):
        assert all(int(o) == r for o in owner_of(p["node_hash"][:, 1], world)), "k-mer on the wrong owner"
    # owner ranges are ordered, so concatenating the sorted partitions is the global sortByHash order
    nh = np.concatenate([p["node_hash"] for p in parts])
    cov = np.concatenate([p["node_cov"] for p in parts])
    tip = np.concatenate([p["node_tip"] for p in parts])
    assert np.array_equal(nh, want.node_hash) and np.array_equal(cov, want.node_cov) and np.array_equal(tip, want.node_tip)
    # edges: every rank holds the edges whose `from` node it owns, with global ids, each partition sorted like the
    # single-GPU export; concatenated they are the oracle's edge table
    id_begin = parts[0]["id_begin"]
    for r, p in enumerate(parts):
        e = p["edges"]
        assert np.all((e[:, 0] >= id_begin[r]) & (e[:, 0] < id_begin[r + 1])), "edge on the wrong owner"
    from oracle_lib import sorted_edges
    edges = np.concatenate([p["edges"] for p in parts])
    assert np.array_equal(sorted_edges(edges), sorted_edges(want.edges))
