#!/usr/bin/env python
"""Parity of the sharded table on N real GPUs (launched with torchrun, one rank per GPU): the concatenation of the ranks'
node and edge partitions must equal the oracle's single table on the same reads.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py
"""
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dist.init_process_group("gloo")          # host plumbing (NCCL id hand-out, gathering the results); the data path is mbgfe_exchange over NCCL
    import workloads
    from dist_helpers import owner_of
    from oracle_lib import Oracle, sorted_edges
    from verkko_b200.frontend import NO_TABLE, MbgFrontEnd
    ok = True
    for mode, k, w, n_reads in (("collapse-msat", 1001, 100, 64), ("hpc", 501, 50, 24)):
        bases, offsets = workloads.coverage_reads(n_reads=n_reads)
        n = len(offsets) - 1
        lo, hi = rank * n // world, (rank + 1) * n // world
        fe = MbgFrontEnd(k, w, mode, device=local, flags=NO_TABLE, rank=rank, world=world)
        fe.add_reads(bases[int(offsets[lo]):int(offsets[hi])], offsets[lo:hi + 1] - offsets[lo])
        fe.run()
        uid = [fe.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        fe.comm_init_nccl(uid[0])
        fe.exchange()
        mine = dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            want = Oracle().run(mode, k, w, (bases, offsets))
            begin = np.concatenate([[0], np.cumsum([len(p["node_cov"]) for p in gathered])])
            for r, p in enumerate(gathered):
                assert all(int(o) == r for o in owner_of(p["node_hash"][:, 1], world)), "k-mer on the wrong owner"
                e = p["edges"]
                assert np.all((e[:, 0] >= begin[r]) & (e[:, 0] < begin[r + 1])), "edge on the wrong owner"
            nh = np.concatenate([p["node_hash"] for p in gathered])
            good = (np.array_equal(nh, want.node_hash) and np.array_equal(np.concatenate([p["node_cov"] for p in gathered]), want.node_cov)
                    and np.array_equal(np.concatenate([p["node_tip"] for p in gathered]), want.node_tip)
                    and np.array_equal(sorted_edges(np.concatenate([p["edges"] for p in gathered])), sorted_edges(want.edges)))
            print(f"multi-gpu parity world={world} {mode} k={k} w={w}: {len(nh)} nodes, {len(want.edges)} edges -> {'OK' if good else 'MISMATCH'}")
            ok = ok and good
        fe.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0 and not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
