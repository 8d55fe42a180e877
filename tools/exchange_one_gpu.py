#!/usr/bin/env python
"""Two ranks of the bench workload as two handles on ONE GPU (mbgfe_exchange_group), so that the exchange kernels can be looked at
with ncu (which must not wrap a multi-rank command).   python tools/exchange_one_gpu.py [--scale 0.5] [--world 2]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=0.5)
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    from verkko_b200.frontend import NO_TABLE, MbgFrontEnd
    wl = dict(bench.WORKLOADS["chr20"])
    fes, keep = [], []
    for r in range(args.world):
        fe = MbgFrontEnd(wl["k"], wl["w"], "collapse-msat", flags=NO_TABLE, rank=r, world=args.world)
        bases, offsets, pinned = bench.make_reads(wl, args.world, r, args.scale, pinned_alloc=fe.pinned)
        keep.append((bases, offsets, pinned))
        fe.add_reads(bases, offsets)
        fe.upload()
        fe.process()
        fes.append(fe)
    os.environ["MBGFE_EXCHANGE_TRACE"] = "1"
    for _ in range(args.steps):
        MbgFrontEnd.exchange_group(fes)
    print(json.dumps([{k: fe.stats()[k] for k in ("selected", "distinct", "distinct_edges", "ms_exchange")} for fe in fes]))
    # the one-table path on the same instances of rank 0, for comparison
    fe = MbgFrontEnd(wl["k"], wl["w"], "collapse-msat")
    fe.add_reads(keep[0][0], keep[0][1])
    fe.upload()
    fe.process()
    fe.process()
    print(json.dumps({k: fe.stats()[k] for k in ("selected", "distinct", "distinct_edges", "ms_nodes", "ms_edges")}))


if __name__ == "__main__":
    main()
