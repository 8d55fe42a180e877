#!/usr/bin/env python
"""Run a few resident passes of the hot path on one GPU -- the command ncu wraps (tools/profile_step.py is what the
launch lists and `--set full` captures under profiles/ were taken from; a number printed under ncu is never a bench value).

  python tools/profile_step.py [--workload chr20] [--scale 0.1] [--steps 2] [--mode collapse-msat] [--k 1001] [--w 100]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="chr20", choices=list(bench.WORKLOADS))
    ap.add_argument("--scale", type=float, default=0.1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--mode", default="collapse-msat")
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--w", type=int, default=0)
    ap.add_argument("--library", default=None, help="an experimental build of libmbgfe.so (tools/build_variant.sh)")
    ap.add_argument("--digest", action="store_true", help="also print an order-free digest of the finished table (to compare builds)")
    args = ap.parse_args()
    wl = dict(bench.WORKLOADS[args.workload])
    if args.k:
        wl["k"] = args.k
    if args.w:
        wl["w"] = args.w
    from verkko_b200 import MbgFrontEnd
    fe = MbgFrontEnd(wl["k"], wl["w"], args.mode, library=args.library)
    bases, offsets, pinned = bench.make_reads(wl, 1, 0, args.scale, pinned_alloc=fe.pinned)
    fe.add_reads(bases, offsets)
    fe.upload()
    for _ in range(args.steps):
        fe.process()
    st = fe.stats()
    fe.parts()
    st2 = fe.stats()
    st.update({k: st2[k] for k in ("collapsed_symbols", "codes", "parts", "hashed_parts")})
    if args.digest:
        import numpy as np
        nh, cov, ed = fe.node_hashes(), fe.node_coverage(), fe.edges()
        st["digest"] = {"nodes": int(len(nh)), "hash_xor": int(np.bitwise_xor.reduce(nh.reshape(-1))), "cov_sum": int(cov.sum(dtype=np.uint64)),
                        "edges": int(len(ed)), "edge_sum": int(ed.astype(np.uint64).sum(dtype=np.uint64))}
    print(json.dumps(st))


if __name__ == "__main__":
    main()
