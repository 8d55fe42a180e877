#!/usr/bin/env python
"""How much would the resident step gain if the tail of one half of the batch (K3b, K4: latency-bound, issue ~50 %) ran beside the
front of the other half (K1..K3a)?  Two handles with half a batch each, process() called from two host threads at once, against
the same two calls one after the other.   python tools/overlap_probe.py"""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    from verkko_b200.frontend import MbgFrontEnd
    wl = dict(bench.WORKLOADS["chr20"])
    fes, keep = [], []
    for r in range(2):
        fe = MbgFrontEnd(wl["k"], wl["w"], "collapse-msat")
        bases, offsets, pinned = bench.make_reads(wl, 2, r, 0.5, pinned_alloc=fe.pinned)
        keep.append((bases, offsets, pinned))
        fe.add_reads(bases, offsets)
        fe.upload()
        fe.process()
        fes.append(fe)
    import torch
    def both_sequential():
        for fe in fes:
            fe.process()
    def both_concurrent():
        ts = [threading.Thread(target=fe.process) for fe in fes]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
    for name, fn in (("sequential", both_sequential), ("concurrent", both_concurrent), ("sequential", both_sequential), ("concurrent", both_concurrent)):
        fn()
        torch.cuda.synchronize()
        t0 = time.time()
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        print(name, round((time.time() - t0) / 5 * 1e3, 3), "ms per pair of half batches", [round(fe.stats()["ms_process"], 2) for fe in fes])


if __name__ == "__main__":
    main()
