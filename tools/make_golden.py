#!/usr/bin/env python
"""Writes tests/golden/*.npz from the REAL reference (oracle/_ref/libmbgref.so, built from /root/reference by
`make -C oracle ref`).  Run in the build container; the fixtures travel with the repo to the GPU box where
/root/reference does not exist."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import workloads  # noqa: E402
from oracle_lib import Reference, pack_reads  # noqa: E402


def main():
    ref = Reference()
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    a = workloads.read_a()
    cases = {
        "kat_a_collapse_msat_k1001_w100": ("collapse-msat", 1001, 100, [a]),
        "kat_b_hpc_k1001_w100": ("hpc", 1001, 100, [a]),
        "kat_c_no_k1001_w100": ("no", 1001, 100, [a[:8000]]),
        "kat_d_collapse_msat_k501_w50": ("collapse-msat", 501, 50, [a]),
        "edge_cases_collapse_msat_k1001_w100": ("collapse-msat", 1001, 100, workloads.edge_case_reads()),
        "edge_cases_msat_k301_w30": ("msat", 301, 30, workloads.edge_case_reads()),
        "edge_cases_no_k63_w33": ("no", 63, 33, [r[:4000] for r in workloads.edge_case_reads()]),
        "edge_cases_collapse_dinuc_k255_w7": ("collapse-dinuc", 255, 7, [r[:6000] for r in workloads.edge_case_reads()]),
        "coverage_collapse_msat_k1001_w100": ("collapse-msat", 1001, 100, workloads.coverage_reads()),
        "coverage_dinuc_k501_w50": ("dinuc", 501, 50, workloads.coverage_reads(n_reads=6)),
        "repeats_collapse_msat_k63_w33": ("collapse-msat", 63, 33, [workloads.repeat_rich(3, 5000), workloads.repeat_rich(4, 700)]),
        "repeats_collapse_k31_w1": ("collapse", 31, 1, [workloads.repeat_rich(5, 3000)]),
    }
    for name, (mode, k, w, reads) in cases.items():
        bases, offsets = reads if isinstance(reads, tuple) else pack_reads(reads)
        r = ref.run(mode, k, w, (bases, offsets))
        np.savez_compressed(os.path.join(out, name + ".npz"), mode=mode, k=k, w=w, bases=bases, offsets=offsets, **{f: r[f] for f in (
            "parts", "codes", "poses", "positions", "hashes", "node_hash", "node_cov", "node_tip", "edges")})
        print(name, len(bases), "bases", len(r.parts), "parts", len(r.positions), "selected", len(r.node_cov), "nodes")


if __name__ == "__main__":
    main()
