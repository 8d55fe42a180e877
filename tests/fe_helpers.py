"""run the C-ABI front end (real CUDA library, or the GPU-less emulator build) and collect its arrays"""
from oracle_lib import RunResult
from verkko_b200.frontend import KEEP_CODES, KEEP_POSES, MbgFrontEnd


def run_frontend(mode, k, w, reads, library=None, flags=KEEP_CODES | KEEP_POSES, table=True, variants=None, **kw):
    fe = MbgFrontEnd(k, w, mode, flags=flags, library=library, **kw)
    try:
        if isinstance(reads, tuple):
            fe.add_reads(*reads)
        else:
            fe.add_read_list(reads)
        if variants is not None:
            fe.set_variants(variants.hashes, variants.offsets, variants.len_begin, variants.lengths)
        fe.run()
        res = RunResult(parts=fe.parts(), codes=fe.codes(), poses=fe.poses(), positions=fe.positions(), hashes=fe.hashes())
        if table:
            res.update(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
        res["stats"] = fe.stats()
        return res
    finally:
        fe.close()


def run_frontend_unitigs(mode, k, w, reads, min_cov, min_unitig_cov, library=None, one_cov=False, paths=False, variants=None, **kw):
    """table + mbgfe_build_unitigs -> (RunResult of the table, UnitigResult)"""
    from oracle_lib import UnitigResult
    fe = MbgFrontEnd(k, w, mode, library=library, **kw)
    try:
        if isinstance(reads, tuple):
            fe.add_reads(*reads)
        else:
            fe.add_read_list(reads)
        if variants is not None:
            fe.set_variants(variants.hashes, variants.offsets, variants.len_begin, variants.lengths)
        fe.run()
        table = RunResult(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
        u = UnitigResult(fe.build_unitigs(min_cov, min_unitig_cov, one_cov_heuristic=one_cov))
        u["stats"] = fe.stats()
        if paths:      # row f2: mbgfe_thread_reads, assembled into ReadPaths by the host mirror
            from oracle_lib import ReadPaths
            rp = fe.read_paths()
            u["paths"] = ReadPaths(heads=rp["heads"], path=rp["path"], poses=rp["poses"])
            u["thread_arrays"] = rp["arrays"]
            u["thread_stats"] = fe.stats()
        return table, u
    finally:
        fe.close()


def check_thinning(fe_result_unitigs, thin, offsets, want, k, what=""):
    """mbgfe_thin_unitigs against the oracle's thinned UnitigResult `want` (oracle.run_unitigs(..., thin=True)): the kept set,
    and for neighbours inside a thinned unitig the overlap k - (offset difference) must be the oracle's sequence overlap"""
    import numpy as np
    ug = fe_result_unitigs
    mine = np.zeros(len(ug["kept"]), np.uint8)
    mine[np.flatnonzero(ug["kept"])[thin == 1]] = 1
    assert np.array_equal(mine, want.kept), f"{what}: thinned k-mer set differs"
    new_id = np.cumsum(thin) - 1                       # filtered id -> thinned id (ids stay in hash order)
    ov = {}
    for f, ffw, t, tfw, o in want.overlaps.tolist():
        ov[(f, ffw, t, tfw)] = o
    b = ug["unitig_begin"].astype(np.int64)
    checked = 0
    for i in range(len(b) - 1):
        seq = [x for x in ug["unitig_nodes"][b[i]:b[i + 1]].tolist() if thin[x >> 1]]
        for x, y in zip(seq[:-1], seq[1:]):
            f, ffw, t, tfw = int(new_id[x >> 1]), x & 1, int(new_id[y >> 1]), y & 1
            if t < f or (t == f and not tfw and not ffw):
                f, ffw, t, tfw = t, 1 - tfw, f, 1 - ffw
            assert ov.get((f, ffw, t, tfw)) == k - (int(offsets[y >> 1]) - int(offsets[x >> 1])), f"{what}: overlap of neighbours {x}, {y} in unitig {i}"
            checked += 1
    return checked


def run_thinned_pass(mode, k, w, reads, min_cov, min_unitig_cov, oracle, library=None, one_cov=True, what=""):
    """runMBG's pass 3 + 4 under MBG's defaults on the device: table, getUnitigGraph, filterKmersToUnitigKmers APPLIED
    (MBGFE_T_*: reads.filter(kept), renumbered unitigs, new sequence overlaps, MBG/MBG.cpp:1065-1185) and getReadPaths on the
    thinned list (MBG/MBG.cpp:1586) -- against the oracle's sequential restatement of the same sequence (pinned on the
    reference in tests/test_oracle.py::test_unitig_kmer_thinning_equals_reference).  Returns the number of read paths compared."""
    import numpy as np
    from oracle_lib import ReadPaths, UnitigResult, compare_paths, compare_unitigs, sorted_edges
    fe = MbgFrontEnd(k, w, mode, library=library)
    try:
        if isinstance(reads, tuple):
            fe.add_reads(*reads)
        else:
            fe.add_read_list(reads)
        fe.run()
        fe.build_unitigs(min_cov, min_unitig_cov, one_cov_heuristic=one_cov)
        fe.thin_unitigs()
        th = fe.thinned()
        rp = fe.read_paths()
    finally:
        fe.close()
    _, want = oracle.run_unitigs(mode, k, w, reads, min_cov, min_unitig_cov, one_cov=one_cov, thin=True, paths=True)
    assert np.array_equal(th["node_hash"], want.node_hash) and np.array_equal(th["node_cov"], want.node_cov) and np.array_equal(th["node_tip"], want.node_tip), f"{what}: thinned list"
    assert np.array_equal(sorted_edges(th["edges"]), sorted_edges(want.edges)), f"{what}: edges of the thinned list"
    got = UnitigResult(kept=None, node_hash=th["node_hash"], node_cov=th["node_cov"], node_tip=th["node_tip"], edges=th["edges"],
                       unitig_begin=th["unitig_begin"], unitig_nodes=th["unitig_nodes"])
    compare_unitigs(got, UnitigResult({f: want[f] for f in ("kept", "node_hash", "node_cov", "node_tip", "edges", "unitig_begin", "unitig_nodes")}), what)
    # sequence overlaps: what the edges hold, then the new ones on top (HashList::addSequenceOverlap replaces)
    ov = {(int(r[0]), int(r[1]), int(r[2]), int(r[3])): int(r[5]) for r in th["edges"]}
    for r in th["new_overlaps"]:
        ov[(int(r[0]), int(r[1]), int(r[2]), int(r[3]))] = int(r[5])
    assert ov == {(int(r[0]), int(r[1]), int(r[2]), int(r[3])): int(r[4]) for r in want.overlaps}, f"{what}: sequence overlaps after the thinning"
    paths = ReadPaths(heads=rp["heads"], path=rp["path"], poses=rp["poses"])
    compare_paths(paths, got, want.paths, want, what)
    return len(want.paths.heads), int(len(th["new_overlaps"]))
