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
