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
