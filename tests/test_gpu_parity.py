"""Parity tests proper: the CUDA path, called through the C ABI, against the oracle and the committed golden
fixtures, plus size-independent properties at larger sizes.  Bit-exact: everything on this path is integer work."""
import glob
import os

import numpy as np
import pytest

import workloads
from fe_helpers import run_frontend
from oracle_lib import MODES, RunResult, compare_runs, lcg_bases
from verkko_b200.frontend import MbgFrontEnd

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_device_tables_known_answers(oracle):
    from verkko_b200 import MbgFrontEnd
    fe = MbgFrontEnd(1001, 100)
    h, c = fe.kat_tables()
    assert [int(x) for x in h[:4]] == [0x77cfa1eef01bca90, 0x5bca7c69b794f8ce, 0xb795033f6f2a0674, 0x135fddf6a6bfbbdd]
    assert all(int(c[i]) == oracle.complement(i) for i in range(30948))
    assert all(int(h[i]) == oracle.char_hash(i) for i in range(0, 30948, 13))
    codes = np.array([(i * 7) % 4 for i in range(31)], dtype=np.uint16)
    assert fe.kat_kmer_hash(codes) == (0xc8ec520a381594d1, 0x6515f3d84acb3010)
    rng = np.random.default_rng(3)
    for k in (31, 33, 35, 37, 1001, 2001):
        codes = rng.integers(0, 30948, k).astype(np.uint16)
        assert fe.kat_kmer_hash(codes) == oracle.kmer_hash(codes), k
    fe.close()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))))
def test_golden_fixtures(path):
    z = np.load(path)
    mode, k, w = str(z["mode"]), int(z["k"]), int(z["w"])
    got = run_frontend(mode, k, w, (z["bases"], z["offsets"]))
    want = RunResult({f: z[f] for f in ("parts", "codes", "poses", "positions", "hashes", "node_hash", "node_cov", "node_tip", "edges")})
    compare_runs(got, want, what=os.path.basename(path))


@pytest.mark.parametrize("mode", list(MODES))
def test_all_modes_against_oracle(oracle, mode):
    reads = [workloads.read_a()] + workloads.edge_case_reads() + [workloads.repeat_rich(9, 30000)]
    for k, w in ((1001, 100), (63, 33)):
        compare_runs(run_frontend(mode, k, w, reads), oracle.run(mode, k, w, reads), what=f"{mode} k={k} w={w}")


@pytest.mark.parametrize("k,w", [(31, 1), (501, 50), (1001, 17), (2001, 200), (3001, 100), (5001, 50), (255, 225)])
def test_k_w_sweep_against_oracle(oracle, k, w):
    reads = workloads.coverage_reads(seed=5, genome_len=60000, n_reads=10, len_min=15000, len_max=20000)
    compare_runs(run_frontend("collapse-msat", k, w, reads), oracle.run("collapse-msat", k, w, reads), what=f"k={k} w={w}")


def test_repeat_rich_reads(oracle):
    reads = [workloads.repeat_rich(s, n) for s, n in ((1, 50000), (2, 20000), (3, 4097), (4, 4096), (5, 8191), (6, 131)) ]
    for mode in ("collapse-msat", "msat", "dinuc"):
        compare_runs(run_frontend(mode, 101, 10, reads), oracle.run(mode, 101, 10, reads), what=mode)


def test_palindrome_rule(oracle):
    """MBG/ReadHelper.h:755-785: droppable palindromic k-mers vanish, a non-droppable one is an error (reference: abort)"""
    from verkko_b200 import MbgfeError
    reads = [workloads.repeat_rich(9, 30000), workloads.palindrome_read(1000)]
    compare_runs(run_frontend("dinuc", 63, 33, reads), oracle.run("dinuc", 63, 33, reads), what="droppable")
    with pytest.raises(RuntimeError, match="palindromic"):
        oracle.run("dinuc", 63, 33, [workloads.palindrome_read(1021)])
    with pytest.raises(MbgfeError) as e:
        run_frontend("dinuc", 63, 33, [workloads.palindrome_read(1021)])
    assert e.value.code == -4


def test_long_and_ragged_reads(oracle):
    from verkko_b200 import synth
    g = synth.genome(21, 400000)
    reads = synth.reads(3, g, 40, 30, 120000, 5000)   # 30 bases .. 120 kb, 0.5 % errors (HERRO/duplex-shaped)
    compare_runs(run_frontend("collapse-msat", 1001, 100, reads), oracle.run("collapse-msat", 1001, 100, reads), what="ragged")


def test_ecoli_shaped_coverage(oracle):
    """config #1 shape at reduced genome size: 24x coverage, 15 kb reads, full table parity"""
    from verkko_b200 import synth
    g = synth.genome(7, 400000)
    reads = synth.reads(11, g, 640, 11000, 19000, 1000)
    got = run_frontend("collapse-msat", 1001, 100, reads)
    want = oracle.run("collapse-msat", 1001, 100, reads)
    compare_runs(got, want, what="ecoli-shaped")
    assert got["stats"]["selected"] == int(want.node_cov.sum())


def test_empty_and_tiny_inputs():
    r = run_frontend("collapse-msat", 1001, 100, [])
    assert len(r.parts) == 0 and len(r.node_cov) == 0
    r = run_frontend("collapse-msat", 1001, 100, [b"ACGT" * 5, b"", b"N" * 100])
    assert len(r.parts) == 0 and len(r.positions) == 0


def test_properties_at_bench_scale():
    """size-independent properties on a chr20-shaped shard too large for the oracle: repeatability, coverage sums,
    tips, ordering"""
    from verkko_b200 import MbgFrontEnd, synth
    g = synth.genome(7, 8_000_000)
    bases, offsets = synth.reads(11, g, 13000, 15000, 20000, 1000)     # ~0.23 Gbp, 30x
    fe = MbgFrontEnd(1001, 100, "collapse-msat")
    fe.add_reads(bases, offsets)
    fe.run()
    st = fe.stats()
    cov, tip, nh, edges = fe.node_coverage(), fe.node_tip(), fe.node_hashes(), fe.edges()
    parts = fe.parts()
    assert int(cov.sum()) == st["selected"] == int(parts["n_sel"].sum())
    assert len(cov) == st["distinct"]
    order = np.lexsort((nh[:, 0], nh[:, 1]))
    assert np.array_equal(order, np.arange(len(nh))), "node table must be sorted by canonical hash"
    assert int(edges[:, 4].sum()) == st["selected"] - int((parts["n_sel"] > 0).sum())
    assert (edges[:, 5] >= 1).all() and (edges[:, 5] < 1001).all()
    assert tip.sum() <= 2 * (parts["n_sel"] > 0).sum()
    fe.process()                                                        # a second pass over the resident reads
    assert np.array_equal(cov, fe.node_coverage()) and np.array_equal(nh, fe.node_hashes()) and np.array_equal(edges, fe.edges())
    fe.close()


def test_file_input(tmp_path, oracle):
    import gzip
    reads = [workloads.read_a(), workloads.read_with_other_characters().lower(), workloads.read_a()[3000:9000]]
    fa = tmp_path / "reads.fa"
    with open(fa, "wb") as f:
        for i, r in enumerate(reads):
            f.write(b">read%d some description\r\n" % i)
            for j in range(0, len(r), 70):
                f.write(r[j:j + 70] + b"\n")
    fq = tmp_path / "reads.fastq.gz"
    with gzip.open(fq, "wb") as f:
        for i, r in enumerate(reads):
            f.write(b"@q%d\n" % i + r + b"\n+\n" + b"I" * len(r) + b"\n")
    from verkko_b200 import MbgFrontEnd
    want = oracle.run("collapse-msat", 1001, 100, [r.upper() for r in reads] * 2)
    fe = MbgFrontEnd(1001, 100, "collapse-msat")
    fe.add_file(str(fa))
    fe.add_file(str(fq))
    fe.run()
    assert fe.read_name(0) == "read0 some description" and fe.read_name(3) == "q0"
    assert np.array_equal(fe.hashes(), want.hashes) and np.array_equal(fe.node_coverage(), want.node_cov)
    fe.close()


@pytest.mark.gpu
def test_pipelined_run_in_small_pieces(oracle):
    """mbgfe_run with small H2D pieces (many pieces, tail launched per piece, sizes extrapolated from the first one)"""
    reads = workloads.coverage_reads(n_reads=24)
    got = run_frontend("collapse-msat", 1001, 100, reads, piece_bytes=20000)
    compare_runs(got, oracle.run("collapse-msat", 1001, 100, reads), what="small pieces")
    assert got["stats"]["table_grows"] == 0


@pytest.mark.gpu
def test_pipelined_run_estimate_overflow(oracle):
    """extrapolated sizes far too small (first piece collapses to almost nothing): refused writes are counted and the
    exact tail reruns; results are the reference's"""
    poor = (b"A" * 320 + b"C" * 320) * 20         # exactly one piece
    reads = [poor] + [workloads.repeat_rich(s, 9000) for s in (3, 4)] + [workloads.read_a()[:9000] for _ in range(6)]
    got = run_frontend("collapse", 31, 1, reads, piece_bytes=12800)
    compare_runs(got, oracle.run("collapse", 31, 1, reads), what="overflow fallback")
    assert got["stats"]["table_grows"] >= 1


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w", [("msat", 63, 33), ("collapse-msat", 31, 1), ("dinuc", 101, 20), ("collapse-msat", 1001, 100)])
def test_tandem_repeats_longer_than_a_tile(oracle, mode, k, w):
    reads = workloads.long_tandem_reads() + [workloads.repeat_rich(21, 20000), workloads.repeat_rich(22, 5000)]
    compare_runs(run_frontend(mode, k, w, reads), oracle.run(mode, k, w, reads), what=f"long tandem {mode}")


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w", [("collapse-msat", 1001, 100), ("collapse-msat", 301, 30), ("hpc", 63, 33), ("no", 101, 20)])
def test_hpc_variant_rehash(oracle, mode, k, w):
    """row a11: pass 3 of runMBG, hashes of k-mers with registered hpc variants are perturbed by their clustered run
    lengths (MBG/ReadHelper.h:680-737); the oracle's version is pinned on the reference's file-driven iterator in
    tests/test_oracle.py::test_hpc_variants_equal_reference"""
    from oracle_lib import make_variants
    reads = workloads.coverage_reads(n_reads=24) if k < 1001 else workloads.coverage_reads(genome_len=60000, n_reads=24, len_min=12000, len_max=20000)
    plain = oracle.run(mode, k, w, reads)
    variants = make_variants(plain, k, 300)
    got = run_frontend(mode, k, w, reads, variants=variants)
    want = oracle.run(mode, k, w, reads, variants=variants)
    compare_runs(got, want, what=f"variants {mode} k={k}")
    assert not np.array_equal(want.hashes, plain.hashes)
    # the library keeps poses by itself when variants are registered, and clearing the set restores the plain hashes
    fe = MbgFrontEnd(k, w, mode, flags=0)
    fe.add_reads(*workloads.as_pair(reads))
    fe.set_variants(variants.hashes, variants.offsets, variants.len_begin, variants.lengths)
    fe.run()
    assert np.array_equal(fe.hashes(), want.hashes) and np.array_equal(fe.node_hashes(), want.node_hash)
    fe.set_variants(np.zeros((0, 2), np.uint64), [], [0], [])
    fe.reset()
    fe.add_reads(*workloads.as_pair(reads))
    fe.run()
    assert np.array_equal(fe.hashes(), plain.hashes)
    fe.close()


def check_unitigs(got, want, what):
    from oracle_lib import compare_unitigs
    assert np.array_equal(got.kept, want.kept), f"{what}: kept flags differ"
    compare_unitigs(got, want, what)
    b = got.unitig_begin.astype(np.int64)
    ids = got.unitig_nodes >> 1
    assert len(np.unique(ids)) == len(ids), f"{what}: a node sits in two unitigs"
    of = got.node_unitig[ids]
    assert np.array_equal(of, np.repeat(np.arange(len(b) - 1), np.diff(b))), f"{what}: node -> unitig map"


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 1001, 100, 1, 2), ("collapse-msat", 301, 30, 1, 2), ("hpc", 63, 33, 2, 2), ("collapse-msat", 101, 20, 2, 3.5),
                                           ("no", 63, 5, 1, 1), ("hpc", 301, 30, 3, 2)])
def test_unitig_filter_against_oracle(oracle, mode, k, w, a, u):
    """rows a14 / f1: abundance filter (edges >= -a, chain mean >= -u), HashList::filter and the unitig chains, on the
    device, against the oracle (itself pinned on the reference's getUnitigGraph in tests/test_oracle.py)"""
    from fe_helpers import run_frontend_unitigs
    big = workloads.coverage_reads(genome_len=120000, n_reads=200, len_min=9000, len_max=14000, error_ppm=3000) if k < 1001 else \
        workloads.coverage_reads(genome_len=200000, n_reads=300, len_min=12000, len_max=20000, error_ppm=2000)
    for reads in (big, workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)]):
        table, got = run_frontend_unitigs(mode, k, w, reads, a, u)
        otable, want = oracle.run_unitigs(mode, k, w, reads, a, u)
        assert np.array_equal(table.node_hash, otable.node_hash)
        check_unitigs(got, want, f"unitigs {mode} k={k} a={a} u={u}")


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 1001, 100, 1, 2), ("collapse-msat", 301, 30, 1, 2), ("hpc", 63, 33, 1, 3), ("collapse-msat", 101, 20, 2, 2), ("no", 63, 5, 1, 2)])
def test_onecov_heuristic_against_oracle(oracle, mode, k, w, a, u):
    """row a14: getUnitigGraph with oneCovHeuristic (removeOnecovNodes twice, MBG/MBG.cpp:527-635) -- what runMBG builds
    its final graph with under Verkko's flags -- on the device, against the oracle (pinned on the reference's own
    getUnitigGraph(..., true) in tests/test_oracle.py)"""
    from fe_helpers import run_frontend_unitigs
    big = workloads.coverage_reads(genome_len=120000, n_reads=200, len_min=9000, len_max=14000, error_ppm=3000) if k < 1001 else \
        workloads.coverage_reads(genome_len=200000, n_reads=300, len_min=12000, len_max=20000, error_ppm=2000)
    changed = 0
    for reads in (big, workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)],
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        table, got = run_frontend_unitigs(mode, k, w, reads, a, u, one_cov=True)
        otable, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=True)
        assert np.array_equal(table.node_hash, otable.node_hash)
        check_unitigs(got, want, f"onecov unitigs {mode} k={k} a={a} u={u}")
        _, plain = oracle.run_unitigs(mode, k, w, reads, a, u)
        changed += int(not np.array_equal(plain.kept, want.kept))
    assert changed > 0 or (mode, k) != ("collapse-msat", 301), "the heuristic never changed the kept set"


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w,a,u,one_cov", [("collapse-msat", 1001, 100, 1, 2, True), ("collapse-msat", 301, 30, 1, 2, True), ("hpc", 301, 30, 2, 2, False), ("hpc", 63, 33, 1, 3, True),
                                                  ("collapse-msat", 101, 20, 2, 2, False), ("no", 63, 5, 1, 1, False)])
def test_read_threading_against_oracle(oracle, mode, k, w, a, u, one_cov):
    """row f2: mbgfe_thread_reads -- getKmerLocator + iterateReadPaths (MBG/KmerMatcher.cpp:3-31, KmerMatcher.h:12-144)
    as a batched lookup over the resident k-mer instances -- against the oracle's sequential restatement (pinned on the
    reference's own functions in tests/test_oracle.py); compared as the oriented k-mers and read positions every path covers"""
    from fe_helpers import run_frontend_unitigs
    from oracle_lib import compare_paths
    big = workloads.coverage_reads(genome_len=60000, n_reads=100, len_min=9000, len_max=14000, error_ppm=3000) if k < 1001 else \
        workloads.coverage_reads(genome_len=100000, n_reads=150, len_min=12000, len_max=20000, error_ppm=2000)
    n_paths = 0
    for reads in (big, workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)],
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        _, got = run_frontend_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
        _, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
        compare_paths(got.paths, got, want.paths, want, f"read paths {mode} k={k} a={a} u={u}")
        n_paths += len(want.paths.heads)
        th = got.thread_arrays
        located = th["unitig"] != 0xffffffff
        assert np.all(th["step"][~located] <= 1) and np.all(th["step"][located] >= 2)
    assert n_paths > 20


@pytest.mark.gpu
def test_read_threading_properties_at_scale():
    """size-independent properties of mbgfe_thread_reads on a larger graph (2 Mbp, 3000 reads): every located instance
    names the unitig slot that holds exactly its k-mer; position + rest + 1 is the unitig's size; a step-4 instance
    follows its predecessor inside the same oriented unitig; every path start is a located k-mer"""
    from verkko_b200.frontend import MbgFrontEnd
    reads = workloads.coverage_reads(genome_len=2000000, n_reads=3000, len_min=15000, len_max=20000, error_ppm=1000)
    fe = MbgFrontEnd(1001, 100, "collapse-msat")
    try:
        fe.add_reads(*reads) if isinstance(reads, tuple) else fe.add_read_list(reads)
        fe.run()
        hashes = fe.hashes().copy()
        u = fe.build_unitigs(1, 2, one_cov_heuristic=True)
        th = fe.thread_reads()
    finally:
        fe.close()
    unitig, upos, urest, step = (th[x].astype(np.int64) for x in ("unitig", "unitig_pos", "unitig_rest", "step"))
    located = unitig != 0xffffffff
    assert located.sum() > 100000 and np.all(step[located] >= 2) and np.all(step[~located] <= 1)
    b = u["unitig_begin"].astype(np.int64)
    size = np.diff(b)
    ut, o = unitig[located] >> 1, unitig[located] & 1
    assert np.all(upos[located] + urest[located] + 1 == size[ut])
    slot = b[ut] + np.where(o == 1, upos[located], size[ut] - 1 - upos[located])
    node = u["unitig_nodes"].astype(np.int64)[slot]
    # the k-mer: canonical hash = the smaller of (hi, lo) and its halves swapped (MBG/HashList.cpp:85-101)
    lo, hi = hashes[located, 0], hashes[located, 1]
    fw = (hi < lo) | ((hi == lo) & (lo < hi))
    canon = np.where(fw[:, None], np.stack([lo, hi], 1), np.stack([hi, lo], 1))
    assert np.array_equal(u["node_hash"][node >> 1], canon), "a located instance points at a unitig slot that holds another k-mer"
    assert np.array_equal((node & 1) == o, fw), "orientation of a located instance inside its unitig"
    cont = np.flatnonzero(step == 4)
    prev = cont - 1
    while True:                       # the previous instance that is in the filtered list (step != 0); never leaves the part
        skip = step[prev] == 0
        if not skip.any():
            break
        prev[skip] -= 1
    assert np.array_equal(unitig[cont], unitig[prev]) and np.array_equal(upos[cont], upos[prev] + 1)
    cross = np.flatnonzero(step == 5)
    assert len(cross) > 0 and np.all(upos[cross] == 0)


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w", [("collapse-msat", 301, 30), ("hpc", 63, 33)])
def test_pass3_variants_unitigs_and_read_paths(oracle, mode, k, w):
    """pass 3 of runMBG end to end on the device: hpc variants registered (row a11), table, getUnitigGraph with
    oneCovHeuristic (a14 / f1) and read threading (f2) -- the calls of MBG/MBG.cpp:1570-1590 -- against the oracle"""
    from fe_helpers import run_frontend_unitigs
    from oracle_lib import compare_paths, make_variants
    reads = workloads.coverage_reads(genome_len=60000, n_reads=100, len_min=9000, len_max=14000, error_ppm=3000)
    variants = make_variants(oracle.run(mode, k, w, reads), k, 300)
    table, got = run_frontend_unitigs(mode, k, w, reads, 1, 2, one_cov=True, paths=True, variants=variants)
    otable, want = oracle.run_unitigs(mode, k, w, reads, 1, 2, variants=variants, one_cov=True, paths=True)
    assert np.array_equal(table.node_hash, otable.node_hash)
    check_unitigs(got, want, f"pass 3 unitigs {mode} k={k}")
    compare_paths(got.paths, got, want.paths, want, f"pass 3 read paths {mode} k={k}")
    _, plain = oracle.run_unitigs(mode, k, w, reads, 1, 2, one_cov=True)
    assert not np.array_equal(plain.node_hash, want.node_hash), "the variants did not change any hash"


@pytest.mark.gpu
@pytest.mark.parametrize("mode,k,w,a,u", [("no", 63, 5, 1, 2), ("hpc", 63, 33, 1, 1), ("collapse-msat", 101, 20, 2, 2)])
def test_unitig_filter_circular_genomes(oracle, mode, k, w, a, u):
    """chains that are cycles (no tip, no branch): the ranking detects them by its round limit, cuts them in front of
    their smallest node and ranks again; the reference starts them at that node too (MBG/MBG.cpp:682-690, 746-760)"""
    from fe_helpers import run_frontend_unitigs
    reads = workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]
    _, got = run_frontend_unitigs(mode, k, w, reads, a, u)
    _, want = oracle.run_unitigs(mode, k, w, reads, a, u)
    check_unitigs(got, want, f"circular {mode} k={k}")
    # circular unitigs start at their smallest node in forward orientation
    b = got.unitig_begin.astype(np.int64)
    for i in range(len(b) - 1):
        seq = got.unitig_nodes[b[i]:b[i + 1]]
        assert (seq[0] >> 1) == (seq >> 1).min() or (seq[-1] >> 1) > (seq[0] >> 1)


@pytest.mark.gpu
def test_unitig_filter_properties_at_scale():
    """size-independent properties on a larger graph: kept nodes are exactly those of chains with mean coverage >= u;
    unitigs partition the kept nodes; consecutive unitig nodes are joined by a covered edge"""
    from fe_helpers import run_frontend_unitigs
    reads = workloads.coverage_reads(genome_len=2000000, n_reads=3000, len_min=15000, len_max=20000, error_ppm=1000)
    table, got = run_frontend_unitigs("collapse-msat", 1001, 100, reads, 1, 2)
    assert len(got.kept) == len(table.node_cov) and got.kept.sum() == len(got.node_cov)
    assert np.array_equal(got.node_hash, table.node_hash[got.kept == 1]) and np.array_equal(got.node_cov, table.node_cov[got.kept == 1])
    b = got.unitig_begin.astype(np.int64)
    assert b[-1] == len(got.unitig_nodes) == len(got.node_cov)
    assert np.array_equal(np.sort(got.unitig_nodes >> 1), np.arange(len(got.node_cov)))
    e = got.edges.astype(np.int64)
    links = set(zip((2 * e[:, 0] + e[:, 1]).tolist(), (2 * e[:, 2] + e[:, 3]).tolist())) | set(zip(((2 * e[:, 2] + e[:, 3]) ^ 1).tolist(), ((2 * e[:, 0] + e[:, 1]) ^ 1).tolist()))
    seq = got.unitig_nodes.astype(np.int64)
    inner = np.ones(len(seq), bool)
    inner[b[1:-1] - 1] = False
    inner[-1] = False
    pairs = list(zip(seq[:-1][inner[:-1]].tolist(), seq[1:][inner[:-1]].tolist()))
    assert all(p in links for p in pairs)
    # mean coverage of every kept chain is >= 2: check through unitig means being >= 2 is NOT implied (unitigs of the
    # filtered list can be shorter), but every dropped node must have coverage-1 neighbours only in dropped chains
    assert got.stats["ms_unitigs"] > 0


@pytest.mark.gpu
@pytest.mark.parametrize("pct", ["100", "70", "0"])
def test_host_packed_upload_pieces(oracle, monkeypatch, pct):
    """the two upload routes of mbgfe_run mixed piece by piece (ASCII + K0 on the device / 2-bit packing on host threads),
    with lower-case reads and characters outside ACGT inside host-packed pieces"""
    monkeypatch.setenv("MBGFE_HOST_PACK", pct)
    reads = [workloads.read_a()] + workloads.edge_case_reads()
    reads += [lcg_bases(5, 9000).replace(b"ACGTA", b"ACNNA", 3), lcg_bases(6, 7000).lower(), lcg_bases(8, 4097)]
    for mode, k, w in (("collapse-msat", 301, 30), ("hpc", 63, 33), ("no", 101, 20)):
        got = run_frontend(mode, k, w, reads, piece_bytes=3200)
        compare_runs(got, oracle.run(mode, k, w, reads), what=f"host pack {pct}% {mode}")
    big = workloads.coverage_reads(genome_len=400000, n_reads=600, len_min=12000, len_max=20000)
    compare_runs(run_frontend("collapse-msat", 1001, 100, big, piece_bytes=1 << 20), oracle.run("collapse-msat", 1001, 100, big), what=f"host pack {pct}% big")


def test_select_shapes_on_gpu(oracle):
    """K3a (rolling form) over every residue of s mod 8, windows of 1 .. 32 blocks and every r -- the shapes of
    tests/test_emu_kernels.py::test_emulated_select_shapes on the real library"""
    from test_emu_kernels import SELECT_SHAPES
    from fe_helpers import run_frontend
    from oracle_lib import compare_runs
    import workloads
    a = workloads.read_a()
    for k, w in SELECT_SHAPES:
        reads = [a[:3000 + 7 * i] for i in range(10)] + [workloads.repeat_rich(k + w, 4000), workloads.long_tandem_reads()[0][:6000], a[:k + 40], a[:k], a[:k - 1]]
        compare_runs(run_frontend("collapse-msat", k, w, reads), oracle.run("collapse-msat", k, w, reads), what=f"select shape k={k} w={w}")


def test_random_inputs_on_gpu(oracle):
    """the seeded random configurations of tests/test_emu_kernels.py::test_emulated_random_inputs, and more of them, on the real library"""
    for seed in range(48):
        mode, k, w, reads = workloads.random_case(seed)
        compare_runs(run_frontend(mode, k, w, reads), oracle.run(mode, k, w, reads), what=f"random seed={seed} {mode} k={k} w={w}")


def test_graph_stage_random_on_gpu(oracle):
    """rows a14 / f1 / f2 on the seeded random read sets of tests/test_emu_kernels.py::test_emulated_graph_stage_random (and as many
    again), on the real library: kept flags, unitigs, unitig edges, read paths"""
    from fe_helpers import run_frontend_unitigs
    from oracle_lib import compare_paths, compare_unitigs
    for seed in range(24):
        mode, k, w, reads, a, u, one_cov = workloads.random_graph_case(seed)
        _, got = run_frontend_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
        _, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
        assert np.array_equal(got.kept, want.kept), f"seed {seed}: kept flags differ"
        compare_unitigs(got, want, f"random graph stage seed={seed} {mode} k={k} w={w} a={a} u={u}")
        compare_paths(got.paths, got, want.paths, want, f"random read paths seed={seed}")

