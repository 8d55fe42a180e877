"""CPU tests of the checker itself: known answers, the reference compiled here, committed golden fixtures."""
import glob
import os

import numpy as np
import pytest

import workloads
from oracle_lib import MODES, RunResult, compare_runs, lcg_bases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_tables_known_answers(oracle):
    # SURVEY.md 8(c) "tables" row, minted from the reference
    assert oracle.lib.orc_max_code() == 30948
    assert [oracle.char_hash(c) for c in range(4)] == [0x77cfa1eef01bca90, 0x5bca7c69b794f8ce, 0xb795033f6f2a0674, 0x135fddf6a6bfbbdd]
    for a, b in ((0, 3), (1, 2), (4, 34), (5, 35), (6, 26), (36, 225), (228, 1248), (1252, 6367), (6372, 30942), (30947, 6377)):
        assert oracle.complement(a) == b
    for c in range(0, 30948, 7):
        assert oracle.complement(oracle.complement(c)) == c
    assert [oracle.lib.orc_code_motif_length(c) for c in (4, 36, 228, 1252, 6372)] == [2, 3, 4, 5, 6]


def test_hash_known_answers(oracle):
    b = lcg_bases(7, 902)
    codes = np.array([b"ACGT".index(x) for x in b], dtype=np.uint16)
    assert oracle.rolling_hash(902, codes) == (0x7868dd6946eb7eeb, 0xac082fb93b68b8f2)
    codes = np.array([(i * 7) % 4 for i in range(31)], dtype=np.uint16)
    assert oracle.kmer_hash(codes) == (0xc8ec520a381594d1, 0x6515f3d84acb3010)


def test_msat_known_answer(oracle):
    s = "ACGTACGATCAGCTAGCTAGCATCGATCGACTACACACACACACACACAGTCAGTCATGATGATGATGATGATCAGCTAGCATCGAGCATCGATCGACTAGC"
    sym = np.array(["ACGT".index(ch) for ch in s], dtype=np.uint16)
    want = [0, 1, 2, 3, 0, 1, 2, 0, 3, 1, 387, 445, 1, 3, 7, 1, 0, 951, 0, 205, 1, 0, 2, 1, 3, 0, 2, 1, 0, 3, 1, 2, 0, 2, 1, 445, 1, 3, 0, 2, 1]
    codes, poses = oracle.msat(sym)
    assert codes.tolist() == want
    assert poses.tolist()[:21] == [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 21, 30, 31, 32, 47, 48, 49, 56, 57, 73]
    assert poses.tolist()[-6:] == [97, 98, 99, 100, 101, 102]
    c2, p2 = oracle.msat_scanform(sym)
    assert c2.tolist() == want and p2.tolist() == poses.tolist()


KAT = {  # (codes, selected, first positions, last position, h[0] hi:lo, xor hi:lo)
    ("collapse-msat", 1001, 100): (9861, 180, [3, 69, 137], 8784, (0x37b28bbafe4ca20e, 0xb05b9c088ddff306), (0x62db18cc977053eb, 0x646835421a56e8fb)),
    ("hpc", 1001, 100): (14906, 283, [16, 115, 129], 13899, (0xa44ec27ba0fa75e8, 0xeb8e6c3091dc1055), (0xcb6fa6b01ee5b63c, 0x7e4a10eae7660319)),
    ("no", 1001, 100): (20000, 381, [67, 71, 170], 18934, (0x7675d69525c130df, 0xc8805d6414fdc0cc), (0x19cfb4d85d721e61, 0x09086f49deabb55d)),
    ("collapse-msat", 501, 50): (9861, 363, [22, 27, 76], 9335, (0x0dd66565f7f7a450, 0x78de3737236c25db), (0xe4266bf336b7323d, 0x9308f9cdd482ffe1)),
}


@pytest.mark.parametrize("key", list(KAT))
def test_read_known_answers(oracle, key):
    mode, k, w = key
    n_codes, n_sel, first, last, h0, hx = KAT[key]
    r = oracle.run(mode, k, w, [workloads.read_a()])
    assert len(r.codes) == n_codes and len(r.positions) == n_sel
    assert r.positions[:3].tolist() == first and int(r.positions[-1]) == last
    assert (int(r.hashes[0, 1]), int(r.hashes[0, 0])) == h0
    x = np.bitwise_xor.reduce(r.hashes, axis=0)
    assert (int(x[1]), int(x[0])) == hx


def test_split_at_n_known_answer(oracle):
    rb = bytearray(workloads.read_a())
    rb[5000:5002] = b"NN"
    r = oracle.run("collapse-msat", 1001, 100, [bytes(rb)])
    assert r.parts["part_offset"].tolist() == [0, 3761]
    assert r.parts["n_codes"].tolist() == [2512, 7349]
    assert r.parts["n_sel"].tolist() == [36, 122]
    assert int(r.poses[2512]) == 3760 and int(r.poses[-1]) == 14907


@pytest.mark.parametrize("mode", list(MODES))
def test_oracle_equals_reference(oracle, reference, mode):
    for k, w, reads in ((1001, 100, [workloads.read_a()] + workloads.edge_case_reads()), (301, 30, workloads.coverage_reads()),
                        (63, 33, [workloads.repeat_rich(3, 5000), workloads.repeat_rich(4, 700)])):
        compare_runs(oracle.run(mode, k, w, reads), reference.run(mode, k, w, reads), what=f"{mode} k={k} w={w}")


def test_msat_scanform_equals_reference_algorithm(oracle):
    """the scan-parallel formulation the CUDA kernel uses == the sequential reference algorithm (on run-length input)"""
    rng = np.random.default_rng(5)
    for it in range(400):
        s = np.frombuffer(workloads.repeat_rich(100 + it, int(rng.integers(40, 600))), dtype=np.uint8)
        sym = np.searchsorted(np.frombuffer(b"ACGT", dtype=np.uint8), s).astype(np.uint16)
        keep = np.ones(len(sym), bool)
        keep[1:] = sym[1:] != sym[:-1]
        sym = sym[keep]
        if len(sym) < 32:
            continue
        for mm in (2, 6):
            a, b = oracle.msat(sym, max_motif=mm), oracle.msat_scanform(sym, max_motif=mm)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (it, mm)


def test_syncmer_rule_bruteforce(oracle):
    r = oracle.run("collapse-msat", 201, 20, [workloads.read_a()[:6000]], table=False)
    p = r.parts[0]
    codes = r.codes[int(p["code_begin"]):int(p["code_begin"]) + int(p["n_codes"])]
    assert np.array_equal(oracle.syncmers_bruteforce(codes, 201, 20), r.positions)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))))
def test_golden_fixtures(oracle, path):
    """fixtures written by tools/make_golden.py from the reference compiled in the build container"""
    z = np.load(path)
    mode, k, w = str(z["mode"]), int(z["k"]), int(z["w"])
    got = oracle.run(mode, k, w, (z["bases"], z["offsets"]))
    want = RunResult({f: z[f] for f in ("parts", "codes", "poses", "positions", "hashes", "node_hash", "node_cov", "node_tip", "edges")})
    compare_runs(got, want, what=os.path.basename(path))


def compare_hashed(a, b, what):
    """a: RunResult with every part; b: hashed parts only (the file-driven reference iterator), same order"""
    pa = a.parts[a.parts["hashed"] == 1]
    assert len(pa) == len(b.parts), f"{what}: hashed part count"
    for f in ("read_idx", "part_offset", "n_codes", "n_sel"):
        assert np.array_equal(pa[f], b.parts[f]), f"{what}: parts.{f}"
    assert np.array_equal(a.positions, b.positions), f"{what}: positions"
    assert np.array_equal(a.hashes, b.hashes), f"{what}: hashes"
    assert np.array_equal(a.node_hash, b.node_hash) and np.array_equal(a.node_cov, b.node_cov) and np.array_equal(a.node_tip, b.node_tip), f"{what}: nodes"
    from oracle_lib import sorted_edges
    assert np.array_equal(sorted_edges(a.edges), sorted_edges(b.edges)), f"{what}: edges"


@pytest.mark.parametrize("mode,k,w", [("collapse-msat", 301, 30), ("hpc", 301, 30), ("hpc", 63, 33), ("no", 101, 20)])
def test_hpc_variants_equal_reference(oracle, reference, tmp_path, mode, k, w):
    """row a11: the pass-3 re-hash with registered hpc variants (MBG/ReadHelper.h:680-737) against the reference's
    file-driven iterator; both strands of a k-mer must still meet in one node"""
    from oracle_lib import make_variants
    reads = workloads.coverage_reads(n_reads=16)
    plain = oracle.run(mode, k, w, reads)
    variants = make_variants(plain, k, 200)
    assert len(variants.offsets) > 100
    got = oracle.run(mode, k, w, reads, variants=variants)
    want = reference.run_variants(mode, k, w, reads, variants, tmp_path / "reads.fa")
    compare_hashed(got, want, f"variants {mode} k={k}")
    assert not np.array_equal(got.hashes, plain.hashes)
    assert np.array_equal(got.positions, plain.positions)


@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 301, 30, 1, 2), ("hpc", 301, 30, 2, 2), ("hpc", 63, 33, 1, 3), ("collapse-msat", 101, 20, 2, 3.5), ("no", 63, 5, 1, 1)])
def test_unitig_filter_equals_reference(oracle, reference, mode, k, w, a, u):
    """rows a14 / f1: getUnitigGraph's abundance filter, HashList::filter and the unitig chains against the reference's
    own getUnitigGraph (MBG/MBG.cpp:629-760) on the same reads"""
    from oracle_lib import compare_unitigs
    for reads in (workloads.coverage_reads(n_reads=40, error_ppm=4000), workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)],
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        table, got = oracle.run_unitigs(mode, k, w, reads, a, u)
        want = reference.run_unitigs(mode, k, w, reads, a, u)
        compare_unitigs(got, want, what=f"unitigs {mode} k={k} a={a} u={u}")
        assert got.kept.sum() == len(got.node_cov) and len(got.kept) == len(table.node_cov)


@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 301, 30, 1, 2), ("hpc", 301, 30, 1, 2), ("hpc", 63, 33, 1, 3), ("collapse-msat", 101, 20, 2, 2), ("no", 63, 5, 1, 2)])
def test_onecov_heuristic_equals_reference(oracle, reference, mode, k, w, a, u):
    """row a14: removeOnecovNodes (MBG/MBG.cpp:527-627), the two passes getUnitigGraph runs first under Verkko's flags
    (oneCovHeuristic, MBG.cpp:631-635, 1573), against the reference's own getUnitigGraph"""
    from oracle_lib import compare_unitigs
    changed = 0
    for reads in (workloads.coverage_reads(n_reads=40, error_ppm=4000), workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  workloads.coverage_reads(seed=5, genome_len=30000, n_reads=60, error_ppm=2000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)],
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        table, got = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=True)
        want = reference.run_unitigs(mode, k, w, reads, a, u, one_cov=True)
        compare_unitigs(got, want, what=f"onecov unitigs {mode} k={k} a={a} u={u}")
        _, plain = oracle.run_unitigs(mode, k, w, reads, a, u)
        changed += int(not np.array_equal(plain.kept, got.kept))
    assert changed > 0 or a > 1, "the heuristic never changed the kept set: the workloads do not exercise it"


@pytest.mark.parametrize("mode,k,w,a,u,one_cov", [("collapse-msat", 301, 30, 1, 2, True), ("hpc", 301, 30, 2, 2, False), ("hpc", 63, 33, 1, 3, True), ("collapse-msat", 101, 20, 2, 2, False),
                                                  ("no", 63, 5, 1, 2, False), ("no", 63, 5, 1, 1, True)])
def test_read_paths_equal_reference(oracle, reference, mode, k, w, a, u, one_cov):
    """row f2: getKmerLocator + iterateReadPaths (MBG/KmerMatcher.cpp:3-31, KmerMatcher.h:12-144) of every part through
    the unitig graph, against the reference's own functions on its own graph; compared as the oriented k-mers and read
    positions every path covers (unitig numbering drops out)"""
    from oracle_lib import compare_paths
    n_paths = 0
    for reads in (workloads.coverage_reads(n_reads=40, error_ppm=4000), workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  [workloads.repeat_rich(3, 6000), workloads.repeat_rich(3, 6000), workloads.repeat_rich(4, 3000)],
                  workloads.circular_reads() + workloads.circular_reads(seed=10, genome_len=2000, read_len=1500, step=250) + [lcg_bases(77, 1500)]):
        _, got = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
        want = reference.run_unitigs(mode, k, w, reads, a, u, one_cov=one_cov)
        compare_paths(got.paths, got, want.paths, want, what=f"read paths {mode} k={k} a={a} u={u}")
        n_paths += len(got.paths.heads)
    assert n_paths > 20


@pytest.mark.parametrize("mode,k,w,a,u", [("collapse-msat", 301, 30, 1, 2), ("hpc", 63, 33, 1, 3), ("collapse-msat", 101, 20, 1, 2), ("no", 63, 5, 1, 2)])
def test_unitig_kmer_thinning_equals_reference(oracle, reference, mode, k, w, a, u):
    """the rest of runMBG's pass 3 (MBG/MBG.cpp:1575-1581) under MBG's defaults: getUnitigs(filterUnitigsByCoverage),
    filterKmersToUnitigKmers with filterWithinUnitig = true, sortKmersByHashes, getReadPaths -- the oracle's restatement of
    the thinning (orc_thin_unitigs) and its read paths on the thinned list against the reference's own calls"""
    from oracle_lib import compare_paths, compare_unitigs
    thinned = 0
    for reads in (workloads.coverage_reads(n_reads=40, error_ppm=4000), workloads.coverage_reads(seed=3, genome_len=20000, n_reads=12, error_ppm=10000),
                  workloads.coverage_reads(seed=5, genome_len=30000, n_reads=60, error_ppm=2000)):
        _, full = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=True)
        _, got = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=True, paths=True, thin=True)
        want = reference.run_unitigs(mode, k, w, reads, a, u, one_cov=True, pass3=True)
        compare_unitigs(got, want, what=f"thinned unitigs {mode} k={k}")
        # ids are in hash order on both sides after sortKmersByHashes: the overlaps compare directly
        rows = lambda x: x[np.lexsort(x.T[::-1])]
        assert np.array_equal(rows(got.overlaps), rows(want.overlaps)), "sequence overlaps after the thinning"
        compare_paths(got.paths, got, want.paths, want, what=f"read paths on the thinned list {mode} k={k}")
        thinned += len(full.node_cov) - len(got.node_cov)
    assert thinned > 0, "nothing was thinned"


def test_table_digest_equals_reference_hashlist_digest(oracle, reference, tmp_path):
    """the order-free table digest (bench.py `parity_sample`, tests/test_gpu_configs.py): computed from exported arrays
    (oracle table here) it equals what oracle/ref_shim.cpp computes straight from the reference's HashList built from a
    FASTA file -- with one worker thread all ten words, with four all but the two the reference leaves to a race
    (tips) or to arrival order (overlap of a pair seen with two spacings, H6)"""
    from oracle_lib import reference_file_digest, table_digest, write_fasta
    bases, offsets = workloads.as_pair(workloads.coverage_reads(genome_len=60000, n_reads=60, len_min=9000, len_max=14000, error_ppm=3000))
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    for mode, k, w in (("collapse-msat", 501, 50), ("hpc", 63, 33)):
        run = oracle.run(mode, k, w, (bases, offsets))
        mine = table_digest(run.node_hash, run.node_cov, run.node_tip, run.edges)
        _, ref1 = reference_file_digest(mode, k, w, fa, 1)
        assert mine == ref1, (mode, mine, ref1)
        _, ref4 = reference_file_digest(mode, k, w, fa, 4)
        for f in ("selected", "distinct", "distinct_edges", "node_xor", "node_sum", "edge_cov", "edge_sum", "edge_xor"):
            assert mine[f] == ref4[f], (mode, f)
    # and it notices a single changed coverage or a flipped edge orientation
    cov = run.node_cov.copy(); cov[3] += 1
    assert table_digest(run.node_hash, cov, run.node_tip, run.edges)["node_xor"] != mine["node_xor"]
    e = run.edges.copy(); e[0, 1] ^= 1
    assert table_digest(run.node_hash, run.node_cov, run.node_tip, e)["edge_sum"] != mine["edge_sum"]


@pytest.mark.parametrize("r,big_r,guesswork", [(0, 0, False), (1500, 400, True)])
def test_t4_backend_from_tables_gives_the_reference_gfa(oracle, reference, tmp_path, r, big_r, guesswork):
    """T4 on the checker's side: tables in the library's result layout -- here the oracle's, the GPU test hands
    over the device's -- put into the reference's own HashList / UnitigGraph (oracle/ref_shim.cpp:refshim_backend) and the rest
    of runMBG run unchanged give the GFA and GAF of `MBG -t 1` byte for byte.  Stage 0 replaces loadReadsAsHashesMultithread,
    stage 1 everything up to sortKmersByHashes."""
    import hashlib
    from oracle_lib import reference_backend, reference_cli, write_fasta
    bases, offsets = workloads.as_pair(workloads.coverage_reads(genome_len=40000, n_reads=60, len_min=6000, len_max=10000, error_ppm=2000))
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    mode, k, w, a, u = "collapse-msat", 301, 30, 1, 2
    want_gfa, want_gaf = reference_cli(fa, tmp_path / "ref.gfa", tmp_path / "ref.gaf", mode, k, w, a, u, r, big_r, guesswork)
    assert want_gfa.count(b"\nS\t") + want_gfa.startswith(b"S\t") > 0 and len(want_gaf) > 0
    table = oracle.run(mode, k, w, (bases, offsets))
    gfa0, gaf0 = reference_backend(0, fa, tmp_path / "s0.gfa", tmp_path / "s0.gaf", mode, k, w, a, u, r, big_r, guesswork,
                                   node_hash=table.node_hash, node_cov=table.node_cov, node_tip=table.node_tip, edges=table.edges)
    assert hashlib.md5(gfa0).hexdigest() == hashlib.md5(want_gfa).hexdigest(), "stage 0 GFA"
    assert hashlib.md5(gaf0).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 0 GAF"
    one_cov = u >= 2 and r > 0 and guesswork
    from oracle_lib import sort_unitigs
    _, th = oracle.run_unitigs(mode, k, w, (bases, offsets), a, u, one_cov=one_cov, thin=True)
    th = sort_unitigs(th)                   # the oracle numbers unitigs in walk order, the library like UnitigGraph::sort
    ov = np.zeros((len(th.overlaps), 6), np.uint32)
    ov[:, :4], ov[:, 5] = th.overlaps[:, :4], th.overlaps[:, 4]
    gfa1, gaf1 = reference_backend(1, fa, tmp_path / "s1.gfa", tmp_path / "s1.gaf", mode, k, w, a, u, r, big_r, guesswork,
                                   node_hash=th.node_hash, node_cov=th.node_cov, node_tip=th.node_tip, edges=th.edges, new_overlaps=ov,
                                   unitig_begin=th.unitig_begin, unitig_nodes=th.unitig_nodes, unitig_edges=th.unitig_edges)
    assert hashlib.md5(gfa1).hexdigest() == hashlib.md5(want_gfa).hexdigest(), "stage 1 GFA"
    assert hashlib.md5(gaf1).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 1 GAF"


def test_t4_backend_with_hpc_variants_verkko_flags(oracle, reference, tmp_path):
    """T4 under Verkko's flag SET (--hpc-variant-onecopy-coverage on, resolution on, guesswork, -a 1 -u 2, collapse-msat): the
    variant table comes from the reference's own passes 1-2 (row f3 is not built), pass 3 tables with the variants applied -- here
    the oracle's, the GPU test hands over the device's -- go through the reference's back end: GFA and GAF of MBG -t 1"""
    import hashlib
    from oracle_lib import reference_backend, reference_cli, reference_hpc_variants, write_fasta
    bases, offsets = workloads.diploid_msat_reads()
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    mode, k, w, a, u, r, big_r, onecopy = "collapse-msat", 301, 30, 1, 2, 4000, 1000, 8
    want_gfa, want_gaf = reference_cli(fa, tmp_path / "ref.gfa", tmp_path / "ref.gaf", mode, k, w, a, u, r, big_r, True, extra=("--hpc-variant-onecopy-coverage", str(onecopy)))
    variants = reference_hpc_variants(fa, mode, k, w, a, u, onecopy)
    assert len(variants.offsets) > 10, "the workload must produce hpc variants"
    table = oracle.run(mode, k, w, (bases, offsets), variants=variants)
    plain = oracle.run(mode, k, w, (bases, offsets))
    assert not np.array_equal(table.node_hash, plain.node_hash), "the variants must change hashes"
    gfa0, gaf0 = reference_backend(0, fa, tmp_path / "s0.gfa", tmp_path / "s0.gaf", mode, k, w, a, u, r, big_r, True, variants=variants,
                                   node_hash=table.node_hash, node_cov=table.node_cov, node_tip=table.node_tip, edges=table.edges)
    assert hashlib.md5(gfa0).hexdigest() == hashlib.md5(want_gfa).hexdigest() and hashlib.md5(gaf0).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 0"
    from oracle_lib import sort_unitigs
    _, th = oracle.run_unitigs(mode, k, w, (bases, offsets), a, u, variants=variants, one_cov=True, thin=True)
    th = sort_unitigs(th)
    ov = np.zeros((len(th.overlaps), 6), np.uint32)
    ov[:, :4], ov[:, 5] = th.overlaps[:, :4], th.overlaps[:, 4]
    gfa1, gaf1 = reference_backend(1, fa, tmp_path / "s1.gfa", tmp_path / "s1.gaf", mode, k, w, a, u, r, big_r, True, variants=variants,
                                   node_hash=th.node_hash, node_cov=th.node_cov, node_tip=th.node_tip, edges=th.edges, new_overlaps=ov,
                                   unitig_begin=th.unitig_begin, unitig_nodes=th.unitig_nodes, unitig_edges=th.unitig_edges)
    assert hashlib.md5(gfa1).hexdigest() == hashlib.md5(want_gfa).hexdigest() and hashlib.md5(gaf1).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "stage 1"


def test_dropin_executable_argv_and_failure_modes(reference, tmp_path):
    """oracle/_ref/MBG_b200 (integration build: MBG's main.cpp + back end around libmbgfe, oracle/mbg_b200.cpp) without a GPU: options
    outside the B200 front end run the reference's own runMBG (renamed, not removed) and give MBG's files; the B200 path itself fails
    loudly instead of falling back; MBG's own argument validation is untouched (main.cpp:158-198)."""
    import subprocess
    import torch
    from oracle_lib import REF_MBG, write_fasta
    exe = os.path.join(os.path.dirname(REF_MBG), "MBG_b200")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/MBG_b200 not built (make -C oracle ref dropin)")
    bases, offsets = workloads.as_pair(workloads.coverage_reads(genome_len=40000, n_reads=60, len_min=6000, len_max=10000, error_ppm=2000))
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    args = ["-i", str(fa), "-k", "301", "-w", "30", "-a", "1", "-u", "2", "-t", "1", "--error-masking=collapse-msat", "-r", "1500", "-R", "400", "--do-unsafe-guesswork-resolutions", "--keep-gaps"]
    subprocess.run([REF_MBG] + args + ["-o", str(tmp_path / "a.gfa"), "--output-sequence-paths", str(tmp_path / "a.gaf")], check=True, stderr=subprocess.DEVNULL)
    r = subprocess.run([exe] + args + ["-o", str(tmp_path / "b.gfa"), "--output-sequence-paths", str(tmp_path / "b.gaf")], stderr=subprocess.PIPE)
    assert r.returncode == 0 and b"reference path" in r.stderr
    assert (tmp_path / "a.gfa").read_bytes() == (tmp_path / "b.gfa").read_bytes() and (tmp_path / "a.gaf").read_bytes() == (tmp_path / "b.gaf").read_bytes()
    r = subprocess.run([exe, "-i", str(fa), "-o", str(tmp_path / "c.gfa"), "-k", "300", "-w", "30"], stderr=subprocess.PIPE)      # even k: MBG's own check
    assert r.returncode != 0
    if not torch.cuda.is_available():
        r = subprocess.run([exe] + args[:-1] + ["-o", str(tmp_path / "d.gfa")], stderr=subprocess.PIPE)
        assert r.returncode != 0 and b"no CPU path" in r.stderr
