"""Kernel logic checked WITHOUT a GPU: the CUDA sources compiled against tests/emu/cuda_emu.h (an execution-model
emulator used for debugging; see that header) and driven through the same C ABI, compared with the oracle.
These are not the parity tests proper -- those are tests/test_gpu_parity.py on a real B200."""
import numpy as np
import pytest

import workloads
from fe_helpers import run_frontend
from oracle_lib import compare_runs


@pytest.mark.parametrize("mode,k,w", [("collapse-msat", 1001, 100), ("hpc", 501, 50), ("no", 63, 33), ("msat", 255, 7), ("collapse-dinuc", 301, 30)])
def test_emulated_kernels_match_oracle(oracle, emu_library, mode, k, w):
    reads = [workloads.read_a()[:9000]] + workloads.edge_case_reads()
    compare_runs(run_frontend(mode, k, w, reads, library=emu_library), oracle.run(mode, k, w, reads), what=f"{mode} k={k} w={w}")


# (k, w): every residue of s = k - w + 1 mod 8 (where the incoming characters of K3a's rolled hashes sit inside their 16-code window,
# odd residues through the 16-bit funnel), windows inside one block of 8 (w <= 8), spanning 2, 3, .. blocks, every r = (w - 1) mod 8
SELECT_SHAPES = [(31, 1), (33, 2), (35, 3), (37, 4), (39, 5), (41, 6), (43, 7), (45, 8), (63, 9), (65, 10), (67, 12), (99, 16), (101, 17), (127, 24), (201, 31),
                 (255, 40), (301, 57), (401, 64), (501, 95), (1001, 100), (1001, 128), (1001, 256)]


@pytest.mark.parametrize("k,w", SELECT_SHAPES)
def test_emulated_select_shapes(oracle, emu_library, k, w):
    """K3a in rolling form over the shapes its index arithmetic distinguishes (s mod 8, a = (w-1) div 8 incl. 0, r = (w-1) mod 8),
    on reads that end 0..9 codes past a block, with tandem repeats (equal hashes inside a window: ties are kept)"""
    a = workloads.read_a()
    reads = [a[:3000 + 7 * i] for i in range(10)] + [workloads.repeat_rich(k + w, 4000), workloads.long_tandem_reads()[0][:6000], a[:k + 40], a[:k], a[:k - 1]]
    for mode in ("hpc", "collapse-msat"):
        got = run_frontend(mode, k, w, reads, library=emu_library)
        compare_runs(got, oracle.run(mode, k, w, reads), what=f"select shape {mode} k={k} w={w} s%8={(k - w + 1) % 8}")


@pytest.mark.parametrize("seed", range(24))
def test_emulated_random_inputs(oracle, emu_library, seed):
    """seeded random configurations against the oracle (workloads.random_case): any mode, odd k from 31, any legal w, reads of
    repeats and plain sequence, lengths around k and around the tile boundaries, stray characters, lower case, duplicates"""
    mode, k, w, reads = workloads.random_case(seed)
    compare_runs(run_frontend(mode, k, w, reads, library=emu_library), oracle.run(mode, k, w, reads), what=f"random seed={seed} {mode} k={k} w={w}")


def test_emulated_kernels_table_with_coverage(oracle, emu_library):
    reads = workloads.coverage_reads(n_reads=6)
    compare_runs(run_frontend("collapse-msat", 501, 50, reads, library=emu_library), oracle.run("collapse-msat", 501, 50, reads), what="coverage")


def test_emulated_tables_sized_from_a_sketch(oracle, emu_library, monkeypatch):
    """the exact tail sizes the node and edge tables from a distinct-count sketch of the hashes (two slots per expected distinct
    key instead of two per instance); an estimate that comes out far too low is mended by doubling the overflowed tables"""
    reads = workloads.coverage_reads(n_reads=30, genome_len=20000)          # ~15x coverage: far fewer distinct k-mers than instances
    want = oracle.run("collapse-msat", 301, 30, reads)
    from verkko_b200.frontend import MbgFrontEnd

    def run():      # upload + process: the exact tail (mbgfe_run sizes its tables from the first piece, before it can know)
        fe = MbgFrontEnd(301, 30, "collapse-msat", library=emu_library)
        try:
            fe.add_reads(*workloads.as_pair(reads))
            fe.upload()
            fe.process()
            return dict(node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges(), stats=fe.stats())
        finally:
            fe.close()

    def same(got):
        return all(np.array_equal(got[f], want[f]) for f in ("node_hash", "node_cov", "node_tip", "edges"))

    monkeypatch.setenv("MBGFE_NO_SKETCH", "1")
    plain = run()
    monkeypatch.delenv("MBGFE_NO_SKETCH")
    monkeypatch.setenv("MBGFE_SKETCH_MIN", "0")
    got = run()
    assert same(plain) and same(got)
    assert got["stats"]["table_slots"] < plain["stats"]["table_slots"], (got["stats"]["table_slots"], plain["stats"]["table_slots"])
    assert got["stats"]["table_slots"] >= 2 * len(want.node_hash)
    monkeypatch.setenv("MBGFE_SKETCH_DIV", "16")
    low = run()
    assert same(low)
    assert low["stats"]["table_grows"] >= 1 and low["stats"]["table_slots"] >= len(want.node_hash)


def test_emulated_palindrome_rule(oracle, emu_library):
    """droppable palindromes vanish like in the reference; a non-droppable one is an error in both"""
    from verkko_b200 import MbgfeError
    reads = [workloads.repeat_rich(9, 12000), workloads.palindrome_read(1000)]
    compare_runs(run_frontend("dinuc", 63, 33, reads, library=emu_library), oracle.run("dinuc", 63, 33, reads), what="droppable")
    with pytest.raises(RuntimeError, match="palindromic"):
        oracle.run("dinuc", 63, 33, [workloads.palindrome_read(1021)])
    with pytest.raises(MbgfeError) as e:
        run_frontend("dinuc", 63, 33, [workloads.palindrome_read(1021)], library=emu_library)
    assert e.value.code == -4


def test_emulated_pipelined_run_in_small_pieces(oracle, emu_library):
    """mbgfe_run with tiny H2D pieces: many pieces, per-piece tail launches, sizes extrapolated from the first piece"""
    reads = workloads.coverage_reads(n_reads=6)
    got = run_frontend("collapse-msat", 501, 50, reads, library=emu_library, piece_bytes=9000)
    compare_runs(got, oracle.run("collapse-msat", 501, 50, reads), what="small pieces")
    assert got["stats"]["table_grows"] == 0


def test_emulated_pipelined_run_estimate_overflow(oracle, emu_library):
    """first piece = a read whose collapse leaves almost nothing (so the extrapolated sizes are far too small): the
    kernels refuse to write past them and the exact tail reruns"""
    poor = (b"A" * 320 + b"C" * 320) * 5          # exactly one piece
    reads = [poor] + [workloads.read_a()[:2500] for _ in range(3)]
    got = run_frontend("collapse", 31, 1, reads, library=emu_library, piece_bytes=3200)
    compare_runs(got, oracle.run("collapse", 31, 1, reads), what="overflow fallback")
    assert got["stats"]["table_grows"] >= 1


def test_emulated_tandem_repeats_longer_than_a_tile(oracle, emu_library):
    reads = workloads.long_tandem_reads()[:3] + [workloads.repeat_rich(22, 5000)]
    compare_runs(run_frontend("msat", 63, 33, reads, library=emu_library), oracle.run("msat", 63, 33, reads), what="long tandem")


@pytest.mark.parametrize("mode,k,w,flags", [("collapse-msat", 301, 30, None), ("hpc", 63, 33, 0)])
def test_emulated_hpc_variant_rehash(oracle, emu_library, mode, k, w, flags):
    """row a11 (MBG/ReadHelper.h:680-737); flags=0: the library keeps poses on its own when variants are registered"""
    from oracle_lib import make_variants
    reads = workloads.coverage_reads(n_reads=10)
    variants = make_variants(oracle.run(mode, k, w, reads), k, 150)
    kw = {} if flags is None else {"flags": flags}
    got = run_frontend(mode, k, w, reads, library=emu_library, variants=variants, **kw)
    want = oracle.run(mode, k, w, reads, variants=variants)
    if flags is None:
        compare_runs(got, want, what="emu variants")
    else:
        assert np.array_equal(got.hashes, want.hashes) and np.array_equal(got.node_hash, want.node_hash) and np.array_equal(got.node_cov, want.node_cov)


def test_emulated_noalloc_kmer_hash(oracle, emu_library):
    """mbgfe_kat_kmer_hash_noalloc on the emulator build: host logic (limit, clearing of the shared scratch between calls);
    the GPU version is tests/test_zz_no_alloc.py"""
    from verkko_b200.frontend import MbgFrontEnd, MbgfeError
    fe = MbgFrontEnd(1001, 100, library=emu_library)
    try:
        rng = np.random.default_rng(2)
        for k in (2001, 31, 16384, 33):
            codes = rng.integers(0, 30948, k).astype(np.uint16)
            assert fe.kat_kmer_hash_noalloc(codes) == fe.kat_kmer_hash(codes) == oracle.kmer_hash(codes), k
        with pytest.raises(MbgfeError, match="MBGFE_NOALLOC_MAX_K"):
            fe.kat_kmer_hash_noalloc(np.zeros(16385, dtype=np.uint16))
    finally:
        fe.close()


def test_emulated_hpc_variants_keep_the_pipelined_run(oracle, emu_library, tmp_path):
    """pass 3 of runMBG (reads + registered hpc variants) goes through mbgfe_run's pipelined upload, from memory and from a
    FASTA file: the re-hash runs per piece between K3b and the table inserts"""
    from oracle_lib import make_variants, write_fasta
    from verkko_b200.frontend import MbgFrontEnd
    mode, k, w = "collapse-msat", 301, 30
    reads = workloads.coverage_reads(n_reads=12)
    variants = make_variants(oracle.run(mode, k, w, reads), k, 150)
    want = oracle.run(mode, k, w, reads, variants=variants)
    got = run_frontend(mode, k, w, reads, library=emu_library, variants=variants)
    compare_runs(got, want, what="emu variants, pipelined from memory")
    fa = str(tmp_path / "reads.fa")
    bases, offsets = workloads.as_pair(reads)
    write_fasta(fa, bases, offsets)
    fe = MbgFrontEnd(k, w, mode, library=emu_library)
    try:
        fe.add_file(fa)
        fe.set_variants(variants.hashes, variants.offsets, variants.len_begin, variants.lengths)
        fe.run()
        assert np.array_equal(fe.node_hashes(), want.node_hash) and np.array_equal(fe.node_coverage(), want.node_cov)
        assert np.array_equal(fe.edges(), want.edges)
        assert fe.stats()["h2d_bytes"] < len(bases) // 3, "the file route ships 2-bit reads"
    finally:
        fe.close()


def check_unitigs(got, want, what):
    from oracle_lib import compare_unitigs
    assert np.array_equal(got.kept, want.kept), f"{what}: kept flags differ"
    compare_unitigs(got, want, what)
    # internal consistency of the unitig arrays: a partition of the nodes with coverage >= -a
    b = got.unitig_begin.astype(np.int64)
    ids = got.unitig_nodes >> 1
    assert len(np.unique(ids)) == len(ids)
    for u in range(len(b) - 1):
        assert np.all(got.node_unitig[ids[b[u]:b[u + 1]]] == u)


@pytest.mark.parametrize("mode,k,w,a,u,which", [("hpc", 63, 33, 2, 2, "repeats"), ("no", 63, 5, 1, 2, "circular")])      # "coverage": the onecov / threading tests below
def test_emulated_unitig_filter(oracle, emu_library, mode, k, w, a, u, which):
    """rows a14 / f1 on the emulator build (the GPU versions are tests/test_gpu_parity.py::test_unitig_filter*)"""
    from fe_helpers import run_frontend_unitigs
    reads = {"coverage": lambda: workloads.coverage_reads(n_reads=16, error_ppm=4000),
             "repeats": lambda: [workloads.repeat_rich(3, 4000), workloads.repeat_rich(3, 4000), workloads.repeat_rich(4, 2000)],
             "circular": lambda: workloads.circular_reads(genome_len=1500, read_len=1200, step=300) + [workloads.lcg_bases(77, 800)]}[which]()
    _, got = run_frontend_unitigs(mode, k, w, reads, a, u, library=emu_library)
    _, want = oracle.run_unitigs(mode, k, w, reads, a, u)
    check_unitigs(got, want, f"emu unitigs {mode} k={k} a={a} u={u}")
    if which == "circular":
        assert len(want.unitig_begin) == 2 and want.unitig_begin[1] > 100      # one circular unitig; the lone read is filtered out


@pytest.mark.parametrize("mode,k,w,a,u,which", [("collapse-msat", 301, 30, 1, 2, "coverage"), ("hpc", 63, 33, 1, 3, "noisy")])
def test_emulated_onecov_heuristic(oracle, emu_library, mode, k, w, a, u, which):
    """row a14: removeOnecovNodes (MBG/MBG.cpp:527-627) on the emulator build, against the oracle (itself pinned on the
    reference's getUnitigGraph(…, oneCovHeuristic = true) in tests/test_oracle.py)"""
    from fe_helpers import run_frontend_unitigs
    reads = {"coverage": lambda: workloads.coverage_reads(n_reads=16, error_ppm=4000),
             "noisy": lambda: workloads.coverage_reads(seed=3, genome_len=8000, n_reads=10, error_ppm=10000),
             "circular": lambda: workloads.circular_reads(genome_len=1500, read_len=1200, step=300) + [workloads.lcg_bases(77, 800)]}[which]()
    _, got = run_frontend_unitigs(mode, k, w, reads, a, u, library=emu_library, one_cov=True)
    _, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=True)
    check_unitigs(got, want, f"emu onecov unitigs {mode} k={k} a={a} u={u}")
    if which == "coverage":
        _, plain = oracle.run_unitigs(mode, k, w, reads, a, u)
        assert not np.array_equal(plain.kept, want.kept), "the heuristic did not change anything on this workload"


@pytest.mark.parametrize("mode,k,w,a,u,one_cov,which", [("collapse-msat", 301, 30, 1, 2, True, "coverage"), ("no", 63, 5, 1, 1, False, "circular")])
def test_emulated_read_threading(oracle, emu_library, mode, k, w, a, u, one_cov, which):
    """row f2: mbgfe_thread_reads (getKmerLocator + iterateReadPaths as a batched lookup) on the emulator build, against
    the oracle's sequential restatement (pinned on the reference's own functions in tests/test_oracle.py)"""
    from fe_helpers import run_frontend_unitigs
    from oracle_lib import compare_paths
    reads = {"coverage": lambda: workloads.coverage_reads(n_reads=16, error_ppm=4000),
             "noisy": lambda: workloads.coverage_reads(seed=3, genome_len=8000, n_reads=10, error_ppm=10000),
             "circular": lambda: workloads.circular_reads(genome_len=1500, read_len=1200, step=300) + [workloads.lcg_bases(77, 800)]}[which]()
    _, got = run_frontend_unitigs(mode, k, w, reads, a, u, library=emu_library, one_cov=one_cov, paths=True)
    _, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
    compare_paths(got.paths, got, want.paths, want, f"emu read paths {mode} k={k}")
    assert len(want.paths.heads) > 0


@pytest.mark.parametrize("seed", range(12))
def test_emulated_graph_stage_random(oracle, emu_library, seed):
    """rows a14 / f1 / f2 on seeded random read sets against the oracle: a genome with repeats read at 6 .. 14x with 0.2 .. 1.5 %
    errors (tips, bubbles and coverage-1 k-mers for the filters), random mode / k / w / -a / -u, with and without the one-coverage
    heuristic: kept flags, unitigs, unitig edges and the read paths"""
    from fe_helpers import run_frontend_unitigs
    from oracle_lib import compare_paths
    mode, k, w, reads, a, u, one_cov = workloads.random_graph_case(seed)
    _, got = run_frontend_unitigs(mode, k, w, reads, a, u, library=emu_library, one_cov=one_cov, paths=True)
    _, want = oracle.run_unitigs(mode, k, w, reads, a, u, one_cov=one_cov, paths=True)
    check_unitigs(got, want, f"random graph stage seed={seed} {mode} k={k} w={w} a={a} u={u}")
    compare_paths(got.paths, got, want.paths, want, f"random read paths seed={seed}")


def test_emulated_graph_stage_on_empty_and_tiny_tables(oracle, emu_library):
    """edge cases of rows a14 / f1 / f2: no k-mer at all, a single read whose k-mers all have coverage 1 (everything is
    filtered), reads shorter than 32 bases"""
    from fe_helpers import run_frontend_unitigs
    for reads, flags in (([workloads.lcg_bases(5, 200)], (True,)), ([workloads.lcg_bases(5, 2000)], (False, True))):
        for one_cov in flags:
            table, got = run_frontend_unitigs("collapse-msat", 301, 30, reads, 1, 2, library=emu_library, one_cov=one_cov, paths=True)
            otable, want = oracle.run_unitigs("collapse-msat", 301, 30, reads, 1, 2, one_cov=one_cov, paths=True)
            assert len(table.node_cov) == len(otable.node_cov) and np.array_equal(got.kept, want.kept)
            assert len(got.node_cov) == len(want.node_cov) == 0 and len(got.unitig_begin) == 1 and len(got.paths.heads) == len(want.paths.heads) == 0


def test_emulated_unitig_numbering_is_unitiggraph_sort(reference, emu_library):
    """the numbering and orientation mbgfe_build_unitigs returns is what UnitigGraph::sort (MBG/UnitigGraph.cpp:278-340)
    leaves after sortByHash: node ids in hash order, every unitig read from its end with the smaller node id (a single
    k-mer forwards), unitigs ordered by that first node -- checked against the reference's own unitigs with that rule
    applied (linear unitigs; circular ones are cut elsewhere)"""
    from fe_helpers import run_frontend_unitigs
    reads = workloads.coverage_reads(seed=3, genome_len=8000, n_reads=10, error_ppm=10000)
    _, got = run_frontend_unitigs("hpc", 63, 33, reads, 1, 2, library=emu_library, one_cov=True)
    want = reference.run_unitigs("hpc", 63, 33, reads, 1, 2, one_cov=True)
    order = np.lexsort((want.node_hash[:, 0], want.node_hash[:, 1]))
    new_id = np.empty(len(order), np.int64)
    new_id[order] = np.arange(len(order))
    b = want.unitig_begin.astype(np.int64)
    seqs = []
    for i in range(len(b) - 1):
        x = want.unitig_nodes[b[i]:b[i + 1]].astype(np.int64)
        x = 2 * new_id[x >> 1] + (x & 1)
        if len(x) == 1:
            x = x | 1
        elif (x[-1] >> 1) < (x[0] >> 1):
            x = x[::-1] ^ 1
        seqs.append(x)
    seqs.sort(key=lambda x: min(x[0] >> 1, x[-1] >> 1))
    gb = got.unitig_begin.astype(np.int64)
    assert len(gb) - 1 == len(seqs) > 5 and any(len(x) == 1 for x in seqs)
    for i, x in enumerate(seqs):
        assert np.array_equal(got.unitig_nodes[gb[i]:gb[i + 1]].astype(np.int64), x), f"unitig {i}"


def test_emulated_unitig_kmer_thinning(oracle, emu_library):
    """filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185, MBG's default thinning) on the emulator build: next-pointers +
    orbit marking by pointer doubling against the oracle's sequential walk (pinned on the reference in tests/test_oracle.py)"""
    from fe_helpers import check_thinning
    from verkko_b200.frontend import MbgFrontEnd
    reads = workloads.coverage_reads(n_reads=16, error_ppm=4000)
    fe = MbgFrontEnd(301, 30, "collapse-msat", library=emu_library)
    try:
        fe.add_reads(*reads) if isinstance(reads, tuple) else fe.add_read_list(reads)
        fe.run()
        ug = fe.build_unitigs(1, 2, one_cov_heuristic=True)
        thin, off = fe.thin_unitigs()
        thin, off = thin.copy(), off.copy()
    finally:
        fe.close()
    _, want = oracle.run_unitigs("collapse-msat", 301, 30, reads, 1, 2, one_cov=True, thin=True)
    assert check_thinning(ug, thin, off, want, 301, "emu thinning") > 10 and thin.sum() < len(thin)


def test_emulated_thinning_applied_and_read_paths_on_the_thinned_list(oracle, emu_library):
    """row f2 as runMBG takes it by default: read threading AFTER filterKmersToUnitigKmers, on the emulator build"""
    from fe_helpers import run_thinned_pass
    reads = workloads.coverage_reads(n_reads=16, error_ppm=4000)
    n_paths, n_new = run_thinned_pass("collapse-msat", 301, 30, reads, 1, 2, oracle, library=emu_library, what="emu thinned pass")
    assert n_paths > 5 and n_new > 5


@pytest.mark.parametrize("seed", range(8))
def test_emulated_thinned_pass_random(oracle, emu_library, seed):
    """MBG's default order on seeded random read sets: unitigs -> filterKmersToUnitigKmers applied -> read threading on the thinned
    list (fe_helpers.run_thinned_pass checks the thinned tables, the new overlaps and the read paths against the oracle)"""
    from fe_helpers import run_thinned_pass
    from verkko_b200 import synth
    rng = np.random.default_rng(900 + seed)
    mode = ["collapse-msat", "hpc", "no", "collapse-dinuc"][seed % 4]
    k = 31 + 2 * int(rng.integers(10, 140))
    w = int(rng.integers(3, min(40, k - 30) + 1))
    g = synth.genome(int(rng.integers(1, 1 << 20)), int(rng.integers(8000, 20000)))
    lmin = max(2 * k + 100, 1500)
    cov = int(rng.integers(8, 15))
    reads = synth.reads(int(rng.integers(1, 1 << 20)), g, cov * len(g) // (lmin + 1000), lmin, lmin + 2000, int(rng.integers(1000, 8000)))
    run_thinned_pass(mode, k, w, reads, 1, int(rng.integers(1, 3)), oracle, library=emu_library, what=f"random thinned pass seed={seed} {mode} k={k} w={w}")


@pytest.mark.parametrize("pct,mode,k,w", [("100", "hpc", 63, 33), ("50", "collapse-msat", 301, 30)])
def test_emulated_host_packed_pieces(oracle, emu_library, monkeypatch, pct, mode, k, w):
    """the two upload routes of mbgfe_run (ASCII + K0 on the device, 2-bit packing on host threads) mixed piece by piece,
    with lower case and characters outside ACGT in host-packed pieces"""
    monkeypatch.setenv("MBGFE_HOST_PACK", pct)
    reads = [workloads.read_a()] + workloads.edge_case_reads()
    reads += [workloads.lcg_bases(5, 9000).replace(b"ACGTA", b"ACNNA", 3), workloads.lcg_bases(6, 7000).lower(), workloads.lcg_bases(8, 4097)]
    got = run_frontend(mode, k, w, reads, library=emu_library, piece_bytes=3200)      # (0 %: every other emulator test)
    compare_runs(got, oracle.run(mode, k, w, reads), what=f"host pack {pct}% {mode}")


@pytest.mark.parametrize("mode,k,w", [("hpc", 63, 33), ("no", 101, 20), ("dinuc", 63, 33), ("collapse", 31, 1), ("msat", 63, 33), ("collapse-msat", 101, 20)])
def test_emulated_collapse_by_table_without_positions(oracle, emu_library, mode, k, w):
    """k_collapse_lut (the K1 that runs when per-base positions are not wanted: flags = 0) in every run-length mode, ragged lengths
    around the 4 096-base tile and the 128-base lane share; in the msat modes also K2's store-per-symbol emission of single codes
    (the path without positions), with tandem repeats longer than a tile"""
    a = workloads.read_a()
    reads = [a, a[:4096], a[:4097], a[:4095 + 128], a[:129], a[:127], a[:33], workloads.read_with_other_characters(), workloads.repeat_rich(7, 9000), b"A" * 5000 + a[:300]]
    if "msat" in mode:
        reads += workloads.long_tandem_reads()[:2] + [workloads.repeat_rich(22, 5000)]
    got = run_frontend(mode, k, w, reads, library=emu_library, flags=0)
    want = oracle.run(mode, k, w, reads)
    assert np.array_equal(got.positions, want.positions) and np.array_equal(got.hashes, want.hashes)
    assert np.array_equal(got.node_hash, want.node_hash) and np.array_equal(got.node_cov, want.node_cov)
    for f in ("read_idx", "part_offset", "raw_len", "n_codes", "n_sel"):
        assert np.array_equal(got.parts[f], want.parts[f]), f
