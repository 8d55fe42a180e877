"""Seeded inputs shared by the CPU and GPU tests."""
import numpy as np

from oracle_lib import lcg_bases, pack_reads


def read_a():
    """SURVEY.md 8(c) KAT read: LCG(1), 20 000 bases."""
    return lcg_bases(1, 20000)


def read_with_other_characters():
    rb = bytearray(read_a())
    rb[5000:5002] = b"NN"
    rb[9000:9003] = b"NRN"
    rb[0:1] = b"N"
    rb[-1:] = b"n"
    rb[12000:12040] = b"N" * 40
    rb[15000:15010] = b"ACGTNACGTN"
    return bytes(rb)


def edge_case_reads():
    a = read_a()
    return [read_with_other_characters(), a[:31], a[:32], b"ACGT" * 10, b"N" * 50, b"", a[100:163], b"acgtacgtac" * 8 + a[:3000].lower()]


def repeat_rich(seed, n):
    """microsatellite / homopolymer rich sequence"""
    rng = np.random.default_rng(seed)
    out = []
    total = 0
    while total < n:
        t = rng.integers(0, 3)
        if t == 0:
            piece = rng.integers(0, 4, rng.integers(1, 40))
        else:
            m = int(rng.integers(1, 8))
            motif = rng.integers(0, 4, m)
            c = int(rng.integers(2, 30))
            piece = np.tile(motif, c)[: m * c - int(rng.integers(0, m))]
        out.append(piece)
        total += len(piece)
    return bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[np.concatenate(out)[:n]])


def coverage_reads(seed=11, genome_len=30000, n_reads=12, len_min=8000, len_max=12000, error_ppm=1000):
    from verkko_b200 import synth
    g = synth.genome(7, genome_len)
    return synth.reads(seed, g, n_reads, len_min, len_max, error_ppm)


def as_pair(reads):
    return reads if isinstance(reads, tuple) else pack_reads(reads)


def palindrome_read(seed):
    """X + (AT)^6 + revcomp(X): in dinuc mode the AT run is one self-complementary code, so the k-mer centred on it is
    a reverse-complement palindrome.  seed 1021 (k=63, w=33): the palindrome is selected with neighbours >= k apart,
    where the reference aborts (MBG/ReadHelper.h:775-777); seed 1000: dropped silently / not selected."""
    comp = {65: 84, 67: 71, 71: 67, 84: 65}
    x = lcg_bases(seed, 400)
    return x + b"AT" * 6 + bytes(comp[c] for c in reversed(x))


def long_tandem_reads(seed=5):
    """tandem repeats much longer than a K2 tile (1024 symbols) at assorted alignments, between random flanks: one code
    spans several tiles, its slot leaves with an earlier tile than the one that computes it"""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)

    def rnd(n):
        return bytes(acgt[rng.integers(0, 4, n)])
    reads = []
    for motif, copies, lead in ((b"AC", 900, 700), (b"ACG", 800, 1000), (b"ACGT", 300, 1019), (b"ACGTC", 500, 33), (b"ACGTCA", 400, 2040), (b"AC", 2000, 5)):
        reads.append(rnd(lead) + motif * copies + rnd(1500))
    return reads


def circular_reads(seed=9, genome_len=3000, read_len=2500, step=300, both_strands=True):
    """error-free reads tiled around a circular random genome: the k-mer graph is one cycle without tips or branches
    (plus, with extra=..., whatever the caller appends), the case getUnitigGraph handles in its last loops"""
    g = lcg_bases(seed, genome_len)
    comp = {65: 84, 67: 71, 71: 67, 84: 65}
    gg = g + g + g
    reads = []
    for i, off in enumerate(range(0, genome_len, step)):
        r = gg[off:off + read_len]
        if both_strands and i % 2:
            r = bytes(comp[c] for c in reversed(r))
        reads.append(r)
    return reads


def diploid_msat_reads(seed=3, genome_len=150000, n_loci=60, coverage_each=15, len_min=6000, len_max=10000, error_ppm=500):
    """reads from two haplotypes that differ in the copy number of inserted microsatellites (motif 2-4, 5-8 copies vs +2): what
    --hpc-variant-onecopy-coverage looks for (run-length clusters with support on both alleles, MBG/ConsensusMaker.cpp:482-537)"""
    from verkko_b200 import synth
    rng = np.random.default_rng(seed)
    base = synth.genome(7, genome_len).tobytes()
    loci = sorted(rng.choice(np.arange(2000, genome_len - 2000), n_loci, replace=False).tolist())
    acgt = b"ACGT"
    h1, h2, last = [], [], 0
    for p in loci:
        m = int(rng.integers(2, 5))
        motif = bytes(acgt[i] for i in rng.permutation(4)[:m])
        c = int(rng.integers(5, 9))
        h1 += [base[last:p], motif * c]
        h2 += [base[last:p], motif * (c + 2)]
        last = p
    h1.append(base[last:]); h2.append(base[last:])
    haps = [np.frombuffer(b"".join(h), dtype=np.uint8).copy() for h in (h1, h2)]
    n = int(genome_len * coverage_each / ((len_min + len_max) / 2))
    parts = [synth.reads(11 + i, haps[i], n, len_min, len_max, error_ppm) for i in range(2)]
    bases = np.concatenate([parts[0][0], parts[1][0]])
    offsets = np.concatenate([parts[0][1], parts[1][1][1:] + parts[0][1][-1]])
    return bases, offsets


def random_case(seed):
    """(mode, k, w, reads) of a seeded random configuration: any mode, odd k from 31, any legal w, reads of repeats and plain
    sequence, lengths around k and around the 32-base / 1 024-symbol tile boundaries, stray characters, lower case, and the
    first two reads once more (coverage 2 on their k-mers, their edges seen twice)"""
    from oracle_lib import MODES
    rng = np.random.default_rng(1000 + seed)
    mode = list(MODES)[int(rng.integers(0, len(MODES)))]
    k = 31 + 2 * int(rng.integers(0, 120))
    w = int(rng.integers(1, k - 30 + 1))
    reads = []
    for _ in range(int(rng.integers(3, 9))):
        n = int(rng.choice([k - 1, k, k + 1, k + 33, 1024, 1025, 2048 + 31, int(rng.integers(40, 6000))]))
        body = repeat_rich(int(rng.integers(1, 1 << 30)), n) if rng.random() < 0.5 else lcg_bases(int(rng.integers(1, 1 << 20)), n)
        body = bytearray(body)
        if rng.random() < 0.3 and n > 10:
            for _ in range(int(rng.integers(1, 4))):
                at = int(rng.integers(0, n - 3))
                body[at:at + int(rng.integers(1, 4))] = b"NRN"[:int(rng.integers(1, 4))]
        if rng.random() < 0.2:
            body = bytearray(bytes(body).lower())
        reads.append(bytes(body[:n]))
    return mode, k, w, reads + reads[:2]


def random_graph_case(seed):
    """(mode, k, w, reads, a, u, one_cov) of a seeded random read set for the graph stage: a genome (every third one with a repeat of
    a few hundred bases) read at 6 .. 14x with 0.2 .. 1.5 % errors -- tips, bubbles and coverage-1 k-mers for the filters"""
    from verkko_b200 import synth
    rng = np.random.default_rng(500 + seed)
    mode = ["collapse-msat", "hpc", "no", "msat"][seed % 4]
    k = 31 + 2 * int(rng.integers(10, 140))
    w = int(rng.integers(3, min(60, k - 30) + 1))
    glen = int(rng.integers(6000, 20000))
    g = synth.genome(int(rng.integers(1, 1 << 20)), glen)
    if seed % 3 == 0:
        g = np.concatenate([g, g[1000:1600], synth.genome(int(rng.integers(1, 1 << 20)), 3000)])
    cov = int(rng.integers(6, 15))
    lmin = max(2 * k + 100, 1500)
    reads = synth.reads(int(rng.integers(1, 1 << 20)), g, cov * len(g) // (lmin + 1000), lmin, lmin + 2000, int(rng.integers(2000, 15000)))
    return mode, k, w, reads, int(rng.integers(1, 3)), int(rng.integers(1, 4)), bool(seed & 1)

