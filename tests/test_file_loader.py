"""Row a1 (MBG/fastqloader.h:38-147): the streaming FASTA loader of mbgfe_run -- pread + scan + 2-bit packing in one pass on
worker threads, record structure stitched from the scan events, the device image in file coordinates -- against the oracle fed
the reads the reference's parser would hand over.  The CPU half runs on the GPU-less emulator build (host logic is the same
code); the GPU half runs the real library."""
import os

import numpy as np
import pytest

import workloads
from oracle_lib import compare_runs, lcg_bases
from verkko_b200.frontend import KEEP_CODES, KEEP_POSES, MbgFrontEnd


def reference_parse(data):
    """streamFastqFastaFromStream (MBG/fastqloader.h:38-80) restated: names and upper-cased sequences"""
    names, seqs = [], []
    lines = data.split(b"\n")
    i, n = 0, len(lines)
    if data.endswith(b"\n"):
        n -= 1                       # getline on the trailing newline yields no further line
    cur = None
    while i < n:
        line = lines[i]
        i += 1
        if len(line) == 0:
            continue
        if line[:1] == b">":
            if line.endswith(b"\r"):
                line = line[:-1]
            names.append(line[1:].decode())
            seqs.append(b"")
            cur = len(seqs) - 1
            continue
        if cur is None:
            continue
        if line.endswith(b"\r"):
            line = line[:-1]
        seqs[cur] += line.upper()
    return names, seqs


def run_file(paths, mode, k, w, library=None, **kw):
    from oracle_lib import RunResult
    fe = MbgFrontEnd(k, w, mode, flags=KEEP_CODES | KEEP_POSES, library=library, **kw)
    try:
        for p in paths:
            fe.add_file(str(p))
        fe.run()
        res = RunResult(parts=fe.parts(), codes=fe.codes(), poses=fe.poses(), positions=fe.positions(), hashes=fe.hashes(),
                        node_hash=fe.node_hashes(), node_cov=fe.node_coverage(), node_tip=fe.node_tip(), edges=fe.edges())
        res["stats"] = fe.stats()
        res["names"] = [fe.read_name(i) for i in range(res["stats"]["reads"])]
        return res
    finally:
        fe.close()


def awkward_fasta():
    """single-line records with everything the parser has rules for"""
    a = workloads.read_a()
    recs = [
        (b"plain read", a[:9000]),
        (b"crlf line\r", a[1000:7000] + b"\r"),
        (b"lower case", a[2000:6000].lower()),
        (b"with N", workloads.read_with_other_characters()),
        (b"short", a[:31]),
        (b"empty", None),
        (b"thirty-two", a[:32]),
        (b"cr inside", a[:3000] + b"\r" + a[3000:6000]),
        (b"gt inside", a[:2500] + b">" + a[2500:5000]),
        (b"last one without newline", a[5000:12000]),
    ]
    out = b"text before the first header\nACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT\n\n"
    for i, (name, seq) in enumerate(recs):
        out += b">" + name + b"\n"
        if seq is not None:
            out += seq
            if i + 1 < len(recs):
                out += b"\n"
        if i % 3 == 0:
            out += b"\n"            # empty lines between records
    return out


def check_against_oracle(oracle, got, datas, mode, k, w, what):
    names, seqs = [], []
    for d in datas:
        n, s = reference_parse(d)
        names += n
        seqs += s
    want = oracle.run(mode, k, w, seqs)
    compare_runs(got, want, what=what)
    assert got["names"] == names, what
    assert got["stats"]["reads"] == len(seqs) and got["stats"]["raw_bases"] == sum(len(s) for s in seqs)


@pytest.mark.parametrize("piece", [0, 1024, 96])
def test_emulated_streaming_fasta(oracle, emu_library, tmp_path, piece):
    data = awkward_fasta()
    fa = tmp_path / "awkward.fa"
    fa.write_bytes(data)
    second = b">r0\n" + lcg_bases(3, 5000) + b"\n>r1 x\n" + lcg_bases(4, 4000).lower() + b"\n"
    fb = tmp_path / "second.fasta"
    fb.write_bytes(second)
    for mode, k, w in (("collapse-msat", 301, 30), ("hpc", 63, 33)):
        got = run_file([fa, fb], mode, k, w, library=emu_library, piece_bytes=piece)
        check_against_oracle(oracle, got, [data, second], mode, k, w, f"streaming loader piece={piece} {mode}")
        assert got["stats"]["h2d_bytes"] < (len(data) + len(second)) // 3, "the bus must carry packed bases, not the file"


def random_fasta(seed):
    """single-line records with runs of characters outside ACGT of every kind the collapse rule distinguishes: the same character
    repeated, different ones in a row, lower and upper case mixed inside a run, at the start and at the end of a read, next to a
    base, a lone \\r in the middle, \\r\\n line ends, empty records and empty lines"""
    rng = np.random.default_rng(seed)
    out = b""
    for r in range(int(rng.integers(3, 9))):
        n = int(rng.integers(20, 5000))
        seq = bytearray(workloads.repeat_rich(int(rng.integers(1, 1 << 30)), n) if rng.random() < 0.5 else lcg_bases(int(rng.integers(1, 1 << 20)), n))
        for _ in range(int(rng.integers(0, 6))):
            at = int(rng.integers(0, n))
            run = bytes(rng.choice(list(b"NNNRYnrK\r-*"), size=int(rng.integers(1, 7))).astype(np.uint8))
            seq[at:at + len(run)] = run
        seq = bytes(seq[:n])
        if rng.random() < 0.2:
            seq = b"N" * int(rng.integers(1, 4)) + seq
        if rng.random() < 0.2:
            seq = seq + b"n" * int(rng.integers(1, 4))
        if rng.random() < 0.15:
            seq = seq.lower()
        end = b"\r\n" if rng.random() < 0.3 else b"\n"
        out += b">read %d of seed %d" % (r, seed) + end
        if rng.random() < 0.9:
            out += seq + end
        if rng.random() < 0.3:
            out += b"\n"
    return out


@pytest.mark.parametrize("seed", range(16))
def test_emulated_streaming_fasta_random(oracle, emu_library, tmp_path, seed):
    """the stitcher cuts reads at characters outside ACGT itself (no second pass over the image): random files against the oracle fed
    what the reference's parser would hand over, in the modes where the collapse step looks at those characters"""
    data = random_fasta(seed)
    fa = tmp_path / "random.fa"
    fa.write_bytes(data)
    mode, k, w = [("collapse-msat", 101, 20), ("collapse", 31, 1), ("hpc", 63, 33), ("collapse-dinuc", 63, 9)][seed % 4]
    got = run_file([fa], mode, k, w, library=emu_library, piece_bytes=[0, 256, 4096, 96][(seed // 4) % 4])
    check_against_oracle(oracle, got, [data], mode, k, w, f"random fasta seed={seed} {mode}")
    assert got["stats"]["h2d_bytes"] < len(data) // 3 + 4096, "the streaming loader ran (packed bases on the bus), not the plain parser"


def test_emulated_streaming_fasta_falls_back_on_multiline(oracle, emu_library, tmp_path):
    a = workloads.read_a()
    data = b">one\n" + a[:5000] + b"\n>two\n" + b"\n".join(a[5000 + i:5000 + i + 70] for i in range(0, 4000, 70)) + b"\n>three\n" + a[9000:15000] + b"\n"
    fa = tmp_path / "multi.fa"
    fa.write_bytes(data)
    got = run_file([fa], "collapse-msat", 301, 30, library=emu_library, piece_bytes=2048)
    check_against_oracle(oracle, got, [data], "collapse-msat", 301, 30, "multi-line fallback")


@pytest.mark.gpu
@pytest.mark.parametrize("piece", [0, 1 << 16, 4096])
def test_streaming_fasta_on_gpu(oracle, tmp_path, piece):
    from verkko_b200 import synth
    g = synth.genome(7, 300000)
    bases, offsets = synth.reads(11, g, 400, 11000, 19000, 1000)
    from oracle_lib import write_fasta
    fa = tmp_path / "reads.fa"
    write_fasta(fa, bases, offsets)
    data = awkward_fasta()
    fb = tmp_path / "awkward.fasta"
    fb.write_bytes(data)
    got = run_file([fa, fb], "collapse-msat", 1001, 100, piece_bytes=piece)
    check_against_oracle(oracle, got, [fa.read_bytes(), data], "collapse-msat", 1001, 100, f"gpu streaming piece={piece}")
    # characters outside ACGT are cut by the stitcher as the pieces come in: the run never goes back over the image (a second,
    # staged pass would report per-kernel times)
    assert got["stats"]["nonacgt"] > 0 and got["stats"]["ms_collapse"] == 0 and got["stats"]["ms_select"] == 0
    # the same through the plain parser (MBGFE_FILE_LEGACY): identical arrays
    os.environ["MBGFE_FILE_LEGACY"] = "1"
    try:
        plain = run_file([fa, fb], "collapse-msat", 1001, 100)
    finally:
        del os.environ["MBGFE_FILE_LEGACY"]
    compare_runs(got, plain, what="streaming vs plain parser")
