"""The outer boundary of SURVEY.md 8(b): oracle/_ref/MBG_b200 -- MBG's own main.cpp and back end with runMBG replaced by
oracle/mbg_b200.cpp, which runs the front end, the table, getUnitigGraph, the thinning and the read paths through libmbgfe --
called with the argv of Verkko's buildGraph rule (Snakefiles/1-buildGraph.sm:88-100) must write the GFA and GAF `MBG -t 1` writes.
Both executables are integration / test builds under oracle/_ref (compiled from the reference where it lies)."""
import hashlib
import os
import subprocess

import pytest

import workloads
from oracle_lib import ROOT, have_reference, write_fasta

pytestmark = pytest.mark.gpu
MBG = os.path.join(ROOT, "oracle", "_ref", "MBG")
MBG_B200 = os.path.join(ROOT, "oracle", "_ref", "MBG_b200")


def run(exe, fa, out, args):
    r = subprocess.run([exe, "-i", str(fa), "-o", str(out) + ".gfa", "--output-sequence-paths", str(out) + ".gaf"] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    return open(str(out) + ".gfa", "rb").read(), open(str(out) + ".gaf", "rb").read(), r.stderr.decode()


@pytest.mark.skipif(not (have_reference() and os.path.exists(MBG_B200)), reason="oracle/_ref/MBG_b200 not built (make -C oracle ref dropin)")
@pytest.mark.parametrize("k,w,extra", [
    (1001, 100, ["-r", "15000", "-R", "4000", "--hpc-variant-onecopy-coverage", "8", "--do-unsafe-guesswork-resolutions"]),   # Verkko's flag set (onecopy coverage scaled to the 15x test set)
    (301, 30, ["-r", "4000", "-R", "1000", "--do-unsafe-guesswork-resolutions"]),
    (301, 30, ["-r", "0"]),
])
def test_mbg_b200_writes_the_files_mbg_writes(tmp_path, k, w, extra):
    bases, offsets = workloads.diploid_msat_reads(len_min=6000 if k < 1001 else 11000, len_max=10000 if k < 1001 else 19000)
    fa = tmp_path / "hifi-corrected.fasta"
    write_fasta(fa, bases, offsets)
    args = ["-t", "1", "-k", str(k), "-w", str(w), "--kmer-abundance", "1", "--unitig-abundance", "2", "--error-masking=collapse-msat"] + extra
    want_gfa, want_gaf, _ = run(MBG, fa, tmp_path / "ref", args)
    got_gfa, got_gaf, log = run(MBG_B200, fa, tmp_path / "b200", args)
    assert "reference path" not in log, "MBG_b200 fell back to the reference's runMBG"
    assert "total selected k-mers in reads" in log and "distinct selected k-mers in unitigs after filtering" in log and "nodes: " in log
    assert hashlib.md5(got_gfa).hexdigest() == hashlib.md5(want_gfa).hexdigest(), "GFA differs from MBG -t 1"
    assert hashlib.md5(got_gaf).hexdigest() == hashlib.md5(want_gaf).hexdigest(), "GAF differs from MBG -t 1"
    # a gzipped copy goes the same way (Verkko's input is hifi-corrected.fasta.gz)
    import gzip
    gz = tmp_path / "hifi-corrected.fasta.gz"
    with gzip.open(gz, "wb", compresslevel=1) as f:
        f.write(fa.read_bytes())
    gz_gfa, gz_gaf, _ = run(MBG_B200, gz, tmp_path / "b200gz", args)
    assert gz_gfa == want_gfa and gz_gaf == want_gaf
    # unreadable input: exit code 1 like MBG (MBG.cpp:1542-1549)
    r = subprocess.run([MBG_B200, "-i", str(tmp_path / "missing.fa"), "-o", str(tmp_path / "x.gfa"), "-k", "301", "-w", "30"], stderr=subprocess.PIPE)
    assert r.returncode == 1
