"""ctypes access to the CPU checkers (TEST INFRASTRUCTURE; product code never imports this).

* ``Oracle``    -- oracle/libmbg_oracle.so, our plain-C restatement of the reference front end.
* ``Reference`` -- oracle/_ref/libmbgref.so, the real reference objects (compiled from /root/reference by
  oracle/Makefile) behind oracle/ref_shim.cpp.  Present only where it was built; it travels to the GPU box as a
  prebuilt file.
Both return the same flat numpy arrays so results can be compared with ``compare_runs``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libmbg_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libmbgref.so")

MODES = {"no": 0, "hpc": 1, "collapse": 2, "dinuc": 3, "msat": 4, "collapse-dinuc": 5, "collapse-msat": 6}

PART_DTYPE = np.dtype([
    ("read_idx", "<u4"), ("part_offset", "<u4"), ("raw_len", "<u4"), ("n_codes", "<u4"),
    ("code_begin", "<u8"), ("pose_begin", "<u8"), ("sel_begin", "<u8"), ("n_sel", "<u4"), ("hashed", "<u4")])
assert PART_DTYPE.itemsize == 48


def build_oracle():
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("mbg_oracle.c", "mbg_unitig_oracle.c", "mbg_oracle.h")]
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "libmbg_oracle.so"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


def pack_reads(reads):
    """list of bytes/str -> (uint8 array of concatenated bases, uint64 offsets[n+1])"""
    bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
    offsets = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        offsets[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    bases = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, dtype=np.uint8)
    return bases, offsets


def _arr(ptr, count, dtype):
    dtype = np.dtype(dtype)
    if count == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).copy()


class RunResult(dict):
    """parts, codes, poses, positions, hashes[n,2](lo,hi), node_hash[n,2], node_cov, node_tip, edges[n,6]"""
    __getattr__ = dict.__getitem__


def _collect(count, data, with_table):
    res = RunResult()
    res["parts"] = _arr(data(0), count(0), PART_DTYPE)
    res["codes"] = _arr(data(1), count(1), "<u2")
    res["poses"] = _arr(data(2), count(2), "<u4")
    res["positions"] = _arr(data(3), count(3), "<u4")
    res["hashes"] = _arr(data(4), count(3) * 2, "<u8").reshape(-1, 2)
    if with_table:
        res["node_hash"] = _arr(data(5), count(4) * 2, "<u8").reshape(-1, 2)
        res["node_cov"] = _arr(data(6), count(4), "<u4")
        res["node_tip"] = _arr(data(7), count(4), "u1")
        res["edges"] = _arr(data(8), count(5) * 6, "<u4").reshape(-1, 6)
        res["total_selected"] = int(count(6))
    return res


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_uint32, C.c_uint32]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_add_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_add_reads.restype = C.c_int
        L.orc_build_table.argtypes = [C.c_void_p]
        L.orc_set_variants.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_count.argtypes = [C.c_void_p, C.c_int]
        L.orc_count.restype = C.c_uint64
        L.orc_data.argtypes = [C.c_void_p, C.c_int]
        L.orc_data.restype = C.c_void_p
        L.orc_char_hash.argtypes = [C.c_uint64]
        L.orc_char_hash.restype = C.c_uint64
        L.orc_complement.argtypes = [C.c_uint16]
        L.orc_complement.restype = C.c_uint16
        L.orc_code_motif_length.argtypes = [C.c_uint16]
        L.orc_code_motif_length.restype = C.c_uint32
        L.orc_max_code.restype = C.c_uint32
        L.orc_murmur64a.argtypes = [C.c_void_p, C.c_uint64]
        L.orc_murmur64a.restype = C.c_uint64
        L.orc_kmer_hash.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_rolling_hash.argtypes = [C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        for f in (L.orc_msat, L.orc_msat_scanform):
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
            f.restype = C.c_uint64
        L.orc_syncmers_bruteforce.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_syncmers_bruteforce.restype = C.c_uint64

    def run(self, mode, k, w, reads, table=True, variants=None):
        bases, offsets = reads if isinstance(reads, tuple) else pack_reads(reads)
        L = self.lib
        h = L.orc_create(MODES[mode] if isinstance(mode, str) else mode, k, w)
        try:
            if variants is not None:
                v = variants
                L.orc_set_variants(h, v.hashes.ctypes.data, v.offsets.ctypes.data, v.len_begin.ctypes.data, v.lengths.ctypes.data, len(v.offsets))
            rc = L.orc_add_reads(h, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1)
            if rc != 0:
                raise RuntimeError("oracle: palindromic k-mer" if rc == -1 else "oracle: unhashable k-mer")
            if table:
                L.orc_build_table(h)
            return _collect(lambda w_: L.orc_count(h, w_), lambda w_: L.orc_data(h, w_), table)
        finally:
            L.orc_free(h)

    def run_unitigs(self, mode, k, w, reads, min_cov, min_unitig_cov, variants=None, one_cov=False, paths=False, thin=False):
        """table + the filter/unitig step after it -> (RunResult, UnitigResult); one_cov = getUnitigGraph's oneCovHeuristic"""
        bases, offsets = reads if isinstance(reads, tuple) else pack_reads(reads)
        L = self.lib
        L.orc_build_unitigs_ex.restype = C.c_void_p
        L.orc_build_unitigs_ex.argtypes = [C.c_void_p, C.c_uint32, C.c_double, C.c_int]
        L.orc_unitigs_free.argtypes = [C.c_void_p]
        L.orc_unitigs_count.argtypes = [C.c_void_p, C.c_int]
        L.orc_unitigs_count.restype = C.c_uint64
        L.orc_unitigs_data.argtypes = [C.c_void_p, C.c_int]
        L.orc_unitigs_data.restype = C.c_void_p
        h = L.orc_create(MODES[mode] if isinstance(mode, str) else mode, k, w)
        u = None
        try:
            if variants is not None:
                v = variants
                L.orc_set_variants(h, v.hashes.ctypes.data, v.offsets.ctypes.data, v.len_begin.ctypes.data, v.lengths.ctypes.data, len(v.offsets))
            if L.orc_add_reads(h, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1) != 0:
                raise RuntimeError("oracle: palindromic or unhashable k-mer")
            L.orc_build_table(h)
            table = _collect(lambda w_: L.orc_count(h, w_), lambda w_: L.orc_data(h, w_), True)
            u = L.orc_build_unitigs_ex(h, min_cov, float(min_unitig_cov), 1 if one_cov else 0)
            if thin:       # filterKmersToUnitigKmers, MBG's default within-unitig thinning
                L.orc_thin_unitigs.argtypes = [C.c_void_p, C.c_uint32]
                L.orc_thin_unitigs(u, k)
            cnt, dat = (lambda w_: L.orc_unitigs_count(u, w_)), (lambda w_: L.orc_unitigs_data(u, w_))
            res = UnitigResult(kept=_arr(dat(0), cnt(0), "u1"), node_hash=_arr(dat(1), cnt(1) * 2, "<u8").reshape(-1, 2), node_cov=_arr(dat(2), cnt(1), "<u4"),
                               node_tip=_arr(dat(3), cnt(1), "u1"), edges=_arr(dat(4), cnt(2) * 6, "<u4").reshape(-1, 6),
                               unitig_begin=_arr(dat(5), cnt(3) + 1, "<u8"), unitig_nodes=_arr(dat(6), cnt(4), "<u4"),
                               unitig_edges=_arr(dat(7), cnt(5) * 6, "<u4").reshape(-1, 6), overlaps=_arr(dat(8), cnt(6) * 5, "<u4").reshape(-1, 5))
            if paths:      # row f2: getReadPaths over the same run
                L.orc_read_paths.restype = C.c_void_p
                L.orc_read_paths.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
                L.orc_paths_free.argtypes = [C.c_void_p]
                L.orc_paths_count.argtypes = [C.c_void_p, C.c_int]
                L.orc_paths_count.restype = C.c_uint64
                L.orc_paths_data.argtypes = [C.c_void_p, C.c_int]
                L.orc_paths_data.restype = C.c_void_p
                rp = L.orc_read_paths(h, u, k)
                try:
                    res["paths"] = ReadPaths(heads=_arr(L.orc_paths_data(rp, 0), L.orc_paths_count(rp, 0) * 6, "<u8").reshape(-1, 6),
                                             path=_arr(L.orc_paths_data(rp, 1), L.orc_paths_count(rp, 1), "<u4"), poses=_arr(L.orc_paths_data(rp, 2), L.orc_paths_count(rp, 2), "<u4"))
                finally:
                    L.orc_paths_free(rp)
            return table, res
        finally:
            if u:
                L.orc_unitigs_free(u)
            L.orc_free(h)

    def char_hash(self, c):
        return int(self.lib.orc_char_hash(c))

    def complement(self, c):
        return int(self.lib.orc_complement(c))

    def kmer_hash(self, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        lo, hi = C.c_uint64(), C.c_uint64()
        self.lib.orc_kmer_hash(codes.ctypes.data, len(codes), C.byref(lo), C.byref(hi))
        return lo.value, hi.value

    def rolling_hash(self, s, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        fw, bw = C.c_uint64(), C.c_uint64()
        self.lib.orc_rolling_hash(s, codes.ctypes.data, len(codes), C.byref(fw), C.byref(bw))
        return fw.value, bw.value

    def _msat(self, fn, sym, poses, max_motif):
        sym = np.ascontiguousarray(sym, dtype=np.uint16)
        poses = np.ascontiguousarray(poses, dtype=np.uint64)
        oc = np.zeros(len(sym) + 1, dtype=np.uint16)
        op = np.zeros(len(sym) + 2, dtype=np.uint64)
        n = fn(sym.ctypes.data, poses.ctypes.data, len(sym), max_motif, oc.ctypes.data, op.ctypes.data)
        return oc[:n].copy(), op[:n + 1].copy()

    def msat(self, sym, poses=None, max_motif=6):
        poses = np.arange(len(sym) + 1, dtype=np.uint64) if poses is None else poses
        return self._msat(self.lib.orc_msat, sym, poses, max_motif)

    def msat_scanform(self, sym, poses=None, max_motif=6):
        poses = np.arange(len(sym) + 1, dtype=np.uint64) if poses is None else poses
        return self._msat(self.lib.orc_msat_scanform, sym, poses, max_motif)

    def syncmers_bruteforce(self, codes, k, w):
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        out = np.zeros(max(len(codes), 1), dtype=np.uint32)
        n = self.lib.orc_syncmers_bruteforce(codes.ctypes.data, len(codes), k, w, out.ctypes.data)
        return out[:n].copy()


def have_reference():
    return os.path.exists(REF_SO)


class Reference:
    def __init__(self):
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        L.refshim_run.restype = C.c_void_p
        L.refshim_run.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]
        L.refshim_free.argtypes = [C.c_void_p]
        L.refshim_run_variants.restype = C.c_void_p
        L.refshim_run_variants.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_char_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.refshim_count.argtypes = [C.c_void_p, C.c_int]
        L.refshim_count.restype = C.c_uint64
        L.refshim_data.argtypes = [C.c_void_p, C.c_int]
        L.refshim_data.restype = C.c_void_p
        L.refshim_max_code.restype = C.c_uint64
        L.refshim_complement.argtypes = [C.c_uint16]
        L.refshim_complement.restype = C.c_uint16
        L.refshim_code_motif_length.argtypes = [C.c_uint16]
        L.refshim_code_motif_length.restype = C.c_uint64
        L.refshim_char_hash.argtypes = [C.c_uint64]
        L.refshim_char_hash.restype = C.c_uint64
        L.refshim_fasthasher.argtypes = [C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.refshim_kmer_hash.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.refshim_msat.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
        L.refshim_msat.restype = C.c_uint64

    def run(self, mode, k, w, reads, table=True):
        bases, offsets = reads if isinstance(reads, tuple) else pack_reads(reads)
        L = self.lib
        h = L.refshim_run(MODES[mode] if isinstance(mode, str) else mode, k, w, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1, 1 if table else 0)
        try:
            return _collect(lambda w_: L.refshim_count(h, w_), lambda w_: L.refshim_data(h, w_), table)
        finally:
            L.refshim_free(h)

    def run_unitigs(self, mode, k, w, reads, min_cov, min_unitig_cov, one_cov=False, pass3=False):
        """the reference's own getUnitigGraph after its table build (ids = filtered arrival order) -> UnitigResult"""
        bases, offsets = reads if isinstance(reads, tuple) else pack_reads(reads)
        L = self.lib
        L.refshim_unitigs_ex.restype = C.c_void_p
        L.refshim_unitigs_ex.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double, C.c_int]
        h = L.refshim_unitigs_ex(MODES[mode] if isinstance(mode, str) else mode, k, w, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1, min_cov, float(min_unitig_cov),
                                 (1 if one_cov else 0) | (2 if pass3 else 0))      # bit 1: the rest of runMBG's pass 3 (MBG.cpp:1575-1581)
        try:
            r = _collect(lambda w_: L.refshim_count(h, w_), lambda w_: L.refshim_data(h, w_), True)
            begin = np.concatenate([r.parts["sel_begin"], [len(r.positions)]]).astype(np.uint64) if len(r.parts) else np.zeros(1, np.uint64)
            ue = _arr(L.refshim_data(h, 9), L.refshim_count(h, 7) * 6, "<u4").reshape(-1, 6)
            paths = ReadPaths(heads=_arr(L.refshim_data(h, 10), L.refshim_count(h, 8) * 6, "<u8").reshape(-1, 6), path=_arr(L.refshim_data(h, 11), L.refshim_count(h, 9), "<u4"),
                              poses=_arr(L.refshim_data(h, 12), L.refshim_count(h, 10), "<u4"))
            return UnitigResult(kept=None, node_hash=r.node_hash, node_cov=r.node_cov, node_tip=r.node_tip, edges=r.edges, unitig_begin=begin, unitig_nodes=r.positions,
                                unitig_edges=ue, paths=paths, overlaps=_arr(L.refshim_data(h, 13), L.refshim_count(h, 11) * 5, "<u4").reshape(-1, 5))
        finally:
            L.refshim_free(h)

    def run_variants(self, mode, k, w, reads, variants, scratch):
        """pass 3 of runMBG: file-driven iterator with hpc variants; parts = hashed parts in file order"""
        bases, offsets = reads if isinstance(reads, tuple) else pack_reads(reads)
        L, v = self.lib, variants
        h = L.refshim_run_variants(MODES[mode] if isinstance(mode, str) else mode, k, w, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1,
                                   str(scratch).encode(), v.hashes.ctypes.data, v.offsets.ctypes.data, v.len_begin.ctypes.data, v.lengths.ctypes.data, len(v.offsets))
        try:
            return _collect(lambda w_: L.refshim_count(h, w_), lambda w_: L.refshim_data(h, w_), True)
        finally:
            L.refshim_free(h)

    def char_hash(self, c):
        return int(self.lib.refshim_char_hash(c))

    def complement(self, c):
        return int(self.lib.refshim_complement(c))

    def kmer_hash(self, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        lo, hi = C.c_uint64(), C.c_uint64()
        self.lib.refshim_kmer_hash(codes.ctypes.data, len(codes), C.byref(lo), C.byref(hi))
        return lo.value, hi.value

    def rolling_hash(self, s, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        fw, bw = C.c_uint64(), C.c_uint64()
        self.lib.refshim_fasthasher(s, codes.ctypes.data, len(codes), C.byref(fw), C.byref(bw))
        return fw.value, bw.value

    def msat(self, sym, poses=None, max_motif=6):
        sym = np.ascontiguousarray(sym, dtype=np.uint16)
        poses = np.arange(len(sym) + 1, dtype=np.uint64) if poses is None else np.ascontiguousarray(poses, dtype=np.uint64)
        oc = np.zeros(len(sym) + 1, dtype=np.uint16)
        op = np.zeros(len(sym) + 2, dtype=np.uint64)
        n = self.lib.refshim_msat(sym.ctypes.data, poses.ctypes.data, len(sym), max_motif, oc.ctypes.data, op.ctypes.data)
        return oc[:n].copy(), op[:n + 1].copy()


class ReadPaths(dict):
    """getReadPaths' result, flat: heads[n, 6] = (read index, part offset, path entries, readPoses entries, leftClip,
    rightClip), path[] = 2*unitig + fw, poses[] = readPoses"""
    __getattr__ = dict.__getitem__


def expand_paths(rp, u):
    """numbering-free form of ReadPaths: every path as (read, part offset, oriented k-mers it covers as (hash lo, hash hi,
    fw), readPoses).  Unitig numbering, orientation and where a circular unitig was cut all drop out."""
    b = u.unitig_begin.astype(np.int64)
    out = []
    at_p = at_q = 0
    for read, off, n_path, n_poses, left, right in rp.heads.astype(np.int64).tolist():
        seq = []
        for x in rp.path[at_p:at_p + n_path].tolist():
            nodes = u.unitig_nodes[b[x >> 1]:b[(x >> 1) + 1]].astype(np.int64)
            seq.append(nodes if x & 1 else (nodes[::-1] ^ 1))
        seq = np.concatenate(seq) if seq else np.zeros(0, np.int64)
        seq = seq[left:len(seq) - right]
        kmers = tuple(zip(u.node_hash[seq >> 1, 0].tolist(), u.node_hash[seq >> 1, 1].tolist(), (seq & 1).tolist()))
        out.append((read, off, kmers, tuple(rp.poses[at_q:at_q + n_poses].tolist())))
        at_p += n_path
        at_q += n_poses
    return out


def compare_paths(a, ua, b, ub, what=""):
    """ReadPaths a over UnitigResult ua against b over ub, in order (paths come out in input order on one thread)"""
    ea, eb = expand_paths(a, ua), expand_paths(b, ub)
    assert len(ea) == len(eb), f"{what}: {len(ea)} read paths vs {len(eb)}"
    for i, (x, y) in enumerate(zip(ea, eb)):
        assert x == y, f"{what}: read path {i} differs (read {x[0]} part {x[1]}: {len(x[2])} k-mers / {len(x[3])} poses vs read {y[0]} part {y[1]}: {len(y[2])} / {len(y[3])})"


class UnitigResult(dict):
    """kept[u8 per table node] (None for the reference), node_hash/node_cov/node_tip/edges of the filtered list,
    unitig_begin[u+1], unitig_nodes[2*filtered id + fw]"""
    __getattr__ = dict.__getitem__


def normalise_unitigs(u):
    """numbering-free form of a UnitigResult: nodes renumbered by ascending canonical hash, edges canonical in that
    numbering and sorted, unitigs as direction-normalised tuples of (2*id + fw)."""
    order = np.lexsort((u.node_hash[:, 0], u.node_hash[:, 1]))
    new_id = np.empty(len(order), dtype=np.int64)
    new_id[order] = np.arange(len(order))
    e = u.edges.astype(np.int64).copy()
    if len(e):
        e[:, 0] = new_id[e[:, 0]]
        e[:, 2] = new_id[e[:, 2]]
        swap = (e[:, 2] < e[:, 0]) | ((e[:, 2] == e[:, 0]) & (e[:, 1] == 0) & (e[:, 3] == 0))
        f, ffw, t, tfw = e[swap, 0].copy(), e[swap, 1].copy(), e[swap, 2].copy(), e[swap, 3].copy()
        e[swap, 0], e[swap, 1], e[swap, 2], e[swap, 3] = t, 1 - tfw, f, 1 - ffw
        e = sorted_edges(e)
    seqs, flipped, name = [], [], []
    b = u.unitig_begin.astype(np.int64)
    for i in range(len(b) - 1):
        raw = u.unitig_nodes[b[i]:b[i + 1]].astype(np.int64)
        seq = tuple((2 * new_id[raw >> 1] + (raw & 1)).tolist())
        rev = tuple((x ^ 1) for x in reversed(seq))
        seqs.append(min(seq, rev))
        flipped.append(rev < seq)
        name.append(min(x >> 1 for x in seq))           # a unitig is named by its smallest node: nodes sit in one unitig only
    res = dict(node_hash=u.node_hash[order], node_cov=u.node_cov[order], node_tip=u.node_tip[order], edges=e, unitigs=sorted(seqs))
    if "unitig_edges" in u:
        # an oriented unitig end is named by the oriented k-mer at that end (numbering- and rotation-free)
        ue = set()
        for r in u.unitig_edges.astype(np.int64):
            def end_node(t, fw, last):
                raw = u.unitig_nodes[b[t]:b[t + 1]].astype(np.int64)
                seq = [int(2 * new_id[x >> 1] + (x & 1)) for x in raw]
                if not fw:
                    seq = [x ^ 1 for x in reversed(seq)]
                return seq[-1] if last else seq[0]
            if r[0] == r[2] and r[1] == r[3]:
                # a circular unitig closing on itself: which of its k-mer edges closes it (and so the coverage shown)
                # depends on where the walk started, i.e. on the node numbering; named by its smallest node only
                ue.add((-1, name[r[0]], 0))
                continue
            x, y = end_node(r[0], r[1], True), end_node(r[2], r[3], False)
            ue.add(min((x, y, int(r[4])), (y ^ 1, x ^ 1, int(r[4]))))
        res["unitig_edges"] = sorted(ue)
    return res


def compare_unitigs(a, b, what=""):
    """a, b: UnitigResults; equal up to node / unitig numbering (circular unitigs also up to rotation)"""
    na, nb = normalise_unitigs(a), normalise_unitigs(b)
    for f in ("node_hash", "node_cov", "node_tip", "edges"):
        assert na[f].shape == nb[f].shape and np.array_equal(na[f], nb[f]), f"{what}: filtered {f} differ ({na[f].shape} vs {nb[f].shape})"
    if na["unitigs"] != nb["unitigs"]:
        def rot(seq):
            out = []
            for s in (seq, tuple((x ^ 1) for x in reversed(seq))):
                i = min(range(len(s)), key=lambda j: s[j] >> 1)
                out.append(s[i:] + s[:i])
            return min(out)
        sa, sb = set(na["unitigs"]), set(nb["unitigs"])
        links = set((int(r[0]) * 2 + int(r[1]), int(r[2]) * 2 + int(r[3])) for r in na["edges"]) | set(((int(r[2]) * 2 + int(r[3])) ^ 1, (int(r[0]) * 2 + int(r[1])) ^ 1) for r in na["edges"])
        for s in (sa ^ sb):
            assert (s[-1], s[0]) in links, f"{what}: unitig differs and is not circular: {s[:6]}... len {len(s)}"
        assert sorted(rot(s) for s in sa - sb) == sorted(rot(s) for s in sb - sa), f"{what}: circular unitigs differ"
    assert len(na["unitigs"]) == len(nb["unitigs"]), f"{what}: unitig count"
    if "unitig_edges" in na and "unitig_edges" in nb:
        assert na["unitig_edges"] == nb["unitig_edges"], f"{what}: unitig edges differ ({len(na['unitig_edges'])} vs {len(nb['unitig_edges'])})"


class Variants:
    """entry e == one ReadpartIterator::addHpcVariants(hash e, offsets[e], lengths[len_begin[e]:len_begin[e+1]]) call"""

    def __init__(self, hashes, offsets, len_begin, lengths):
        self.hashes = np.ascontiguousarray(hashes, dtype=np.uint64).reshape(-1, 2)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        self.len_begin = np.ascontiguousarray(len_begin, dtype=np.uint64)
        self.lengths = np.ascontiguousarray(lengths, dtype=np.uint32)


def make_variants(run, k, n_kmers, seed=5, per_kmer=3):
    """hpc-variant entries for n_kmers of the k-mers selected in `run` (a RunResult without variants): a few code offsets
    per k-mer (below, at and above k/2), each with a sorted list of run-length clusters that covers every run length seen
    in the reads.  Half of the k-mers are registered under the swapped hash, so both branches of
    MBG/ReadHelper.h:690-727 are taken; a k-mer is registered under one orientation only (the reference asserts that)."""
    rng = np.random.default_rng(seed)
    n_sel = len(run.positions)
    pick = rng.choice(n_sel, size=min(n_kmers, n_sel), replace=False)
    canon = {}
    hashes, offsets, len_begin, lengths = [], [], [0], []
    max_len = int(np.max(np.diff(run.poses.astype(np.int64)))) + 2 if len(run.poses) > 1 else 8
    for i, idx in enumerate(pick):
        lo, hi = int(run.hashes[idx, 0]), int(run.hashes[idx, 1])
        key = (min(lo, hi), max(lo, hi))
        if key in canon or lo == hi:
            continue
        canon[key] = True
        reg = (lo, hi) if i % 2 == 0 else (hi, lo)
        offs = {int(rng.integers(0, k // 2)), k // 2, int(rng.integers(k // 2 + 1, k))} if per_kmer >= 3 else {int(rng.integers(0, k))}
        for off in sorted(offs):
            hashes.append(reg)
            offsets.append(off)
            cl = sorted(set(int(x) for x in rng.integers(1, max_len, size=int(rng.integers(1, 4)))) | {max_len})
            rng.shuffle(cl)
            lengths.extend(cl)
            len_begin.append(len(lengths))
    return Variants(np.array(hashes, dtype=np.uint64).reshape(-1, 2), offsets, len_begin, lengths)


def sorted_edges(edges):
    if len(edges) == 0:
        return edges
    order = np.lexsort((edges[:, 3], edges[:, 2], edges[:, 1], edges[:, 0]))
    return edges[order]


def compare_runs(a, b, table=True, what=""):
    """assert that two RunResults are identical (parts in the same (read, part) order)."""
    assert len(a.parts) == len(b.parts), f"{what}: part count {len(a.parts)} != {len(b.parts)}"
    for f in ("read_idx", "part_offset", "raw_len", "n_codes", "n_sel", "hashed"):
        assert np.array_equal(a.parts[f], b.parts[f]), f"{what}: parts.{f} differ"
    for f in ("codes", "poses", "positions", "hashes"):
        assert a[f].shape == b[f].shape, f"{what}: {f} shape {a[f].shape} != {b[f].shape}"
        assert np.array_equal(a[f], b[f]), f"{what}: {f} differ"
    if table:
        assert np.array_equal(a.node_hash, b.node_hash), f"{what}: node hashes differ"
        assert np.array_equal(a.node_cov, b.node_cov), f"{what}: node coverage differs"
        assert np.array_equal(a.node_tip, b.node_tip), f"{what}: tip flags differ"
        assert np.array_equal(sorted_edges(a.edges), sorted_edges(b.edges)), f"{what}: edges differ"


def lcg_bases(seed, n):
    """SURVEY.md 8(c) generator: x <- x*6364136223846793005 + 1442695040888963407, base = ACGT[(x>>33)&3]."""
    out = bytearray(n)
    x = seed
    for i in range(n):
        x = (x * 6364136223846793005 + 1442695040888963407) & 0xFFFFFFFFFFFFFFFF
        out[i] = b"ACGT"[(x >> 33) & 3]
    return bytes(out)


DIGEST_FIELDS = ("selected", "distinct", "distinct_edges", "node_xor", "node_sum", "tips", "edge_cov", "edge_sum", "edge_xor", "overlap_sum")


def table_digest(node_hash, node_cov, node_tip, edges):
    """order-free digest of a k-mer table from its exported arrays; the same ten words as oracle/ref_shim.cpp:hashlist_digest
    computes from the reference's HashList (mod 2^64)"""
    with np.errstate(over="ignore"):
        nh = np.asarray(node_hash, dtype=np.uint64).reshape(-1, 2)
        cov = np.asarray(node_cov).astype(np.uint64)
        mix = nh[:, 0] * np.uint64(0x9E3779B97F4A7C15) + nh[:, 1] * np.uint64(0xC2B2AE3D27D4EB4F)
        e = np.asarray(edges).reshape(-1, 6).astype(np.uint64)
        three, five, seven, eleven, one, two = (np.uint64(x) for x in (3, 5, 7, 11, 1, 2))

        def v(ids, fw):
            return mix[ids.astype(np.int64)] * np.where(fw == one, three, five)
        out = dict.fromkeys(DIGEST_FIELDS, 0)
        out["selected"] = int(cov.sum(dtype=np.uint64))
        out["distinct"] = len(cov)
        out["distinct_edges"] = len(e)
        out["node_xor"] = int(np.bitwise_xor.reduce(mix * (two * cov + one))) if len(cov) else 0
        out["node_sum"] = int((mix * cov).sum(dtype=np.uint64))
        out["tips"] = int(np.asarray(node_tip).astype(np.uint64).sum())
        if len(e):
            sym = (v(e[:, 0], e[:, 1]) * seven + v(e[:, 2], e[:, 3]) * eleven) + (v(e[:, 2], one - e[:, 3]) * seven + v(e[:, 0], one - e[:, 1]) * eleven)
            out["edge_cov"] = int(e[:, 4].sum(dtype=np.uint64))
            out["edge_sum"] = int((sym * e[:, 4]).sum(dtype=np.uint64))
            out["edge_xor"] = int(np.bitwise_xor.reduce(sym * (two * e[:, 4] + one)))
            out["overlap_sum"] = int((sym * e[:, 5]).sum(dtype=np.uint64))
        return out


def reference_file_digest(mode, k, w, fasta_path, threads):
    """(seconds, digest dict) of the reference's loadReadsAsHashesMultithread over a FASTA/FASTQ file"""
    lib = C.CDLL(REF_SO)
    lib.refshim_time_file_digest.restype = C.c_double
    lib.refshim_time_file_digest.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_char_p, C.c_uint64, C.c_void_p]
    d = (C.c_uint64 * 10)()
    t = lib.refshim_time_file_digest(MODES[mode] if isinstance(mode, str) else mode, k, w, str(fasta_path).encode(), threads, d)
    return t, dict(zip(DIGEST_FIELDS, [int(x) for x in d]))


def write_fasta(path, bases, offsets, line=0):
    """reads -> FASTA (one sequence line per read unless line > 0)"""
    with open(path, "wb") as f:
        for i in range(len(offsets) - 1):
            seq = bases[int(offsets[i]):int(offsets[i + 1])].tobytes()
            f.write(b">r%d\n" % i)
            if line:
                for j in range(0, len(seq), line):
                    f.write(seq[j:j + line] + b"\n")
            else:
                f.write(seq + b"\n")


REF_MBG = os.path.join(ROOT, "oracle", "_ref", "MBG")


def reference_cli(fasta, out_gfa, out_gaf, mode, k, w, a, u, r=0, big_r=0, guesswork=False, threads=1, extra=()):
    """the reference's own executable (oracle/_ref/MBG, built by oracle/Makefile) -> (gfa bytes, gaf bytes)"""
    cmd = [REF_MBG, "-i", str(fasta), "-o", str(out_gfa), "-k", str(k), "-w", str(w), "-a", str(a), "-u", str(u), "-t", str(threads),
           "--error-masking=" + mode, "--output-sequence-paths", str(out_gaf), "-r", str(r), "-R", str(big_r)]
    if guesswork:
        cmd.append("--do-unsafe-guesswork-resolutions")
    cmd += list(extra)
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return open(out_gfa, "rb").read(), open(out_gaf, "rb").read()


def reference_hpc_variants(fasta, mode, k, w, a, u, onecopy_coverage):
    """passes 1 and 2 of runMBG under --hpc-variant-onecopy-coverage, run by the reference -> Variants (row f3 is not built)"""
    lib = C.CDLL(REF_SO)
    lib.refshim_hpc_variants.restype = C.c_uint64
    lib.refshim_hpc_variants.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_char_p, C.c_uint64, C.c_double, C.c_double]
    lib.refshim_hpc_variants_data.restype = C.c_void_p
    lib.refshim_hpc_variants_data.argtypes = [C.c_int]
    lib.refshim_hpc_variants_lengths.restype = C.c_uint64
    saved = os.dup(2)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 2)
    try:
        n = lib.refshim_hpc_variants(MODES[mode], k, w, str(fasta).encode(), a, float(u), float(onecopy_coverage))
    finally:
        os.dup2(saved, 2)
        os.close(saved)
        os.close(devnull)
    return Variants(_arr(lib.refshim_hpc_variants_data(0), 2 * n, "<u8").reshape(-1, 2), _arr(lib.refshim_hpc_variants_data(1), n, "<u4"),
                    _arr(lib.refshim_hpc_variants_data(2), n + 1, "<u8"), _arr(lib.refshim_hpc_variants_data(3), lib.refshim_hpc_variants_lengths(), "<u4"))


def reference_backend(stage, fasta, out_gfa, out_gaf, mode, k, w, a, u, r=0, big_r=0, guesswork=False, *, node_hash, node_cov, node_tip, edges,
                      new_overlaps=None, unitig_begin=None, unitig_nodes=None, unitig_edges=None, variants=None):
    """T4: tables (stage 0) or the thinned list + unitigs (stage 1) handed to the reference's own back end
    (oracle/ref_shim.cpp:refshim_backend) -> (gfa bytes, gaf bytes)"""
    lib = C.CDLL(REF_SO)
    f = lib.refshim_backend_variants
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_char_p, C.c_uint64, C.c_double, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int,
                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                  C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_char_p, C.c_char_p,
                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    v = variants if variants is not None else Variants(np.zeros((0, 2)), [], [0], [])
    nh = np.ascontiguousarray(node_hash, dtype=np.uint64)
    nc = np.ascontiguousarray(node_cov, dtype=np.uint32)
    nt = np.ascontiguousarray(node_tip, dtype=np.uint8)
    ed = np.ascontiguousarray(edges, dtype=np.uint32).reshape(-1, 6)
    no = np.ascontiguousarray(new_overlaps if new_overlaps is not None else np.zeros((0, 6)), dtype=np.uint32).reshape(-1, 6)
    ub = np.ascontiguousarray(unitig_begin if unitig_begin is not None else np.zeros(1), dtype=np.uint64)
    un = np.ascontiguousarray(unitig_nodes if unitig_nodes is not None else np.zeros(0), dtype=np.uint32)
    ue = np.ascontiguousarray(unitig_edges if unitig_edges is not None else np.zeros((0, 6)), dtype=np.uint32).reshape(-1, 6)
    saved = os.dup(2)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 2)
    try:
        rc = f(stage, MODES[mode], k, w, str(fasta).encode(), a, float(u), r, big_r, 1 if guesswork else 0, 0, 0, 1,
               nh.ctypes.data, nc.ctypes.data, nt.ctypes.data, len(nc), ed.ctypes.data, len(ed),
               no.ctypes.data, len(no), ub.ctypes.data, un.ctypes.data, len(ub) - 1, ue.ctypes.data, len(ue), str(out_gfa).encode(), str(out_gaf).encode(),
               v.hashes.ctypes.data, v.offsets.ctypes.data, v.len_begin.ctypes.data, v.lengths.ctypes.data, len(v.offsets))
    finally:
        os.dup2(saved, 2)
        os.close(saved)
        os.close(devnull)
    assert rc == 0
    return open(out_gfa, "rb").read(), open(out_gaf, "rb").read()


def sort_unitigs(u):
    """UnitigGraph::sort (MBG/UnitigGraph.cpp:278-438) on a UnitigResult whose node ids are in hash order: every unitig oriented so
    that its first node id is below its last one (a single k-mer: forward), unitigs ordered by that first id, unitig edges
    renumbered and put back into canon() form.  The device numbers its unitigs this way already; the oracle's come in walk order."""
    b = u.unitig_begin.astype(np.int64)
    n = len(b) - 1
    seqs, flip = [], []
    for i in range(n):
        s = u.unitig_nodes[b[i]:b[i + 1]].astype(np.int64)
        f = (s[-1] >> 1) < (s[0] >> 1) if len(s) > 1 else not (s[0] & 1)
        seqs.append((s[::-1] ^ 1) if f else s)
        flip.append(bool(f))
    order = sorted(range(n), key=lambda i: int(seqs[i][0] >> 1))
    new_id = np.empty(n, np.int64)
    new_id[order] = np.arange(n)
    nodes = np.concatenate([seqs[i] for i in order]).astype(np.uint32) if n else np.zeros(0, np.uint32)
    begin = np.zeros(n + 1, np.uint64)
    begin[1:] = np.cumsum([len(seqs[i]) for i in order])
    e = u.unitig_edges.astype(np.int64).copy()
    for r in e:
        f, ffw, t, tfw = int(r[0]), int(r[1]), int(r[2]), int(r[3])
        ffw, tfw = ffw ^ flip[f], tfw ^ flip[t]
        f, t = int(new_id[f]), int(new_id[t])
        if t < f or (t == f and not tfw and not ffw):
            f, ffw, t, tfw = t, 1 - tfw, f, 1 - ffw
        r[0], r[1], r[2], r[3] = f, ffw, t, tfw
    out = UnitigResult(u)
    out["unitig_begin"], out["unitig_nodes"], out["unitig_edges"] = begin, nodes, e.astype(np.uint32)
    return out
