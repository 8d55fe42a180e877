"""ctypes mirror of include/mbgfe.h.

Mirrors the reference seam ``ReadpartIterator`` + ``loadReadsAsHashesMultithread`` (MBG/ReadHelper.h:162-241,
MBG/MBG.cpp:44-76): construct with (k, w, error masking), add reads, run, read the per-part arrays and the table.
The library is CUDA only; loading or creating a handle fails loudly when the extension or the GPU is missing.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MODES = {"no": 0, "hpc": 1, "collapse": 2, "dinuc": 3, "msat": 4, "collapse-dinuc": 5, "collapse-msat": 6}

KEEP_CODES, KEEP_POSES, NO_TABLE, NO_EDGES = 1, 2, 4, 8
PARTS, CODES, POSES, POSITIONS, HASHES, NODE_HASHES, NODE_COVERAGE, NODE_TIP, EDGES = range(9)
NODE_KEPT, F_NODE_HASHES, F_NODE_COVERAGE, F_NODE_TIP, F_EDGES, UNITIG_BEGIN, UNITIG_NODES, F_NODE_UNITIG, UNITIG_EDGES = range(9, 18)
INST_UNITIG, INST_UNITIG_POS, INST_UNITIG_REST, INST_STEP = range(18, 22)
F_NODE_THIN, F_NODE_UNITIG_OFFSET = 22, 23
T_NODE_HASHES, T_NODE_COVERAGE, T_NODE_TIP, T_EDGES, T_UNITIG_BEGIN, T_UNITIG_NODES, T_NEW_OVERLAPS = range(24, 31)
UNITIG_ONECOV_HEURISTIC = 1      # mbgfe_build_unitigs_ex flag (include/mbgfe.h)

PART_DTYPE = np.dtype([
    ("read_idx", "<u4"), ("part_offset", "<u4"), ("raw_len", "<u4"), ("n_codes", "<u4"),
    ("code_begin", "<u8"), ("pose_begin", "<u8"), ("sel_begin", "<u8"), ("n_sel", "<u4"), ("hashed", "<u4")])


class Params(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("k", C.c_uint32), ("w", C.c_uint32), ("error_masking", C.c_uint32),
                ("kmer_abundance", C.c_uint32), ("flags", C.c_uint32), ("device", C.c_int32), ("rank", C.c_uint32),
                ("world", C.c_uint32), ("piece_bytes", C.c_uint32), ("table_slots", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "reads", "raw_bases", "segments", "collapsed_symbols", "codes", "parts", "hashed_parts", "selected", "distinct",
        "distinct_edges", "table_slots", "table_grows", "nonacgt", "h2d_bytes", "d2h_bytes")] + [(n, C.c_float) for n in (
        "ms_upload", "ms_pack", "ms_collapse", "ms_msat", "ms_select", "ms_hash", "ms_nodes", "ms_edges", "ms_process",
        "ms_export")] + [("launches", C.c_uint32), ("ms_unitigs", C.c_float), ("ms_exchange", C.c_float), ("exchange_launches", C.c_uint32),
                         ("exchange_bytes", C.c_uint64)]


class MbgfeError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"mbgfe error {code}: {text}")
        self.code = code


def library_path():
    """The in-tree CUDA library (no override: tests that want the GPU-less emulator build pass `library=` explicitly)."""
    return os.path.join(HERE, "libmbgfe.so")


_lib = None


def load_library(path=None):
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or library_path()
    if not os.path.exists(path):
        raise MbgfeError(-2, f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(there is no CPU fallback)")
    L = C.CDLL(path)
    L.mbgfe_create.argtypes = [C.POINTER(Params), C.POINTER(C.c_void_p)]
    L.mbgfe_destroy.argtypes = [C.c_void_p]
    L.mbgfe_last_error.argtypes = [C.c_void_p]
    L.mbgfe_last_error.restype = C.c_char_p
    L.mbgfe_alloc_pinned.argtypes = [C.c_uint64]
    L.mbgfe_alloc_pinned.restype = C.c_void_p
    L.mbgfe_free_pinned.argtypes = [C.c_void_p]
    L.mbgfe_add_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    L.mbgfe_add_file.argtypes = [C.c_void_p, C.c_char_p]
    L.mbgfe_read_name.argtypes = [C.c_void_p, C.c_uint64]
    L.mbgfe_read_name.restype = C.c_char_p
    for f in (L.mbgfe_reset, L.mbgfe_upload, L.mbgfe_process, L.mbgfe_run, L.mbgfe_run_append, L.mbgfe_upload_append, L.mbgfe_process_append, L.mbgfe_build_table):
        f.argtypes = [C.c_void_p]
    L.mbgfe_count.argtypes = [C.c_void_p, C.c_int]
    L.mbgfe_count.restype = C.c_uint64
    L.mbgfe_data.argtypes = [C.c_void_p, C.c_int]
    L.mbgfe_data.restype = C.c_void_p
    L.mbgfe_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
    L.mbgfe_timer_start.argtypes = [C.c_void_p]
    L.mbgfe_timer_stop.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    L.mbgfe_comm_unique_id.argtypes = [C.c_void_p]
    L.mbgfe_comm_init_nccl.argtypes = [C.c_void_p, C.c_void_p]
    L.mbgfe_comm_init_callbacks.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.mbgfe_exchange.argtypes = [C.c_void_p]
    L.mbgfe_exchange_group.argtypes = [C.c_void_p, C.c_uint32]
    L.mbgfe_build_unitigs.argtypes = [C.c_void_p, C.c_uint32, C.c_double]
    L.mbgfe_build_unitigs_ex.argtypes = [C.c_void_p, C.c_uint32, C.c_double, C.c_uint32]
    L.mbgfe_thread_reads.argtypes = [C.c_void_p]
    L.mbgfe_thin_unitigs.argtypes = [C.c_void_p]
    L.mbgfe_set_variants.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    L.mbgfe_kat_tables.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.mbgfe_kat_kmer_hash.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    L.mbgfe_kat_kmer_hash_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    if path == library_path():
        _lib = L
    return L


def _array(ptr, count, dtype, copy=True):
    """numpy array over library-owned memory; copy=False gives a view that is valid until the next run on the handle"""
    dtype = np.dtype(dtype)
    if count == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
    a = np.frombuffer(buf, dtype=dtype)
    return a.copy() if copy else a


class PinnedBuffer:
    """numpy view of page-locked host memory from mbgfe_alloc_pinned."""

    def __init__(self, lib, nbytes):
        self.lib, self.nbytes = lib, int(nbytes)
        self.ptr = lib.mbgfe_alloc_pinned(max(self.nbytes, 1))
        if not self.ptr:
            raise MbgfeError(-2, f"cannot allocate {nbytes} bytes of pinned memory")
        self.array = np.frombuffer((C.c_char * max(self.nbytes, 1)).from_address(self.ptr), dtype=np.uint8)[:self.nbytes]

    def free(self):
        if self.ptr:
            self.array = None
            self.lib.mbgfe_free_pinned(self.ptr)
            self.ptr = None


class MbgFrontEnd:
    """ReadpartIterator{k, w, errorMasking} + HashList{k} on one B200."""

    def __init__(self, k, w, error_masking="collapse-msat", kmer_abundance=1, flags=0, device=0, rank=0, world=1,
                 table_slots=0, library=None, piece_bytes=0):
        self.lib = load_library(library)
        mode = MODES[error_masking] if isinstance(error_masking, str) else int(error_masking)
        self.params = Params(2, k, w, mode, kmer_abundance, flags, device, rank, world, piece_bytes, table_slots)
        self.handle = C.c_void_p()
        rc = self.lib.mbgfe_create(C.byref(self.params), C.byref(self.handle))
        if rc != 0:
            raise MbgfeError(rc, self.lib.mbgfe_last_error(None).decode())
        self._keep = []

    def close(self):
        if self.handle:
            self.lib.mbgfe_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise MbgfeError(rc, self.lib.mbgfe_last_error(self.handle).decode())

    # ---- input
    def add_reads(self, bases, offsets):
        """bases: uint8 array (ASCII), offsets: uint64 array of n+1 entries.  Buffers are referenced, not copied."""
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self._keep.append((bases, offsets))
        self._check(self.lib.mbgfe_add_reads(self.handle, bases.ctypes.data, offsets.ctypes.data, len(offsets) - 1))

    def add_read_list(self, reads):
        bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
        offsets = np.zeros(len(bs) + 1, dtype=np.uint64)
        if bs:
            offsets[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
        bases = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, dtype=np.uint8)
        self.add_reads(bases, offsets)

    def add_file(self, path):
        self._check(self.lib.mbgfe_add_file(self.handle, os.fsencode(path)))

    def read_name(self, idx):
        return self.lib.mbgfe_read_name(self.handle, idx).decode()

    def set_variants(self, hashes, offsets, len_begin, lengths):
        """ReadpartIterator::addHpcVariants for every entry at once (MBG/ReadHelper.cpp:63-70); hashes[n,2] = (lo, hi)"""
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64).reshape(-1, 2)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        len_begin = np.ascontiguousarray(len_begin, dtype=np.uint64)
        lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
        self._check(self.lib.mbgfe_set_variants(self.handle, hashes.ctypes.data, offsets.ctypes.data, len_begin.ctypes.data, lengths.ctypes.data, len(offsets)))

    def reset(self):
        self._check(self.lib.mbgfe_reset(self.handle))
        self._keep = []

    # ---- run
    def upload(self):
        self._check(self.lib.mbgfe_upload(self.handle))
        self._keep = []

    def process(self):
        self._check(self.lib.mbgfe_process(self.handle))

    def run(self):
        self._check(self.lib.mbgfe_run(self.handle))
        self._keep = []

    def run_append(self):
        """mbgfe_run on the pending reads, keeping the selected k-mers of earlier batches; tables come from build_table() / exchange()"""
        self._check(self.lib.mbgfe_run_append(self.handle))
        self._keep = []

    def upload_append(self):
        self._check(self.lib.mbgfe_upload_append(self.handle))
        self._keep = []

    def process_append(self):
        self._check(self.lib.mbgfe_process_append(self.handle))

    def build_table(self):
        self._check(self.lib.mbgfe_build_table(self.handle))

    # ---- results
    def count(self, what):
        return int(self.lib.mbgfe_count(self.handle, what))

    def _data(self, what, dtype, per=1, copy=True):
        n = self.count(what)
        ptr = self.lib.mbgfe_data(self.handle, what)
        if n and not ptr:
            self._check(-6)
        return _array(ptr, n * per, dtype, copy)

    def count_selected(self):
        return self.count(POSITIONS)

    def parts(self):
        return self._data(PARTS, PART_DTYPE)

    def codes(self):
        return self._data(CODES, "<u2")

    def poses(self):
        return self._data(POSES, "<u4")

    def positions(self):
        return self._data(POSITIONS, "<u4")

    def hashes(self):
        return self._data(HASHES, "<u8", 2).reshape(-1, 2)

    def node_hashes(self, copy=True):
        return self._data(NODE_HASHES, "<u8", 2, copy).reshape(-1, 2)

    def node_coverage(self, copy=True):
        return self._data(NODE_COVERAGE, "<u4", 1, copy)

    def node_tip(self, copy=True):
        return self._data(NODE_TIP, "u1", 1, copy)

    def edges(self, copy=True):
        return self._data(EDGES, "<u4", 6, copy).reshape(-1, 6)

    def build_unitigs(self, kmer_abundance, unitig_abundance, one_cov_heuristic=False):
        """getUnitigGraph (MBG/MBG.cpp:629-760, no keepGaps) on the device -> dict of the filtered HashList and the
        unitig chains (names as in include/mbgfe.h).  one_cov_heuristic = the function's last argument
        (removeOnecovNodes, MBG/MBG.cpp:527-627; on for runMBG's final graph under Verkko's flags)"""
        self._check(self.lib.mbgfe_build_unitigs_ex(self.handle, int(kmer_abundance), float(unitig_abundance), UNITIG_ONECOV_HEURISTIC if one_cov_heuristic else 0))
        return dict(kept=self._data(NODE_KEPT, "u1"), node_hash=self._data(F_NODE_HASHES, "<u8", 2).reshape(-1, 2),
                    node_cov=self._data(F_NODE_COVERAGE, "<u4"), node_tip=self._data(F_NODE_TIP, "u1"),
                    edges=self._data(F_EDGES, "<u4", 6).reshape(-1, 6), unitig_begin=self._data(UNITIG_BEGIN, "<u8"),
                    unitig_nodes=self._data(UNITIG_NODES, "<u4"), node_unitig=self._data(F_NODE_UNITIG, "<u4"),
                    unitig_edges=self._data(UNITIG_EDGES, "<u4", 6).reshape(-1, 6))

    def thread_reads(self):
        """getReadPaths (MBG/MBG.cpp:1235-1251; KmerMatcher.h:12-144) as a batched lookup -> dict of the per-instance
        arrays of include/mbgfe.h (unitig, unitig_pos, unitig_rest, step), in positions() order"""
        self._check(self.lib.mbgfe_thread_reads(self.handle))
        return dict(unitig=self._data(INST_UNITIG, "<u4"), unitig_pos=self._data(INST_UNITIG_POS, "<u4"), unitig_rest=self._data(INST_UNITIG_REST, "<u4"),
                    step=self._data(INST_STEP, "u1"))

    def thin_unitigs(self):
        """filterKmersToUnitigKmers (MBG/MBG.cpp:1065-1185, filterWithinUnitig = true) -> (kept flag, offset inside the
        unitig) per filtered node"""
        self._check(self.lib.mbgfe_thin_unitigs(self.handle))
        return self._data(F_NODE_THIN, "u1"), self._data(F_NODE_UNITIG_OFFSET, "<u4")

    def thinned(self):
        """the list, unitigs and new sequence overlaps after the thinning was applied (MBG/MBG.cpp:1166-1184); needs thin_unitigs()"""
        return dict(node_hash=self._data(T_NODE_HASHES, "<u8", 2).reshape(-1, 2), node_cov=self._data(T_NODE_COVERAGE, "<u4"), node_tip=self._data(T_NODE_TIP, "u1"),
                    edges=self._data(T_EDGES, "<u4", 6).reshape(-1, 6), unitig_begin=self._data(T_UNITIG_BEGIN, "<u8"), unitig_nodes=self._data(T_UNITIG_NODES, "<u4"),
                    new_overlaps=self._data(T_NEW_OVERLAPS, "<u4", 6).reshape(-1, 6))

    def read_paths(self):
        """the ReadPath list of getReadPaths (MBG/MBG.cpp:1235-1251) put together from thread_reads() the way
        INTEGRATION.md 2c shows: (heads[n, 6] = read index, part offset, path entries, readPoses entries, leftClip,
        rightClip; path[] = 2*unitig + fw; poses[] = readPoses).  Host loop over the instances: meant for inspection and
        tests, the arrays themselves are the product."""
        th = self.thread_reads()
        parts, positions = self.parts(), self.positions()
        unitig, upos, urest, step = th["unitig"], th["unitig_pos"], th["unitig_rest"], th["step"]
        heads, path, poses = [], [], []
        for p in parts:
            if not p["hashed"]:
                continue
            cur_path, cur_poses, left, right = [], [], 0, 0

            def emit():
                heads.append((int(p["read_idx"]), int(p["part_offset"]), len(cur_path), len(cur_poses), left, right))
                path.extend(cur_path)
                poses.extend(cur_poses)
            b = int(p["sel_begin"])
            for i in range(b, b + int(p["n_sel"])):
                s = int(step[i])
                if s == 0:
                    continue
                if s == 1:                       # KmerMatcher.h:30-47: path ends, readPoses stays
                    if cur_path:
                        emit()
                    cur_path, left, right = [], 0, 0
                elif s == 2:
                    cur_path.append(int(unitig[i])); cur_poses.append(int(positions[i])); left, right = int(upos[i]), int(urest[i])
                elif s == 3:
                    emit()
                    cur_path, cur_poses, left, right = [int(unitig[i])], [int(positions[i])], int(upos[i]), int(urest[i])
                elif s == 4:
                    cur_poses.append(int(positions[i])); right = int(urest[i])
                else:
                    cur_path.append(int(unitig[i])); cur_poses.append(int(positions[i])); right = int(urest[i])
            if cur_path:
                emit()
        return dict(heads=np.array(heads, dtype=np.uint64).reshape(-1, 6), path=np.array(path, dtype=np.uint32), poses=np.array(poses, dtype=np.uint32), arrays=th)

    def stats(self):
        s = Stats()
        self._check(self.lib.mbgfe_get_stats(self.handle, C.byref(s)))
        return {n: getattr(s, n) for n, _ in Stats._fields_ if n != "reserved"}

    def timer_start(self):
        self._check(self.lib.mbgfe_timer_start(self.handle))

    def timer_stop(self):
        ms = C.c_float()
        self._check(self.lib.mbgfe_timer_stop(self.handle, C.byref(ms)))
        return float(ms.value)

    def pinned(self, nbytes):
        return PinnedBuffer(self.lib, nbytes)

    # ---- multi-GPU exchange (include/mbgfe.h): the whole thing runs inside the library
    def unique_id(self):
        """ncclGetUniqueId (rank 0), as 128 bytes to hand to every rank"""
        buf = (C.c_char * 128)()
        self._check(self.lib.mbgfe_comm_unique_id(buf))
        return bytes(buf)

    def comm_init_nccl(self, unique_id):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        self._check(self.lib.mbgfe_comm_init_nccl(self.handle, buf))

    def comm_init_callbacks(self, alltoallv, allgather):
        """the caller's collectives on the library's device buffers: alltoallv(send_ptr, send_bytes[world], recv_ptr, recv_bytes[world]),
        allgather(send_ptr, recv_ptr, bytes_per_rank); both return 0 on success"""
        world = self.params.world
        A2A = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p, C.POINTER(C.c_uint64))
        AG = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64)
        self._cb = (A2A(lambda ctx, s, sb, r, rb: int(alltoallv(s or 0, [sb[i] for i in range(world)], r or 0, [rb[i] for i in range(world)]) or 0)),
                    AG(lambda ctx, s, r, n: int(allgather(s or 0, r or 0, n) or 0)))
        self._check(self.lib.mbgfe_comm_init_callbacks(self.handle, self._cb[0], self._cb[1], None))

    def exchange(self):
        self._check(self.lib.mbgfe_exchange(self.handle))

    @staticmethod
    def exchange_group(handles):
        """all ranks in this process (one handle per GPU, or several on one GPU): peer copies instead of a communicator"""
        arr = (C.c_void_p * len(handles))(*[h.handle for h in handles])
        rc = handles[0].lib.mbgfe_exchange_group(arr, len(handles))
        if rc:
            for h in handles:
                h._check(rc) if h.lib.mbgfe_last_error(h.handle) else None
            handles[0]._check(rc)

    # ---- known-answer access
    def kat_tables(self):
        h = np.zeros(30948, dtype=np.uint64)
        c = np.zeros(30948, dtype=np.uint16)
        self._check(self.lib.mbgfe_kat_tables(self.handle, h.ctypes.data, c.ctypes.data))
        return h, c

    def kat_kmer_hash(self, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        out = np.zeros(2, dtype=np.uint64)
        self._check(self.lib.mbgfe_kat_kmer_hash(self.handle, codes.ctypes.data, len(codes), out.ctypes.data))
        return int(out[0]), int(out[1])

    def kat_kmer_hash_noalloc(self, codes):
        """the same through mbgfe_kat_kmer_hash_noalloc: static device scratch, no allocation; len(codes) <= NOALLOC_MAX_K"""
        codes = np.ascontiguousarray(codes, dtype=np.uint16)
        out = np.zeros(2, dtype=np.uint64)
        self._check(self.lib.mbgfe_kat_kmer_hash_noalloc(self.handle, codes.ctypes.data, len(codes), out.ctypes.data))
        return int(out[0]), int(out[1])
