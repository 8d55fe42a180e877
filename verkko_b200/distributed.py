"""Multi-GPU sharding of the k-mer table: reads are sharded across ranks, every selected k-mer is sent
to the rank that owns its canonical hash range with one all-to-all-v (NCCL over NVLink on GPUs), each rank counts its
partition, answers with global node ids, and the edge observations follow to the owner of their `from` node.
torch.distributed is plumbing only; records are produced and consumed by the library's kernels in its device buffers.

The same code runs on the GPU-less emulator build with the gloo backend (tests), where "device" pointers are host
pointers.
"""
import ctypes as C
import time

import numpy as np
import torch
import torch.distributed as dist

RECORD_WORDS = 4       # node records: 32 bytes = canonical hash lo, hi, flags, reserved (int64 words)
EDGE_RECORD_WORDS = 2  # edge records: 16 bytes = canonical (from, to) key, (instance << 24) | overlap

_TYPESTR = {torch.int64: ("<i8", np.int64, C.c_int64), torch.int32: ("<i4", np.int32, C.c_int32)}


class _DevicePointer:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


def wrap_pointer(ptr, n, on_gpu, dtype=torch.int64):
    """tensor over n elements at ptr (zero-copy)"""
    typestr, npdt, cdt = _TYPESTR[dtype]
    if n == 0:
        return torch.empty(0, dtype=dtype, device="cuda" if on_gpu else "cpu")
    if on_gpu:
        return torch.as_tensor(_DevicePointer(ptr, n, typestr), device="cuda")
    buf = (cdt * n).from_address(ptr)
    return torch.from_numpy(np.frombuffer(buf, dtype=npdt))


def owner_of(hash_hi, world):
    """owner rank of canonical hashes (k_owner_* in csrc/kernels.cuh): equal shares of the distribution of the canonical
    high word, which is the smaller of two uniform words: F(x) = x + x (1 - x) in 64-bit fixed point, owner = floor(world F)"""
    mask = (1 << 64) - 1
    out = []
    for h in np.asarray(hash_hi, dtype=np.uint64).tolist():
        f = h + ((h * (mask - h)) >> 64)
        out.append((f * world) >> 64)
    return np.array(out, dtype=np.int64)


class DistributedCounter:
    """Drives the exchange steps of include/mbgfe.h for one rank."""

    def __init__(self, fe, world, rank, on_gpu=True, edges=True):
        self.fe, self.world, self.rank, self.on_gpu, self.edges = fe, world, rank, on_gpu, edges
        self.last_exchange_ms = 0.0
        # owner count (x2) + pack, table clear + insert;  with edges: 7 export-sort launches for the ranks, id lookup,
        # edge owner count + pack, edge clear + insert
        self.launches = 5 + (12 if edges else 0)
        self.n_recv = 0
        self.sent_bytes = 0
        self.id_begin = None
        self.wall_ms = {}        # host wall clock per phase of the last exchange (where the step's host gaps are)

    def _mark(self, name):
        now = time.perf_counter()
        self.wall_ms[name] = self.wall_ms.get(name, 0.0) + (now - self._t) * 1e3
        self._t = now

    def _dev(self):
        return "cuda" if self.on_gpu else "cpu"

    def _counts_exchange(self, send_counts):
        sc = torch.from_numpy(send_counts.astype(np.int64)).to(self._dev())
        rc = torch.empty_like(sc)
        dist.all_to_all_single(rc, sc)
        return rc.cpu().numpy()

    def _timed(self, fn):
        if not self.on_gpu:
            fn()
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        self.last_exchange_ms += e0.elapsed_time(e1)

    def exchange_and_count(self):
        fe, world = self.fe, self.world
        self.last_exchange_ms = 0.0
        self.wall_ms = {}
        self._t = time.perf_counter()
        # ---- nodes: k-mer records to their owners
        send_counts = fe.exchange_counts().astype(np.int64)
        self._mark("node_counts")
        recv_counts = self._counts_exchange(send_counts)
        self._mark("node_counts_a2a")
        n_send, n_recv = int(send_counts.sum()), int(recv_counts.sum())
        send_ptr = fe.exchange_pack()
        self._mark("node_pack")
        recv_ptr = fe.exchange_reserve(n_recv)
        send_t = wrap_pointer(send_ptr, n_send * RECORD_WORDS, self.on_gpu)
        recv_t = wrap_pointer(recv_ptr, n_recv * RECORD_WORDS, self.on_gpu)
        self._timed(lambda: dist.all_to_all_single(recv_t, send_t, output_split_sizes=(recv_counts * RECORD_WORDS).tolist(),
                                                   input_split_sizes=(send_counts * RECORD_WORDS).tolist()))
        self._mark("node_a2a")
        fe.exchange_insert(recv_ptr, n_recv)
        self._mark("node_insert")
        self.n_recv = n_recv
        self.sent_bytes = n_send * 32
        if not self.edges:
            return
        # ---- ids: partition sizes -> id ranges; owners answer every record with its global node id
        n_local = fe.stats()["distinct"]
        mine = torch.tensor([n_local, fe.count_selected()], dtype=torch.int64, device=self._dev())
        both = [torch.zeros(2, dtype=torch.int64, device=self._dev()) for _ in range(world)]
        dist.all_gather(both, mine)
        sizes = np.array([int(b[0]) for b in both], dtype=np.uint64)
        n_inst = np.array([int(b[1]) for b in both], dtype=np.uint64)
        id_begin = np.zeros(world + 1, dtype=np.uint64)
        id_begin[1:] = np.cumsum(sizes)
        self.id_begin = id_begin
        inst_base = int(n_inst[:self.rank].sum())
        self._mark("sizes_allgather")
        ids_ptr = fe.exchange_node_ids(int(id_begin[self.rank]), recv_ptr, n_recv)
        self._mark("rank_and_ids")
        back_ptr = fe.exchange_scratch(0, n_send * 4)
        ids_t = wrap_pointer(ids_ptr, n_recv, self.on_gpu, torch.int32)
        back_t = wrap_pointer(back_ptr, n_send, self.on_gpu, torch.int32)
        self._timed(lambda: dist.all_to_all_single(back_t, ids_t, output_split_sizes=send_counts.tolist(), input_split_sizes=recv_counts.tolist()))
        self._mark("ids_a2a")
        # ---- edges: observations to the owner of their `from` node
        e_send = fe.exchange_edge_counts(back_ptr, id_begin, inst_base).astype(np.int64)
        self._mark("edge_counts")
        e_recv = self._counts_exchange(e_send)
        ne_send, ne_recv = int(e_send.sum()), int(e_recv.sum())
        self._mark("edge_counts_a2a")
        es_ptr = fe.exchange_edge_pack(e_send)
        self._mark("edge_pack")
        er_ptr = fe.exchange_scratch(1, ne_recv * 16)
        es_t = wrap_pointer(es_ptr, ne_send * EDGE_RECORD_WORDS, self.on_gpu)
        er_t = wrap_pointer(er_ptr, ne_recv * EDGE_RECORD_WORDS, self.on_gpu)
        self._timed(lambda: dist.all_to_all_single(er_t, es_t, output_split_sizes=(e_recv * EDGE_RECORD_WORDS).tolist(),
                                                   input_split_sizes=(e_send * EDGE_RECORD_WORDS).tolist()))
        self._mark("edge_a2a")
        fe.exchange_insert_edges(er_ptr, ne_recv)
        self._mark("edge_insert")
        self.sent_bytes += n_recv * 4 + ne_send * 16

    def download_counts(self):
        cov = self.fe.node_coverage(copy=False)
        self.fe.node_hashes(copy=False)
        total = int(cov.sum(dtype=np.uint64))
        if self.edges:
            total += len(self.fe.edges(copy=False))
        return total
