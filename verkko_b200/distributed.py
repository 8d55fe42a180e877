"""Multi-GPU sharding of the k-mer table: reads are sharded across ranks, every selected k-mer is sent
to the rank that owns its canonical hash range with one all-to-all-v (NCCL over NVLink on GPUs), and each rank counts
its partition.  torch.distributed is plumbing only; the records live in the library's device buffers.

The same code runs on the GPU-less emulator build with the gloo backend (tests), where "device" pointers are host
pointers.
"""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

RECORD_WORDS = 4  # 32-byte records: canonical hash lo, hi, flags, reserved


class _DevicePointer:
    def __init__(self, ptr, n_words):
        self.__cuda_array_interface__ = {"shape": (n_words,), "typestr": "<i8", "data": (ptr, False), "version": 3}


def wrap_pointer(ptr, n_words, on_gpu):
    """int64 tensor over n_words 8-byte words at ptr (zero-copy)"""
    if n_words == 0:
        return torch.empty(0, dtype=torch.int64, device="cuda" if on_gpu else "cpu")
    if on_gpu:
        return torch.as_tensor(_DevicePointer(ptr, n_words), device="cuda")
    buf = (C.c_int64 * n_words).from_address(ptr)
    return torch.from_numpy(np.frombuffer(buf, dtype=np.int64))


def owner_of(hash_hi, world):
    """range partition of the canonical hash space: owner = floor(hi * world / 2^64)"""
    return (np.asarray(hash_hi, dtype=np.uint64).astype(object) * world) >> 64


class DistributedCounter:
    def __init__(self, fe, world, rank, on_gpu=True):
        self.fe, self.world, self.rank, self.on_gpu = fe, world, rank, on_gpu
        self.last_exchange_ms = 0.0
        self.launches = 4  # owner count (x2), owner pack, table clear + insert per step
        self.n_recv = 0

    def exchange_and_count(self):
        fe, world = self.fe, self.world
        dev = "cuda" if self.on_gpu else "cpu"
        send_counts = fe.exchange_counts().astype(np.int64)
        sc = torch.from_numpy(send_counts).to(dev)
        rc = torch.empty_like(sc)
        dist.all_to_all_single(rc, sc)
        recv_counts = rc.cpu().numpy()
        n_send, n_recv = int(send_counts.sum()), int(recv_counts.sum())
        send_ptr = fe.exchange_pack()
        recv_ptr = fe.exchange_reserve(n_recv)
        send_t = wrap_pointer(send_ptr, n_send * RECORD_WORDS, self.on_gpu)
        recv_t = wrap_pointer(recv_ptr, n_recv * RECORD_WORDS, self.on_gpu)
        if self.on_gpu:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        dist.all_to_all_single(recv_t, send_t, output_split_sizes=(recv_counts * RECORD_WORDS).tolist(),
                               input_split_sizes=(send_counts * RECORD_WORDS).tolist())
        if self.on_gpu:
            e1.record()
            torch.cuda.synchronize()
            self.last_exchange_ms = e0.elapsed_time(e1)
        fe.exchange_insert(recv_ptr, n_recv)
        self.n_recv = n_recv
        self.sent_bytes = n_send * 32

    def download_counts(self):
        cov = self.fe.node_coverage(copy=False)
        self.fe.node_hashes(copy=False)
        return int(cov.sum(dtype=np.uint64))
