"""Test-side helpers for the sharded table: the owner rule restated in numpy, and gloo collectives as transport callbacks
(torch.distributed is test plumbing; nothing under verkko_b200/ imports torch)."""
import ctypes as C

import numpy as np


def owner_of(hash_hi, world):
    """owner rank of canonical hashes (owner_of in csrc/kernels.cuh): equal shares of the distribution of the canonical high
    word, the smaller of two uniform words: F(x) = x + x (1 - x) in 64-bit fixed point, owner = floor(world F)"""
    mask = (1 << 64) - 1
    out = []
    for h in np.asarray(hash_hi, dtype=np.uint64).tolist():
        f = h + ((h * (mask - h)) >> 64)
        out.append((f * world) >> 64)
    return np.array(out, dtype=np.int64)


def host_view(ptr, nbytes):
    """uint8 torch tensor over host memory at ptr (the emulator build's "device" buffers are host memory)"""
    import torch
    if nbytes == 0:
        return torch.empty(0, dtype=torch.uint8)
    return torch.from_numpy(np.frombuffer((C.c_uint8 * nbytes).from_address(ptr), dtype=np.uint8))


def gloo_callbacks(world):
    """(alltoallv, allgather) for MbgFrontEnd.comm_init_callbacks over torch.distributed / gloo on host buffers"""
    import torch
    import torch.distributed as dist

    def alltoallv(send_ptr, send_bytes, recv_ptr, recv_bytes):
        s, r = host_view(send_ptr, sum(send_bytes)).clone(), host_view(recv_ptr, sum(recv_bytes))
        dist.all_to_all_single(r, s, output_split_sizes=list(recv_bytes), input_split_sizes=list(send_bytes))
        return 0

    def allgather(send_ptr, recv_ptr, nbytes):
        s, r = host_view(send_ptr, nbytes), host_view(recv_ptr, nbytes * world)
        dist.all_gather(list(r.split(nbytes)), s.clone())
        return 0
    return alltoallv, allgather
