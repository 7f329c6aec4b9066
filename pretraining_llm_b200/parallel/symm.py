"""Symmetric device memory across the ranks of one node.

Every rank allocates the same-sized buffer (``b200_ipc_alloc`` = cudaMalloc), exports it as a
64-byte CUDA IPC handle, and imports everybody else's handle: afterwards each rank holds a
device pointer to every peer's buffer, and kernels read / write peer memory directly over
NVLink 5 / NVSwitch.  ``torch.distributed`` (NCCL or gloo) only carries the handles.

This is the role c10d/NCCL's internal buffers play for the reference's DDP; here the
buffers are the framework's own gradient / weight / signal arrays, addressed by our kernels.
"""
from __future__ import annotations

import ctypes
from typing import List

import torch
import torch.distributed as dist

from ..ops import lib


class _CudaArray:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class SymmBuffer:
    """``nbytes`` of device memory on every rank, peer-mapped everywhere."""

    def __init__(self, nbytes: int, rank: int, world: int, device: torch.device):
        self.nbytes, self.rank, self.world = nbytes, rank, world
        L = lib.load()
        p = ctypes.c_void_p()
        lib.check(L.b200_ipc_alloc(ctypes.byref(p), nbytes), "ipc_alloc")
        self.local_ptr = p.value
        handle = ctypes.create_string_buffer(64)
        lib.check(L.b200_ipc_export(p, handle), "ipc_export")
        handles: List[bytes] = [b""] * world
        dist.all_gather_object(handles, handle.raw)
        self.ptrs: List[int] = []
        self._imported: List[int] = []
        for r in range(world):
            if r == rank:
                self.ptrs.append(self.local_ptr)
                continue
            q = ctypes.c_void_p()
            buf = ctypes.create_string_buffer(handles[r], 64)
            lib.check(L.b200_ipc_import(buf, ctypes.byref(q)), "ipc_import")
            self.ptrs.append(q.value)
            self._imported.append(q.value)
        self._arr = _CudaArray(self.local_ptr, nbytes)
        self.bytes_tensor = torch.as_tensor(self._arr, device=device)   # uint8 view of the local buffer
        self.ptr_array = (ctypes.c_void_p * world)(*self.ptrs)

    def view(self, dtype: torch.dtype) -> torch.Tensor:
        return self.bytes_tensor.view(dtype)

    def close(self) -> None:
        L = lib.load()
        for q in self._imported:
            L.b200_ipc_close(ctypes.c_void_p(q))
        self._imported = []
        if self.local_ptr:
            self.bytes_tensor = None
            L.b200_ipc_free(ctypes.c_void_p(self.local_ptr))
            self.local_ptr = 0
