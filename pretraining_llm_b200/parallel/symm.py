"""Symmetric device memory across the ranks of one node, with NVSwitch multicast where the box has it.

Every rank allocates the same-sized buffer and ends up with a device pointer to every peer's copy, so kernels read
and write peer memory directly over NVLink 5 / NVSwitch.  Two substrates (csrc/comm):

  * **VMM heap** (default, ``vmm.cu``): ``cuMemCreate`` physical memory exported as a POSIX file descriptor, the
    descriptors passed between the rank processes over unix sockets (``SCM_RIGHTS``), imported and mapped with
    ``cuMemMap``.  Because the memory is a VMM allocation it can also be bound into a **multicast object**
    (``cuMulticastCreate`` / ``cuMulticastBindMem``): ``mc_ptr`` is then one address whose ``multimem.st`` stores the
    switch replicates into every rank's copy and whose ``multimem.ld_reduce`` loads return the sum over all ranks,
    reduced inside the switch (NVLS).  ``mc_ptr`` is 0 when the device / driver / fabric has no multicast; the
    kernels then use the per-peer pointers.
  * **CUDA IPC** (fallback, ``dp.cu``): ``cudaMalloc`` + 64-byte ``cudaIpcMemHandle``.  Used when the VMM path is
    unavailable or ``B200_SYMM=ipc``.

Peer and multicast mappings are addressed with generic loads / stores and ``multimem`` instructions only; every TMA
tensor map of this library is built over the GPU's own memory.

``torch.distributed`` (NCCL or gloo) only carries handles, a rendezvous token and barriers.

This is the role c10d/NCCL's internal buffers play for the reference's DDP; here the buffers are the
framework's own gradient / weight / signal arrays, addressed by our kernels.
"""
from __future__ import annotations

import array
import ctypes
import os
import socket
import struct
import sys
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import lib


class _CudaArray:
    """Expose a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


# ------------------------------------------------------------------------------------------------ fd exchange
class FdExchange:
    """All-to-all exchange of file descriptors between the rank processes of one node.

    Each rank binds a datagram socket in the abstract unix namespace (no file to clean up) under a name built from a
    job token that rank 0 draws and broadcasts; a descriptor travels as SCM_RIGHTS ancillary data with the sender's
    rank and a sequence tag in the payload.  One instance per process, reused by every SymmBuffer."""

    _inst: Optional["FdExchange"] = None

    def __init__(self, rank: int, world: int):
        self.rank, self.world = rank, world
        tok = [os.urandom(8).hex() if rank == 0 else None]
        dist.broadcast_object_list(tok, src=0)
        self.token = tok[0]
        self.sock = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
        self.sock.bind(self._addr(rank))
        self.sock.settimeout(120.0)
        self.seq = 0
        self._stash = {}
        dist.barrier()                     # every socket is bound before anyone sends

    @classmethod
    def get(cls, rank: int, world: int) -> "FdExchange":
        if cls._inst is None or cls._inst.world != world or cls._inst.rank != rank:
            cls._inst = cls(rank, world)
        return cls._inst

    def _addr(self, r: int) -> bytes:
        return b"\0b200llm-" + self.token.encode() + b"-" + str(r).encode()

    def _send(self, dst: int, fd: int, tag: int) -> None:
        payload = struct.pack("ii", self.rank, tag)
        self.sock.sendmsg([payload], [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))], 0, self._addr(dst))

    def _recv(self, src: int, tag: int) -> int:
        while (src, tag) not in self._stash:
            msg, anc, _, _ = self.sock.recvmsg(8, socket.CMSG_SPACE(4))
            r, t = struct.unpack("ii", msg)
            fds = array.array("i")
            for level, typ, data in anc:
                if level == socket.SOL_SOCKET and typ == socket.SCM_RIGHTS:
                    fds.frombytes(data[:4])
            if not fds:
                raise RuntimeError("fd exchange: message without a descriptor")
            self._stash[(r, t)] = fds[0]
        return self._stash.pop((src, tag))

    def all_to_all(self, fd: int) -> List[Optional[int]]:
        """Every rank contributes one descriptor; returns the peers' descriptors (None at the own index)."""
        self.seq += 1
        for r in range(self.world):
            if r != self.rank:
                self._send(r, fd, self.seq)
        return [None if r == self.rank else self._recv(r, self.seq) for r in range(self.world)]

    def from_root(self, fd: Optional[int]) -> Optional[int]:
        """Rank 0 hands one descriptor to every other rank (returns it there; rank 0 gets None)."""
        self.seq += 1
        if self.rank == 0:
            for r in range(1, self.world):
                self._send(r, fd, self.seq)
            return None
        return self._recv(0, self.seq)


# ------------------------------------------------------------------------------------------------ capability probe
_CAPS = None


def _agree(flag: bool) -> bool:
    """True only if `flag` is true on every rank."""
    t = torch.tensor([int(flag)], device="cuda", dtype=torch.int32)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def vmm_caps(world: int) -> dict:
    global _CAPS
    if _CAPS is None or _CAPS.get("world") != world:
        out = (ctypes.c_longlong * 5)()
        ok = lib.load().b200_vmm_query(out, world) == 0
        caps = {"world": world, "vmm": bool(ok and out[0] and out[1]), "multicast": bool(ok and out[2] and out[4]),
                "gran": int(out[3]) or (2 << 20), "mc_gran": int(out[4])}
        if os.environ.get("B200_SYMM", "vmm") == "ipc":
            caps["vmm"] = False
        if os.environ.get("B200_MULTICAST", "1") != "1":
            caps["multicast"] = False
        # every rank must take the same path: agree on the minimum
        caps["vmm"] = _agree(caps["vmm"])
        caps["multicast"] = _agree(caps["multicast"]) and caps["vmm"]
        _CAPS = caps
    return _CAPS


class SymmBuffer:
    """``nbytes`` of zero-initialised device memory on every rank, peer-mapped everywhere.

    ``ptrs[r]`` / ``ptr_array``: device pointer to rank r's copy (own rank: the local mapping).
    ``mc_ptr``: multicast address of the whole buffer, or 0.  ``multicast=False`` skips the multicast object for
    buffers no kernel addresses that way (signal pads)."""

    def __init__(self, nbytes: int, rank: int, world: int, device: torch.device, multicast: bool = True):
        self.nbytes, self.rank, self.world = nbytes, rank, world
        self.mc_ptr = 0
        self.kind = "ipc"
        self.size = nbytes
        self._handles: List[int] = []
        self._maps: List[int] = []
        self._imported: List[int] = []
        caps = vmm_caps(world) if world > 1 else {"vmm": False, "multicast": False}
        done = False
        if caps["vmm"]:
            try:
                self._init_vmm(nbytes, rank, world, caps, multicast and caps["multicast"])
                done = True
            except Exception as e:       # noqa: BLE001 - any failure of the new path falls back, loudly, on every rank
                print(f"[symm] VMM heap unavailable on rank {rank} ({e}); using CUDA IPC", file=sys.stderr)
            all_ok = _agree(done)
            if done and not all_ok:
                self._close_vmm()
                done = False
            if not done:
                caps["vmm"] = caps["multicast"] = False      # do not try again for the next buffer
        if not done:
            self._init_ipc(nbytes, rank, world)
        self._arr = _CudaArray(self.local_ptr, nbytes)
        self.bytes_tensor = torch.as_tensor(self._arr, device=device)   # uint8 view of the local buffer
        self.ptr_array = (ctypes.c_void_p * world)(*self.ptrs)

    # -------------------------------------------------------------------------------------------- VMM + multicast
    def _init_vmm(self, nbytes: int, rank: int, world: int, caps: dict, want_mc: bool) -> None:
        L = lib.load()
        gran = max(caps["gran"], caps["mc_gran"] if want_mc else 0, 1)
        size = (max(nbytes, 1) + gran - 1) // gran * gran
        self.size = size
        h = ctypes.c_ulonglong()
        fd = ctypes.c_int(-1)
        lib.check(L.b200_vmm_create(size, ctypes.byref(h), ctypes.byref(fd)), "vmm_create")
        self._handles.append(h.value)
        p = ctypes.c_void_p()
        lib.check(L.b200_vmm_map(h, size, ctypes.byref(p)), "vmm_map(local)")
        self._maps.append(p.value)
        self.local_ptr = p.value
        ex = FdExchange.get(rank, world)
        fds = ex.all_to_all(fd.value)
        self.ptrs: List[int] = []
        for r in range(world):
            if r == rank:
                self.ptrs.append(self.local_ptr)
                continue
            hr = ctypes.c_ulonglong()
            lib.check(L.b200_vmm_import(fds[r], ctypes.byref(hr)), "vmm_import")
            L.b200_vmm_close_fd(fds[r])
            self._handles.append(hr.value)
            q = ctypes.c_void_p()
            lib.check(L.b200_vmm_map(hr, size, ctypes.byref(q)), "vmm_map(peer)")
            self._maps.append(q.value)
            self.ptrs.append(q.value)
        dist.barrier()                     # every peer has imported my descriptor
        L.b200_vmm_close_fd(fd.value)
        torch.as_tensor(_CudaArray(self.local_ptr, size), device="cuda").zero_()    # cuMemCreate does not clear memory
        torch.cuda.synchronize()
        self.kind = "vmm"
        if want_mc:
            try:
                self._bind_multicast(L, ex, h.value, size, rank, world)
            except Exception as e:       # noqa: BLE001
                print(f"[symm] multicast unavailable on rank {rank} ({e}); P2P only", file=sys.stderr)
                self.mc_ptr = 0
            if not _agree(self.mc_ptr != 0):
                self.mc_ptr = 0
                caps["multicast"] = False
        torch.cuda.synchronize()
        dist.barrier()

    def _bind_multicast(self, L, ex: FdExchange, mem_handle: int, size: int, rank: int, world: int) -> None:
        mc = ctypes.c_ulonglong()
        mfd = ctypes.c_int(-1)
        rc = L.b200_mc_create(size, world, ctypes.byref(mc), ctypes.byref(mfd)) if rank == 0 else 0
        created = torch.tensor([int(rc == 0)], device="cuda", dtype=torch.int32)
        dist.broadcast(created, src=0)     # nobody waits for a descriptor that will never come
        if not created.item():
            raise RuntimeError("cuMulticastCreate failed on rank 0: " + (L.b200_last_error() or b"").decode())
        if rank == 0:
            ex.from_root(mfd.value)
        else:
            got = ex.from_root(None)
            lib.check(L.b200_vmm_import(got, ctypes.byref(mc)), "mc_import")
            L.b200_vmm_close_fd(got)
        self._handles.append(mc.value)
        rc_add = L.b200_mc_add_device(mc)
        # ALL devices are added before anyone binds - and nobody binds if somebody could not add
        if not _agree(rc_add == 0):
            raise RuntimeError("cuMulticastAddDevice failed on some rank")
        if rank == 0:
            L.b200_vmm_close_fd(mfd.value)
        lib.check(L.b200_mc_bind(mc, ctypes.c_ulonglong(mem_handle), size), "mc_bind")
        q = ctypes.c_void_p()
        lib.check(L.b200_vmm_map(mc, size, ctypes.byref(q)), "vmm_map(multicast)")
        self._maps.append(q.value)
        self.mc_ptr = q.value

    def _close_vmm(self) -> None:
        L = lib.load()
        for q in self._maps:
            L.b200_vmm_unmap(ctypes.c_void_p(q), self.size)
        for h in self._handles:
            L.b200_vmm_release(ctypes.c_ulonglong(h))
        self._maps, self._handles = [], []
        self.mc_ptr = 0

    # -------------------------------------------------------------------------------------------- CUDA IPC
    def _init_ipc(self, nbytes: int, rank: int, world: int) -> None:
        L = lib.load()
        p = ctypes.c_void_p()
        lib.check(L.b200_ipc_alloc(ctypes.byref(p), max(nbytes, 16)), "ipc_alloc")
        self.local_ptr = p.value
        handle = ctypes.create_string_buffer(64)
        lib.check(L.b200_ipc_export(p, handle), "ipc_export")
        handles: List[bytes] = [b""] * world
        if world > 1:
            dist.all_gather_object(handles, handle.raw)
        self.ptrs = []
        for r in range(world):
            if r == rank:
                self.ptrs.append(self.local_ptr)
                continue
            q = ctypes.c_void_p()
            buf = ctypes.create_string_buffer(handles[r], 64)
            lib.check(L.b200_ipc_import(buf, ctypes.byref(q)), "ipc_import")
            self.ptrs.append(q.value)
            self._imported.append(q.value)
        self.kind = "ipc"

    def view(self, dtype: torch.dtype) -> torch.Tensor:
        return self.bytes_tensor.view(dtype)

    def peer_view(self, r: int, dtype: torch.dtype, device) -> torch.Tensor:
        """torch view of rank r's copy through the peer mapping (off the hot path: checkpoints, evaluation)."""
        return torch.as_tensor(_CudaArray(self.ptrs[r], self.nbytes), device=device).view(dtype)

    def close(self) -> None:
        self.bytes_tensor = None
        if self.kind == "vmm":
            self._close_vmm()
            self.local_ptr = 0
            return
        L = lib.load()
        for q in self._imported:
            L.b200_ipc_close(ctypes.c_void_p(q))
        self._imported = []
        if self.local_ptr:
            L.b200_ipc_free(ctypes.c_void_p(self.local_ptr))
            self.local_ptr = 0
