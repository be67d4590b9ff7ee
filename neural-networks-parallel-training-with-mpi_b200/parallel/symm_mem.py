"""Symmetric (peer-mapped) device memory: one equally sized buffer per rank, every buffer mapped
into every process, optionally also bound to an NVLS multicast address.

This is the transport that stands where the reference has ``MPI.COMM_WORLD`` (ref.py:61-63): after
the one-time handle exchange (over the host communicator) the kernels talk to each other with plain
loads / stores / ``multimem`` instructions on these pointers.  Native side: ``csrc/symm_mem.cu``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops._lib import c_int, c_ll, c_void_p

_lib.register("b200dp_ipc_alloc", [c_ll, C.POINTER(c_void_p), C.c_char_p])
_lib.register("b200dp_ipc_open", [C.c_char_p, C.POINTER(c_void_p)])
_lib.register("b200dp_ipc_close", [c_void_p])
_lib.register("b200dp_ipc_free", [c_void_p])
_lib.register("b200dp_vmm_supported", [C.POINTER(c_int), C.POINTER(c_int)])
_lib.register("b200dp_vmm_alloc", [c_ll, C.POINTER(c_void_p), C.POINTER(c_int), C.POINTER(c_ll)])
_lib.register("b200dp_vmm_import", [c_int, c_int, c_ll, C.POINTER(c_void_p)])
_lib.register("b200dp_vmm_free", [c_void_p, c_ll])
_lib.register("b200dp_mc_create", [c_int, c_ll, C.POINTER(c_int), C.POINTER(C.c_ulonglong)])
_lib.register("b200dp_mc_import", [c_int, c_int, C.POINTER(C.c_ulonglong)])
_lib.register("b200dp_mc_add_device", [C.c_ulonglong])
_lib.register("b200dp_mc_bind", [C.c_ulonglong, c_void_p, c_ll, C.POINTER(c_void_p)])
_lib.register("b200dp_mc_release", [C.c_ulonglong, c_void_p, c_ll])


class _RawCudaMemory:
    """Exposes a raw device pointer through ``__cuda_array_interface__`` so torch can view it."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2}
        self._owner = owner


def vmm_support():
    fd, mc = c_int(0), c_int(0)
    _lib.check(_lib.lib().b200dp_vmm_supported(C.byref(fd), C.byref(mc)))
    return bool(fd.value), bool(mc.value)


class SymmetricBuffer:
    """``nbytes`` on every rank.  ``ptrs[r]`` is rank r's buffer as mapped into this process;
    ``mc_ptr`` (if multicast was requested and is available) addresses all copies at once."""

    def __init__(self, comm, nbytes: int, device, flavour: str = "auto", multicast: bool = False):
        self.comm, self.nbytes, self.device = comm, int(nbytes), torch.device(device)
        self.rank, self.world = comm.rank, comm.size
        self.mc_ptr: Optional[int] = None
        self._mc_handle = None
        self._closed = False
        lib = _lib.lib()
        if flavour == "auto":
            flavour = "ipc"
            if multicast:
                try:
                    fd_ok, mc_ok = vmm_support()
                except _lib.B200dpError:
                    fd_ok = mc_ok = False
                # every rank must take the same branch
                oks = comm.allgather(bool(fd_ok and mc_ok))
                flavour = "vmm" if all(oks) else "ipc"
        self.flavour = flavour
        p = c_void_p()
        if flavour == "ipc":
            handle = C.create_string_buffer(64)
            _lib.check(lib.b200dp_ipc_alloc(self.nbytes, C.byref(p), handle))
            self.local_ptr = p.value
            self.padded = self.nbytes
            handles = comm.allgather(handle.raw)
            self.ptrs: List[int] = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    self.ptrs.append(self.local_ptr)
                else:
                    q = c_void_p()
                    _lib.check(lib.b200dp_ipc_open(h, C.byref(q)))
                    self.ptrs.append(q.value)
        elif flavour == "vmm":
            fd, padded = c_int(-1), c_ll(0)
            _lib.check(lib.b200dp_vmm_alloc(self.nbytes, C.byref(p), C.byref(fd), C.byref(padded)))
            self.local_ptr, self.padded, self._fd = p.value, padded.value, fd.value
            infos = comm.allgather((os.getpid(), fd.value, padded.value))
            self.ptrs = []
            for r, (pid, pfd, pbytes) in enumerate(infos):
                if r == self.rank:
                    self.ptrs.append(self.local_ptr)
                else:
                    q = c_void_p()
                    _lib.check(lib.b200dp_vmm_import(pid, pfd, pbytes, C.byref(q)))
                    self.ptrs.append(q.value)
            comm.barrier()             # nobody closes an exported fd before every peer duplicated it
            if multicast and self.world > 1:
                self._bind_multicast()
        else:
            raise ValueError(f"unknown symmetric-memory flavour {flavour!r}")
        self._mem = _RawCudaMemory(self.local_ptr, self.nbytes, self)

    def _bind_multicast(self):
        lib, comm = _lib.lib(), self.comm
        h = C.c_ulonglong(0)
        ok = True
        try:
            if self.rank == 0:
                fd = c_int(-1)
                _lib.check(lib.b200dp_mc_create(self.world, self.padded, C.byref(fd), C.byref(h)))
                info = (os.getpid(), fd.value)
            else:
                info = None
        except _lib.B200dpError:
            info, ok = None, False
        info = comm.bcast(info, root=0)
        if info is None:
            return
        try:
            if self.rank != 0:
                _lib.check(lib.b200dp_mc_import(info[0], info[1], C.byref(h)))
            _lib.check(lib.b200dp_mc_add_device(h.value))
        except _lib.B200dpError:
            ok = False
        if not all(comm.allgather(ok)):
            return
        q = c_void_p()
        try:
            _lib.check(lib.b200dp_mc_bind(h.value, self.local_ptr, self.padded, C.byref(q)))
        except _lib.B200dpError:
            ok = False
        if all(comm.allgather(ok)):
            self.mc_ptr, self._mc_handle = q.value, h.value

    def tensor(self, dtype, numel: Optional[int] = None, byte_offset: int = 0) -> torch.Tensor:
        """A torch view over this rank's own buffer."""
        raw = torch.as_tensor(self._mem, device=self.device)
        item = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - byte_offset) // item
        return raw[byte_offset:byte_offset + numel * item].view(dtype)

    def close(self):
        if self._closed:
            return
        self._closed = True
        lib = _lib.lib()
        torch.cuda.synchronize()
        self.comm.barrier()
        if self._mc_handle is not None:
            lib.b200dp_mc_release(self._mc_handle, self.mc_ptr, self.padded)
        for r, q in enumerate(self.ptrs):
            if r == self.rank:
                continue
            if self.flavour == "ipc":
                lib.b200dp_ipc_close(q)
            else:
                lib.b200dp_vmm_free(q, self.padded)
        self.comm.barrier()
        if self.flavour == "ipc":
            lib.b200dp_ipc_free(self.local_ptr)
        else:
            lib.b200dp_vmm_free(self.local_ptr, self.padded)
            try:
                os.close(self._fd)
            except OSError:
                pass


class LocalSymmetricWorld:
    """All ranks' buffers on ONE GPU in ONE process - the memory side of the single-GPU emulation
    (``b200dp_allreduce_sgd_emulated``) used by the tests."""

    def __init__(self, world: int, nbytes: int, device):
        self.world, self.nbytes = world, nbytes
        self.buffers = [torch.zeros(nbytes, dtype=torch.uint8, device=device) for _ in range(world)]
        self.ptrs = [b.data_ptr() for b in self.buffers]

    def tensor(self, rank: int, dtype, numel=None, byte_offset=0):
        item = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - byte_offset) // item
        return self.buffers[rank][byte_offset:byte_offset + numel * item].view(dtype)
