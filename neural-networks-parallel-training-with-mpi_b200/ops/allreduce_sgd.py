"""Python face of the fused all-reduce + SGD kernels (``csrc/allreduce_sgd.cu``)."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import MAX_RANKS, c_float, c_int, c_ll, c_void_p

SYNC_MAX_BLOCKS = 160
MODES = {"fused_oneshot": 0, "fused_twoshot": 1, "fused_nvls": 2}
FLAG_BYTES = SYNC_MAX_BLOCKS * MAX_RANKS * 4


class SyncArgs(C.Structure):
    _fields_ = [
        ("rank", c_int), ("world", c_int),
        ("grads", c_void_p * MAX_RANKS), ("params", c_void_p * MAX_RANKS),
        ("shadow", c_void_p * MAX_RANKS), ("flags", c_void_p * MAX_RANKS),
        ("momentum", c_void_p), ("epochs", c_void_p), ("error_flag", c_void_p),
        ("weights", c_float * MAX_RANKS), ("div", c_float), ("lr", c_float), ("mu", c_float),
        ("first_step", c_int), ("grads_bf16", c_int), ("offset", c_ll), ("n", c_ll),
        ("mode", c_int), ("push_master", c_int), ("replicate_tail", c_ll), ("max_blocks", c_int), ("timeout_ms", c_int),
        ("mc_grads", c_void_p), ("mc_shadow", c_void_p), ("mc_params", c_void_p),
    ]


_lib.register("b200dp_allreduce_sgd", [C.POINTER(SyncArgs), c_void_p])
_lib.register("b200dp_allreduce_sgd_emulated", [C.POINTER(SyncArgs), c_int, c_void_p, c_void_p])
_lib.register("b200dp_peer_barrier", [C.POINTER(SyncArgs), c_void_p])
_lib.register("b200dp_broadcast", [C.POINTER(SyncArgs), c_int, c_ll, c_void_p])


def make_args(rank: int, world: int, grads: Sequence[int], params: Sequence[int],
              shadow: Optional[Sequence[int]], flags: Sequence[int], momentum: int, epochs: int,
              error_flag: int, weights: Sequence[float], div: float, grads_bf16: bool,
              mode: int, push_master: bool, max_blocks: int = 0, timeout_ms: int = 10000,
              mc_grads: Optional[int] = None, mc_shadow: Optional[int] = None,
              mc_params: Optional[int] = None) -> SyncArgs:
    a = SyncArgs()
    a.rank, a.world = rank, world
    for r in range(world):
        a.grads[r], a.params[r], a.flags[r] = grads[r], params[r], flags[r]
        a.shadow[r] = shadow[r] if shadow is not None else None
        a.weights[r] = weights[r]
    a.momentum, a.epochs, a.error_flag = momentum, epochs, error_flag
    a.div, a.grads_bf16 = div, int(grads_bf16)
    a.mode, a.push_master, a.max_blocks, a.timeout_ms = mode, int(push_master), max_blocks, timeout_ms
    a.mc_grads, a.mc_shadow, a.mc_params = mc_grads, mc_shadow, mc_params
    return a


def launch(a: SyncArgs, offset: int, n: int, lr: float, mu: float, first_step: bool, stream=None,
           replicate_tail: int = 0) -> None:
    """``replicate_tail``: the last that-many elements of the bucket (a layer's biases) get their fp32
    master value written into every peer even when the master is otherwise sharded (bf16 two-shot /
    NVLS), because kernels read them in fp32 on every rank."""
    a.offset, a.n, a.lr, a.mu, a.first_step = offset, n, lr, mu, int(first_step)
    a.replicate_tail = replicate_tail
    _lib.check(_lib.lib().b200dp_allreduce_sgd(C.byref(a), _lib.stream_ptr(stream)))
    _lib.count()


def launch_emulated(per_rank: List[SyncArgs], offset: int, n: int, lr: float, mu: float, first_step: bool,
                    device_args: torch.Tensor, stream=None, replicate_tail: int = 0) -> None:
    arr = (SyncArgs * len(per_rank))()
    for i, a in enumerate(per_rank):
        a.offset, a.n, a.lr, a.mu, a.first_step = offset, n, lr, mu, int(first_step)
        a.replicate_tail = replicate_tail
        C.memmove(C.byref(arr, i * C.sizeof(SyncArgs)), C.byref(a), C.sizeof(SyncArgs))
    _lib.check(_lib.lib().b200dp_allreduce_sgd_emulated(arr, len(per_rank), device_args.data_ptr(),
                                                        _lib.stream_ptr(stream)))
    _lib.count()


def device_args_bytes() -> int:
    lib = _lib.lib()
    lib.b200dp_sync_args_device_bytes.restype = c_ll
    return int(lib.b200dp_sync_args_device_bytes())


def peer_barrier(a: SyncArgs, stream=None) -> None:
    _lib.check(_lib.lib().b200dp_peer_barrier(C.byref(a), _lib.stream_ptr(stream)))
    _lib.count()


def broadcast(a: SyncArgs, root: int, nbytes: int, stream=None) -> None:
    _lib.check(_lib.lib().b200dp_broadcast(C.byref(a), root, nbytes, _lib.stream_ptr(stream)))
    _lib.count()
