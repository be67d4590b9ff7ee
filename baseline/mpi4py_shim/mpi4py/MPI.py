"""``mpi4py.MPI`` surface used by the reference (ref.py:61-63,87,97,108,133,138,185,199,203):
``COMM_WORLD.Get_rank/Get_size/bcast/Bcast/Scatter/Scatterv/gather/send/recv``, ``MPI.INT``,
``MPI.DOUBLE``.  Lower-case methods pickle Python objects, upper-case methods move numpy buffers,
like the real library.  Importing this module joins the process group (real mpi4py calls MPI_Init
at import).  ``COMM_WORLD.gather_times`` records a perf_counter stamp at every ``gather`` - the
reference calls it exactly once per training step (ref.py:185), which lets a harness time its
steps without touching its source."""
import os
import time

import numpy as np
import torch
import torch.distributed as dist

INT = np.dtype("int32")
DOUBLE = np.dtype("float64")
FLOAT = np.dtype("float32")


def _ensure_group():
    if dist.is_initialized():
        return
    rank = int(os.environ.get("RANK", os.environ.get("OMPI_COMM_WORLD_RANK", "0")))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_SIZE", "1")))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    dist.init_process_group("gloo", rank=rank, world_size=world)


def _buf_and_dtype(spec):
    """mpi4py buffer specs: ``array`` or ``[array, datatype]`` or ``[array, counts, displs, datatype]``."""
    if isinstance(spec, (list, tuple)):
        arr = spec[0]
        dt = spec[-1] if isinstance(spec[-1], np.dtype) else None
        rest = spec[1:-1] if dt is not None else spec[1:]
        return arr, dt, rest
    return spec, None, ()


class Intracomm:
    def __init__(self):
        self.gather_times = []
        self._group = None

    def _g(self):
        if self._group is None:
            _ensure_group()
            # a private gloo group: the harness process may own an NCCL default group
            self._group = dist.new_group(backend="gloo") if dist.get_backend() != "gloo" else dist.group.WORLD
        return self._group

    def Get_rank(self):
        self._g()
        return dist.get_rank()

    def Get_size(self):
        self._g()
        return dist.get_world_size()

    # ---- pickle flavour ----
    def bcast(self, obj, root=0):
        box = [obj]
        dist.broadcast_object_list(box, src=root, group=self._g())
        return box[0]

    def gather(self, sendobj, root=0):
        self.gather_times.append(time.perf_counter())
        g = self._g()
        out = [None] * dist.get_world_size() if dist.get_rank() == root else None
        dist.gather_object(sendobj, out, dst=root, group=g)
        return out

    def send(self, obj, dest, tag=0):
        dist.send_object_list([obj], dst=dest, group=self._g())

    def recv(self, buf=None, source=0, tag=0):
        box = [None]
        dist.recv_object_list(box, src=source, group=self._g())
        return box[0]

    def barrier(self):
        dist.barrier(group=self._g())

    Barrier = barrier

    # ---- buffer flavour ----
    def Bcast(self, buf, root=0):
        arr, dt, _ = _buf_and_dtype(buf)
        if dt is not None and (arr.nbytes % dt.itemsize) != 0:
            # mpi4py refuses a buffer whose byte length is not a multiple of the datatype extent;
            # this is what the reference's int8 count array declared as MPI.INT runs into
            # (ref.py:118,128,133)
            raise ValueError("message: buffer length %d is not a multiple of datatype extent %d"
                             % (arr.nbytes, dt.itemsize))
        t = torch.from_numpy(arr.view(np.uint8).reshape(-1))
        dist.broadcast(t, src=root, group=self._g())

    def Scatter(self, sendbuf, recvbuf, root=0):
        g = self._g()
        rank, world = dist.get_rank(), dist.get_world_size()
        rarr, _, _ = _buf_and_dtype(recvbuf)
        out = torch.from_numpy(rarr.reshape(-1))
        if rank == root:
            sarr, _, _ = _buf_and_dtype(sendbuf)
            flat = torch.from_numpy(np.ascontiguousarray(sarr).reshape(-1))
            chunks = [c.contiguous() for c in flat.chunk(world)]
            dist.scatter(out, chunks, src=root, group=g)
        else:
            dist.scatter(out, None, src=root, group=g)

    def Scatterv(self, sendbuf, recvbuf, root=0):
        g = self._g()
        rank, world = dist.get_rank(), dist.get_world_size()
        rarr, _, _ = _buf_and_dtype(recvbuf)
        out = torch.from_numpy(rarr.reshape(-1))
        if rank == root:
            sarr, _, rest = _buf_and_dtype(sendbuf)
            counts, displs = [np.asarray(x).astype(np.int64) for x in rest[:2]]
            flat = torch.from_numpy(np.ascontiguousarray(sarr).reshape(-1))
            for r in range(world):
                piece = flat[displs[r]:displs[r] + counts[r]].contiguous()
                if r == root:
                    out.copy_(piece)
                else:
                    dist.send(piece, dst=r, group=g)
        else:
            dist.recv(out, src=root, group=g)


COMM_WORLD = Intracomm()


def Finalize():
    if dist.is_initialized():
        dist.destroy_process_group()
