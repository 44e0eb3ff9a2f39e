"""Python handle on a libdgs communicator (NCCL over NVLink on a side stream).

Takes the place of the c10d process group the reference creates at
harness/determined/pytorch/_pytorch_trial.py:296-300 / _trainer.py:216-236.  The 128-byte
ncclUniqueId is generated on the chief and shipped through the control plane that already exists
(DistributedContext.broadcast over ZeroMQ), or through any other callable given as `bcast`.
"""
import ctypes
from typing import Any, Callable, Optional

import torch

from determined_clone_b200 import _lib


def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class Comm:
    def __init__(self, rank: int, size: int, device: torch.device, bcast: Callable[[Any], Any],
                 allgather: Optional[Callable[[Any], Any]] = None) -> None:
        self.rank, self.size, self.device = rank, size, device
        # control-plane allgather of small Python objects (CUDA IPC handles of the peer-memory exchange, _peer.py)
        self.allgather_obj = allgather
        ident = ctypes.create_string_buffer(128)
        if rank == 0:
            _lib.calls.dgs_comm_unique_id(ident)
        raw = bcast(bytes(ident.raw) if rank == 0 else None)
        if not isinstance(raw, (bytes, bytearray)) or len(raw) != 128:
            raise _lib.DgsError("communicator bootstrap: did not receive a 128-byte NCCL unique id")
        handle = ctypes.c_void_p()
        index = device.index if device.index is not None else torch.cuda.current_device()
        _lib.calls.dgs_comm_init(ctypes.create_string_buffer(bytes(raw), 128), rank, size, index, ctypes.byref(handle))
        self._h: Optional[ctypes.c_void_p] = handle

    @classmethod
    def from_distributed(cls, distributed: Any, device: torch.device) -> "Comm":
        return cls(distributed.rank, distributed.size, device, distributed.broadcast, distributed.allgather)

    def _handle(self) -> ctypes.c_void_p:
        if self._h is None:
            raise _lib.DgsError("communicator is closed")
        return self._h

    def allreduce(self, t: torch.Tensor) -> None:
        """In-place sum over ranks, enqueued on the side stream after everything on the current stream."""
        _lib.calls.dgs_comm_allreduce_async(
            self._handle(), ctypes.c_void_p(t.data_ptr()), t.numel(), _lib.dtype_code(t.dtype), _stream()
        )

    def allreduce_raw(self, ptr: int, count: int, dtype_code: int) -> None:
        _lib.calls.dgs_comm_allreduce_async(self._handle(), ctypes.c_void_p(ptr), count, dtype_code, _stream())

    def reduce_scatter_raw(self, ptr: int, count_per_rank: int, dtype_code: int) -> None:
        """In place: this rank's slice of [ptr, ptr + size*count_per_rank) ends up holding the sum over ranks."""
        _lib.calls.dgs_comm_reduce_scatter_async(self._handle(), ctypes.c_void_p(ptr), count_per_rank, dtype_code, _stream())

    def allgather_inplace_raw(self, ptr: int, bytes_per_rank: int) -> None:
        """In place: every rank's slice of [ptr, ptr + size*bytes_per_rank) is filled with its owner's data."""
        _lib.calls.dgs_comm_allgather_async(
            self._handle(), ctypes.c_void_p(ptr + self.rank * bytes_per_rank), ctypes.c_void_p(ptr), bytes_per_rank, _stream()
        )

    def broadcast(self, t: torch.Tensor, root: int = 0) -> None:
        _lib.calls.dgs_comm_broadcast_async(
            self._handle(), ctypes.c_void_p(t.data_ptr()), t.numel() * t.element_size(), root, _stream()
        )

    def allgather(self, send: torch.Tensor, recv: torch.Tensor) -> None:
        nbytes = send.numel() * send.element_size()
        if recv.numel() * recv.element_size() != nbytes * self.size:
            raise _lib.DgsError("allgather: recv must hold size * send bytes")
        _lib.calls.dgs_comm_allgather_async(
            self._handle(), ctypes.c_void_p(send.data_ptr()), ctypes.c_void_p(recv.data_ptr()), nbytes, _stream()
        )

    def l2norm_partial(self, ptr: int, dtype_code: int, n: int, partials_ptr: int) -> None:
        """Norm pass 1 over a slice, on the side stream behind the collectives issued so far."""
        _lib.calls.dgs_comm_l2norm_partial_async(self._handle(), ctypes.c_void_p(ptr), dtype_code, n, ctypes.c_void_p(partials_ptr))

    def wait(self) -> None:
        """Make the current stream wait for every collective issued so far (no host block)."""
        _lib.calls.dgs_comm_wait(self._handle(), _stream())

    def close(self) -> None:
        if self._h is not None:
            _lib.calls.dgs_comm_destroy(self._h)
            self._h = None
