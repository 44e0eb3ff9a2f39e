"""Peer-mapped buffers of one GradSyncEngine: the host side of csrc/dgs_peer.cu.

Every rank exports its wire and parameter buffers (and a 512-byte sync block) with CUDA IPC, the 64-byte handles
travel over the control plane that already exists (DistributedContext.allgather, the ZeroMQ star of the reference's
harness/determined/core/_distributed.py:337-356), and every rank maps the others' allocations.  From then on the
bucket exchange of the data-parallel step (reference harness/determined/pytorch/_pytorch_context.py:295-296,
:813-816: torch DDP's allreduce) is stores and loads on those pointers from this library's own kernels.

Set-up is all-or-none: a rank whose export or open fails (no peer access between the GPUs, an allocator whose memory
cannot be shared) reports it, and the engine stays on the NCCL routes everywhere."""
import ctypes
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

from determined_clone_b200 import _lib

S_PUSHED, S_NORM, S_PARAMS = 0, 1, 2  # sync slots: gradient pushes landed / norm exchange / parameter stores landed


class PeerGroup:
    def __init__(self, rank: int, size: int, device: torch.device, allgather: Callable[[Any], List[Any]],
                 buffers: Dict[str, torch.Tensor], timeout_s: float = 0.0) -> None:
        if size > _lib.PEER_MAX:
            raise _lib.DgsError(f"the peer exchange handles at most {_lib.PEER_MAX} ranks, got {size}")
        self.rank, self.size, self.device, self.timeout_s = rank, size, device, float(timeout_s)
        self.sync = torch.zeros(_lib.PEER_SYNC_BYTES, dtype=torch.uint8, device=device)
        self.buffers = dict(buffers, __sync__=self.sync)  # keeps the exported tensors alive
        self._epoch = [0] * _lib.PEER_SLOTS
        mine: Dict[str, Tuple[bytes, int]] = {}
        err: Optional[str] = None
        try:
            if device.type == "cuda":
                torch.cuda.current_stream().synchronize()  # the zeroed sync block must be in memory before anyone maps it
            for name, t in self.buffers.items():
                handle = ctypes.create_string_buffer(_lib.PEER_HANDLE_BYTES)
                off = ctypes.c_int64(0)
                _lib.calls.dgs_peer_export(ctypes.c_void_p(t.data_ptr()), handle, ctypes.byref(off))
                mine[name] = (bytes(handle.raw), int(off.value))
        except Exception as e:  # noqa: BLE001 -- reported to every rank below
            err = f"rank {rank}: {e}"
        everyone = allgather((err, mine))
        self._allgather = allgather
        self._opened: List[bytes] = []  # one entry per successful dgs_peer_open, for close()
        self.base: Dict[str, List[int]] = {}
        if not any(e for e, _ in everyone):
            try:
                for name, t in self.buffers.items():
                    ptrs = []
                    for r, (_, exported) in enumerate(everyone):
                        if r == rank:
                            ptrs.append(t.data_ptr())
                            continue
                        handle, off = exported[name]
                        out = ctypes.c_void_p()
                        _lib.calls.dgs_peer_open(ctypes.create_string_buffer(handle, _lib.PEER_HANDLE_BYTES), off, ctypes.byref(out))
                        self._opened.append(handle)
                        ptrs.append(int(out.value))
                    self.base[name] = ptrs
            except Exception as e:  # noqa: BLE001
                err = f"rank {rank}: {e}"
        errors = [e for e in allgather(err) if e] + [e for e, _ in everyone if e]
        if errors:
            self._unmap()
            raise _lib.DgsError("peer-memory set-up failed: " + "; ".join(sorted(set(errors))))
        self._tables: Dict[Tuple[str, int], Any] = {}
        self.sync_table = self.table("__sync__")

    def _unmap(self) -> None:
        for handle in self._opened:
            try:
                _lib.calls.dgs_peer_close(ctypes.create_string_buffer(handle, _lib.PEER_HANDLE_BYTES))
            except _lib.DgsError:
                pass
        self._opened = []

    def close(self) -> None:
        """Unmap the peers' buffers from this process (after this rank's own kernels on them have finished).  Not a
        collective, so that a rank tearing down after an error cannot hang the others: the buffers this rank EXPORTED stay
        valid for as long as their tensors live, and every step ends with a barrier behind the peers' last stores into
        them."""
        if self.base is None:
            return
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self._unmap()
        self.base = None  # type: ignore
        self._tables = {}

    def table(self, name: str, byte_off: int = 0) -> Any:
        """(void*)[size]: every rank's pointer to byte `byte_off` of buffer `name`, as mapped in this process."""
        key = (name, byte_off)
        t = self._tables.get(key)
        if t is None:
            t = self._tables[key] = (ctypes.c_void_p * self.size)(*[b + byte_off for b in self.base[name]])
        return t

    def _next(self, slot: int) -> int:
        self._epoch[slot] = (self._epoch[slot] + 1) & 0xFFFFFFFF
        return self._epoch[slot]

    def barrier(self, slot: int, stream: Any) -> None:
        _lib.calls.dgs_peer_barrier(self.sync_table, self.size, self.rank, slot, self._next(slot), self.timeout_s, stream)

    def norm_exchange(self, partials_ptr: Any, n_partials: int, out_norm_ptr: Any, stream: Any) -> None:
        _lib.calls.dgs_peer_norm_exchange(self.sync_table, self.size, self.rank, S_NORM, self._next(S_NORM), partials_ptr,
                                          n_partials, out_norm_ptr, self.timeout_s, stream)

    def check(self) -> None:
        """Raise if a device-side wait gave up (a rank never arrived): better than a hang, and no GPU fault."""
        code = ctypes.c_int(0)
        _lib.calls.dgs_peer_error(ctypes.byref(code))
        if code.value:
            slot, peer = (code.value - 1000) // 16, (code.value - 1000) % 16
            raise _lib.DgsError(
                f"peer exchange: rank {self.rank} waited longer than {self.timeout_s or 120.0:.0f} s for rank {peer} "
                f"(sync slot {slot}); the ranks no longer run the same sequence of steps"
            )
