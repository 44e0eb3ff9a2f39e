"""TEST INFRASTRUCTURE ONLY: a host-memory stand-in for the part of libdgs.so's C ABI that the training path calls
(the communicator entry points over torch.distributed/gloo), so that the Python side of the path -- engine bookkeeping, context, training loop, checkpoints, AMP protocol -- can
be exercised end to end in the CPU test suite, where no kernel can run.

`install(monkeypatch)` swaps `determined_clone_b200._lib.calls` for this emulator and lifts the engine's "CUDA device
required" check for the duration of one test.  Nothing under determined_clone_b200/ imports this module, and the
product keeps failing loudly without a GPU; the kernels themselves are tested against the oracle on a GPU
(tests/test_gpu_*.py).  The arithmetic below restates what each entry point is documented to do in include/dgs.h
with torch CPU ops in fp32; it makes no claim about the kernels' rounding beyond the usual fp32 tolerance.

Pointers arrive as ctypes.c_void_p / ints holding HOST addresses (CPU tensors), which is what makes this possible."""
import ctypes
from typing import Any, Optional

import torch

from determined_clone_b200 import _lib

_U8 = -1  # raw bytes
_TORCH = {_U8: torch.uint8, _lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16, _lib.F64: torch.float64, _lib.I64: torch.int64}


def _addr(p: Any) -> int:
    if p is None:
        return 0
    if isinstance(p, ctypes.c_void_p):
        return p.value or 0
    return int(p)


def _view(p: Any, dtype_code: int, n: int) -> Optional[torch.Tensor]:
    """n elements of the given dtype at host address p, as a tensor sharing that memory."""
    a = _addr(p)
    if a == 0 or n == 0:
        return None if a == 0 else torch.empty(0, dtype=_TORCH[dtype_code])
    dt = _TORCH[dtype_code]
    raw = (ctypes.c_char * (n * torch.empty(0, dtype=dt).element_size())).from_address(a)
    return torch.frombuffer(raw, dtype=dt, count=n)


def _scalar(p: Any) -> Optional[float]:
    v = _view(p, _lib.F32, 1)
    return None if v is None else float(v[0])


class FakeCalls:
    """Same call signatures as _lib.calls (status checking included: errors are raised as DgsError)."""

    def __init__(self) -> None:
        self.launches = 0
        self.log: list = []

    # ---- bookkeeping ------------------------------------------------------------------------------------------
    def dgs_launch_count(self) -> int:
        return self.launches

    def dgs_timer_enable(self, enabled: int) -> None:
        pass

    def dgs_timer_read(self, out: Any, capacity: int, n_out: Any) -> None:
        n_out._obj.value = 0

    # ---- flatten ----------------------------------------------------------------------------------------------
    def dgs_flatten_accumulate(self, segs: Any, n: int, gdt: int, acc: Any, wire: Any, wdt: int, mode: int, scale: float, stream: Any) -> None:
        self.launches += 1
        self.log.append(("flatten", n, mode, float(scale)))
        for k in range(n):
            s = segs[k]
            if s.dst_off % 8:
                raise _lib.DgsError("dgs_flatten_accumulate: dst_off must be a multiple of 8 elements")
            g = _view(s.src, gdt, s.numel).float()
            total = g
            if _addr(acc):
                a = _view(_addr(acc) + 4 * s.dst_off, _lib.F32, s.numel)
                if mode & _lib.READ_ACC:
                    total = a + g
                if not (mode & _lib.WRITE_WIRE):
                    a.copy_(total)
            elif mode & _lib.READ_ACC:
                raise _lib.DgsError("dgs_flatten_accumulate: DGS_READ_ACC without an accumulator")
            if mode & _lib.WRITE_WIRE:
                isz = torch.empty(0, dtype=_TORCH[wdt]).element_size()
                w = _view(_addr(wire) + isz * s.dst_off, wdt, s.numel)
                w.copy_((total * torch.tensor(scale, dtype=torch.float32)).to(w.dtype))

    # ---- norm -------------------------------------------------------------------------------------------------
    def dgs_l2norm_grid(self, n: int) -> int:
        return 1 if n > 0 else 0

    def dgs_l2norm_partial(self, x: Any, dt: int, n: int, partials: Any, stream: Any) -> None:
        self.launches += 1
        if n > 0:
            _view(partials, _lib.F32, 1)[0] = float(_view(x, dt, n).double().pow(2).sum())

    def dgs_l2norm_final(self, partials: Any, n_partials: int, out: Any, stream: Any) -> None:
        self.launches += 1
        total = float(_view(partials, _lib.F32, n_partials).double().sum()) if n_partials else 0.0
        _view(out, _lib.F32, 1)[0] = total ** 0.5 if total == total and total >= 0 else float("nan")

    def dgs_clip_scale(self, g: Any, dt: int, n: int, total_norm: Any, max_norm: float, stream: Any) -> None:
        self.launches += 1
        v = _view(g, dt, n)
        v.copy_((v.float() * self._coef(total_norm, max_norm, None)).to(v.dtype))

    @staticmethod
    def _coef(total_norm: Any, max_norm: float, inv_scale: Any) -> torch.Tensor:
        inv = torch.tensor(_scalar(inv_scale) if _addr(inv_scale) else 1.0, dtype=torch.float32)
        if not _addr(total_norm):
            return inv
        c = torch.tensor(max_norm, dtype=torch.float32) / (torch.tensor(_scalar(total_norm), dtype=torch.float32) * inv + 1e-6)
        return torch.where(c > 1.0, torch.ones((), dtype=torch.float32), c) * inv  # NaN stays NaN

    # ---- steps ------------------------------------------------------------------------------------------------
    def dgs_step_adam(self, p, m, v, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, stream, inv_scale=None) -> None:
        self.launches += 1
        self.log.append(("adam", n_ranges))
        coef = self._coef(total_norm, max_norm, inv_scale)
        P, M, V, G = _view(p, sdt, n), _view(m, sdt, n), _view(v, sdt, n), _view(g, gdt, n)
        f32 = lambda x: torch.tensor(x, dtype=torch.float32)  # noqa: E731
        for i in range(n_ranges):
            r = ranges[i]
            sl = slice(r.begin, r.end)
            pp, mm, vv, gg = P[sl].float(), M[sl].float(), V[sl].float(), G[sl].float() * coef
            if r.flags & _lib.OPT_MAXIMIZE:
                gg = -gg
            if r.flags & _lib.ADAM_DECOUPLED_WD:
                pp = pp * f32(1.0 - r.lr * r.weight_decay)
            else:
                gg = gg + f32(r.weight_decay) * pp
            mm = mm + f32(1.0 - r.beta1) * (gg - mm)
            vv = vv * f32(r.beta2) + f32(1.0 - r.beta2) * (gg * gg)
            denom = vv.sqrt() * f32(1.0 / r.bias_correction2_sqrt) + f32(r.eps)
            pp = pp + f32(-r.step_size) * (mm / denom)
            P[sl], M[sl], V[sl] = pp.to(P.dtype), mm.to(M.dtype), vv.to(V.dtype)

    def dgs_step_adam_scaled(self, p, m, v, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, inv_scale, stream) -> None:
        self.dgs_step_adam(p, m, v, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, stream, inv_scale=inv_scale)

    def dgs_step_sgd(self, p, buf, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, stream, inv_scale=None) -> None:
        self.launches += 1
        self.log.append(("sgd", n_ranges))
        coef = self._coef(total_norm, max_norm, inv_scale)
        P, G = _view(p, sdt, n), _view(g, gdt, n)
        B = _view(buf, sdt, n)
        f32 = lambda x: torch.tensor(x, dtype=torch.float32)  # noqa: E731
        for i in range(n_ranges):
            r = ranges[i]
            sl = slice(r.begin, r.end)
            pp, gg = P[sl].float(), G[sl].float() * coef
            if r.flags & _lib.OPT_MAXIMIZE:
                gg = -gg
            if r.weight_decay != 0:
                gg = gg + f32(r.weight_decay) * pp
            if B is not None:
                bb = gg.clone() if r.flags & _lib.SGD_FIRST_STEP else B[sl].float() * f32(r.momentum) + f32(1.0 - r.dampening) * gg
                B[sl] = bb.to(B.dtype)
                gg = gg + f32(r.momentum) * bb if r.flags & _lib.SGD_NESTEROV else bb
            elif r.momentum != 0:
                raise _lib.DgsError("dgs_step_sgd: momentum without a momentum buffer")
            P[sl] = (pp + f32(-r.lr) * gg).to(P.dtype)

    def dgs_step_sgd_scaled(self, p, buf, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, inv_scale, stream) -> None:
        self.dgs_step_sgd(p, buf, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, stream, inv_scale=inv_scale)


    # ---- communicator: torch.distributed (gloo) on host memory stands in for NCCL; everything completes at once ----
    def dgs_comm_unique_id(self, out128: Any) -> None:
        pass

    def dgs_comm_init(self, ident: Any, rank: int, nranks: int, device: int, out_handle: Any) -> None:
        import torch.distributed as dist

        if not dist.is_initialized() or dist.get_world_size() != nranks or dist.get_rank() != rank:
            raise _lib.DgsError("fake communicator: torch.distributed (gloo) must be initialised with the same rank / size")
        out_handle._obj.value = 1

    def dgs_comm_destroy(self, handle: Any) -> None:
        pass

    def dgs_comm_allreduce_async(self, handle: Any, buf: Any, count: int, dt: int, stream: Any) -> None:
        import torch.distributed as dist

        self.log.append(("allreduce", count))
        dist.all_reduce(_view(buf, dt, count))

    def dgs_comm_reduce_scatter_async(self, handle: Any, buf: Any, count_per_rank: int, dt: int, stream: Any) -> None:
        import torch.distributed as dist

        self.log.append(("reduce_scatter", count_per_rank))
        n, r = dist.get_world_size(), dist.get_rank()
        whole = _view(buf, dt, count_per_rank * n)
        summed = whole.clone()
        dist.all_reduce(summed)
        # in place, as ncclReduceScatter with recvbuff inside sendbuff: only this rank's slice is written
        whole[r * count_per_rank:(r + 1) * count_per_rank] = summed[r * count_per_rank:(r + 1) * count_per_rank]

    def dgs_comm_broadcast_async(self, handle: Any, buf: Any, nbytes: int, root: int, stream: Any) -> None:
        import torch.distributed as dist

        self.log.append(("broadcast", nbytes))
        dist.broadcast(_view(buf, _U8, nbytes), src=root)

    def dgs_comm_allgather_async(self, handle: Any, send: Any, recv: Any, nbytes: int, stream: Any) -> None:
        import torch.distributed as dist

        self.log.append(("allgather", nbytes))
        n = dist.get_world_size()
        mine = _view(send, _U8, nbytes).clone()  # send may alias this rank's slice of recv (in-place all-gather)
        parts = [torch.empty(nbytes, dtype=torch.uint8) for _ in range(n)]
        dist.all_gather(parts, mine)
        out = _view(recv, _U8, nbytes * n)
        for k, part in enumerate(parts):
            out[k * nbytes:(k + 1) * nbytes] = part

    def dgs_comm_l2norm_partial_async(self, handle: Any, x: Any, dt: int, n: int, partials: Any) -> None:
        self.dgs_l2norm_partial(x, dt, n, partials, None)

    def dgs_comm_wait(self, handle: Any, stream: Any) -> None:
        pass

    def dgs_l2norm_final_sumsq(self, partials: Any, n_partials: int, out: Any, stream: Any) -> None:
        self.launches += 1
        _view(out, _lib.F32, 1)[0] = float(_view(partials, _lib.F32, n_partials).double().sum()) if n_partials else 0.0

    def dgs_sqrt_inplace(self, x: Any, stream: Any) -> None:
        v = _view(x, _lib.F32, 1)
        v[0] = float(v[0]) ** 0.5 if float(v[0]) >= 0 else float("nan")


    # ---- single-launch step (include/dgs.h section 3b) -------------------------------------------------------------
    def _direct(self, kind: str, segs, seg_range, n_segs, gdt, acc, scale, p, states, sdt, ranges, n_ranges, clip, max_norm,
                partials, cap, out_norm, stream) -> None:
        self.launches += 1
        self.log.append(("direct_" + kind, n_segs, int(clip)))
        dt = _TORCH[sdt]
        isz = torch.empty(0, dtype=dt).element_size()
        total_sq = 0.0
        grads = []
        for k in range(n_segs):
            sg = segs[k]
            g = _view(sg.src, gdt, sg.numel).float()
            if _addr(acc):
                g = _view(_addr(acc) + 4 * sg.dst_off, _lib.F32, sg.numel) + g
            g = (g * torch.tensor(scale, dtype=torch.float32)).to(dt)
            grads.append(g)
            total_sq += float(g.double().pow(2).sum())
        norm = torch.zeros(1, dtype=torch.float32)
        if clip:
            norm[0] = total_sq ** 0.5
            _view(out_norm, _lib.F32, 1)[0] = float(norm[0])
        for k in range(n_segs):
            sg = segs[k]
            one = type(ranges[0])()
            for f, _t in one._fields_:
                setattr(one, f, getattr(ranges[seg_range[k]], f))
            one.begin, one.end = 0, sg.numel
            arr = (type(one) * 1)(one)
            ptrs = [_addr(x) + isz * sg.dst_off if _addr(x) else None for x in [p] + list(states)]
            tn = norm.data_ptr() if clip else None
            if kind == "adam":
                self.dgs_step_adam(ptrs[0], ptrs[1], ptrs[2], sdt, grads[k].data_ptr(), sdt, sg.numel, arr, 1, tn, max_norm, stream)
            else:
                self.dgs_step_sgd(ptrs[0], ptrs[1], sdt, grads[k].data_ptr(), sdt, sg.numel, arr, 1, tn, max_norm, stream)
            self.launches -= 1

    def dgs_direct_step_adam(self, segs, seg_range, n_segs, gdt, acc, scale, p, m, v, sdt, ranges, n_ranges, clip, max_norm,
                             partials, cap, out_norm, stream) -> None:
        self._direct("adam", segs, seg_range, n_segs, gdt, acc, scale, p, [m, v], sdt, ranges, n_ranges, clip, max_norm, partials, cap, out_norm, stream)

    def dgs_direct_step_sgd(self, segs, seg_range, n_segs, gdt, acc, scale, p, buf, sdt, ranges, n_ranges, clip, max_norm,
                            partials, cap, out_norm, stream) -> None:
        self._direct("sgd", segs, seg_range, n_segs, gdt, acc, scale, p, [buf], sdt, ranges, n_ranges, clip, max_norm, partials, cap, out_norm, stream)

    # ---- peer exchange (include/dgs.h section 6) --------------------------------------------------------------------
    # Remote memory is another PROCESS's host memory here, so a "peer pointer" is a token: (owner rank + 1) << 56 plus
    # the address in the owner's process.  Stores to tokens are queued and delivered when the ranks meet at the next
    # barrier (all_gather_object over gloo) -- which is also when the real kernels' stores are guaranteed visible.
    _TAG = 56

    def _peer_state(self) -> Any:
        if not hasattr(self, "_mail"):
            self._mail = []  # (dest rank, dest address, dtype code, tensor)
        return self._mail

    def dgs_peer_export(self, ptr: Any, handle_out: Any, offset_out: Any) -> None:
        import torch.distributed as dist

        handle_out.raw = bytes([dist.get_rank() + 1]) + bytes(_lib.PEER_HANDLE_BYTES - 1)
        offset_out._obj.value = _addr(ptr)

    def dgs_peer_open(self, handle: Any, offset: int, out: Any) -> None:
        out._obj.value = (handle.raw[0] << self._TAG) + offset

    def dgs_peer_close_all(self) -> None:
        pass

    def dgs_peer_close(self, handle: Any) -> None:
        self.log.append(("peer_close",))

    def dgs_peer_error(self, out: Any) -> None:
        out._obj.value = 0

    def dgs_peer_error_clear(self) -> None:
        pass

    def _store(self, token: int, dt: int, data: torch.Tensor) -> None:
        owner = (token >> self._TAG) - 1
        if owner < 0:  # a local address
            _view(token, dt, data.numel()).copy_(data)
        else:
            self._peer_state().append((owner, token & ((1 << self._TAG) - 1), dt, data.clone()))

    def _deliver(self) -> None:
        import torch.distributed as dist

        boxes = [None] * dist.get_world_size()
        dist.all_gather_object(boxes, self._peer_state())
        me = dist.get_rank()
        for box in boxes:
            for dest, addr, dt, data in box:
                if dest == me:
                    _view(addr, dt, data.numel()).copy_(data)
        self._mail = []

    def dgs_peer_barrier(self, sync: Any, nranks: int, rank: int, slot: int, epoch: int, timeout: float, stream: Any) -> None:
        self.launches += 1
        self.log.append(("peer_barrier", slot, epoch))
        self._deliver()

    def dgs_flatten_push(self, segs: Any, n: int, gdt: int, acc: Any, wire_peers: Any, wdt: int, mode: int, scale: float,
                         begin: int, span: int, nranks: int, rank: int, stream: Any) -> None:
        self.launches += 1
        self.log.append(("flatten_push", n, mode, float(scale)))
        assert mode & _lib.WRITE_WIRE
        isz = torch.empty(0, dtype=_TORCH[wdt]).element_size()
        for k in range(n):
            s = segs[k]
            total = _view(s.src, gdt, s.numel).float()
            if mode & _lib.READ_ACC:
                total = _view(_addr(acc) + 4 * s.dst_off, _lib.F32, s.numel) + total
            vals = (total * torch.tensor(scale, dtype=torch.float32)).to(_TORCH[wdt])
            f0 = s.dst_off
            while f0 < s.dst_off + s.numel:  # split the tensor at slice boundaries
                j = (f0 - begin) // span
                f1 = min(s.dst_off + s.numel, begin + (j + 1) * span)
                self._store(_addr(wire_peers[j]) + isz * (f0 + (rank - j) * span), wdt, vals[f0 - s.dst_off:f1 - s.dst_off])
                f0 = f1

    def dgs_peer_reduce(self, bucket: Any, dt: int, span: int, nranks: int, rank: int, partials: Any, stream: Any) -> None:
        self.launches += 1
        self.log.append(("peer_reduce", span))
        rows = _view(bucket, dt, nranks * span).view(nranks, span)
        acc = rows[0].float().clone()
        for r in range(1, nranks):
            acc += rows[r].float()
        rows[rank].copy_(acc.to(rows.dtype))
        if span > 0:
            _view(partials, _lib.F32, 1)[0] = float(rows[rank].double().pow(2).sum())

    def dgs_peer_reduce_multi(self, wire: Any, dt: int, begins: Any, spans: Any, n_buckets: int, nranks: int, rank: int,
                              partials: Any, capacity: int, stream: Any) -> None:
        isz = torch.empty(0, dtype=_TORCH[dt]).element_size()
        _view(partials, _lib.F32, capacity).zero_()
        total = 0.0
        for b in range(n_buckets):
            one = torch.zeros(1, dtype=torch.float32)
            self.dgs_peer_reduce(_addr(wire) + isz * begins[b], dt, spans[b], nranks, rank, one.data_ptr(), stream)
            total += float(one[0])
        _view(partials, _lib.F32, 1)[0] = total

    def dgs_peer_scatter(self, bases: Any, offs: Any, sizes: Any, n: int, nranks: int, rank: int, stream: Any) -> None:
        self.launches += 1
        self.log.append(("peer_scatter", n))
        for i in range(n):
            mine = _view(_addr(bases[rank]) + offs[i], _U8, sizes[i])
            for r in range(nranks):
                if r != rank:
                    self._store(_addr(bases[r]) + offs[i], _U8, mine)

    def dgs_peer_norm_exchange(self, sync: Any, nranks: int, rank: int, slot: int, epoch: int, partials: Any, n_partials: int,
                               out: Any, timeout: float, stream: Any) -> None:
        import torch.distributed as dist

        self.launches += 1
        self.log.append(("peer_norm_exchange", epoch))
        mine = torch.tensor([float(_view(partials, _lib.F32, n_partials).double().sum()) if n_partials else 0.0], dtype=torch.float64)
        parts = [torch.zeros_like(mine) for _ in range(nranks)]
        dist.all_gather(parts, mine)
        self._deliver()
        total = sum(float(p) for p in parts)
        _view(out, _lib.F32, 1)[0] = total ** 0.5 if total == total and total >= 0 else float("nan")

    def _push_params(self, p: Any, sdt: int, ranges: Any, n_ranges: int, peers: Any, nranks: int, rank: int) -> None:
        """Only elements inside a range were stepped, and only those are stored to the peers (the gaps between this
        rank's slices belong to other ranks, whose fresh values must not be overwritten)."""
        isz = torch.empty(0, dtype=_TORCH[sdt]).element_size()
        for i in range(n_ranges):
            b, e = ranges[i].begin, ranges[i].end
            mine = _view(_addr(p) + isz * b, sdt, e - b)
            for r in range(nranks):
                if r != rank:
                    self._store(_addr(peers[r]) + isz * b, sdt, mine)

    def dgs_step_adam_push(self, p, m, v, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, peers, nranks, rank, stream) -> None:
        assert _addr(peers[rank]) == _addr(p)
        self.dgs_step_adam(p, m, v, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, stream)
        self._push_params(p, sdt, ranges, n_ranges, peers, nranks, rank)

    def dgs_step_sgd_push(self, p, buf, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, peers, nranks, rank, stream) -> None:
        assert _addr(peers[rank]) == _addr(p)
        self.dgs_step_sgd(p, buf, sdt, g, gdt, n, ranges, n_ranges, total_norm, max_norm, stream)
        self._push_params(p, sdt, ranges, n_ranges, peers, nranks, rank)


class _NoStream:
    cuda_stream = 0

    def synchronize(self) -> None:
        pass


def install(monkeypatch: Any = None) -> FakeCalls:
    """Route the engine's C-ABI calls to the emulator and let engines be built on the CPU: for the current test
    when given pytest's monkeypatch, for the rest of the process otherwise (spawned rank processes)."""
    from determined_clone_b200.pytorch import _comm, _gradsync, _pytorch_context

    fake = FakeCalls()
    patches = [(_lib, "calls", fake), (_gradsync, "_check_device", lambda device: None), (_gradsync, "_stream", lambda: None),
               (_comm, "_stream", lambda: None), (torch.cuda, "current_device", lambda: 0),
               (torch.cuda, "current_stream", lambda *a, **k: _NoStream())]
    for obj, name, value in patches:
        if monkeypatch is not None:
            monkeypatch.setattr(obj, name, value)
        else:
            setattr(obj, name, value)
    assert _gradsync._lib is _lib and _pytorch_context._lib is _lib and _comm._lib is _lib  # all reach `calls` through the module
    return fake
