"""TEST INFRASTRUCTURE ONLY: a host-memory stand-in for the part of libdgs.so's C ABI that the single-rank path calls,
so that the Python side of the path -- engine bookkeeping, context, training loop, checkpoints, AMP protocol -- can
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

_TORCH = {_lib.F32: torch.float32, _lib.BF16: torch.bfloat16, _lib.F16: torch.float16, _lib.F64: torch.float64, _lib.I64: torch.int64}


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


def install(monkeypatch: Any) -> FakeCalls:
    """Route the engine's C-ABI calls to the emulator for the current test and let engines be built on the CPU."""
    from determined_clone_b200.pytorch import _gradsync, _pytorch_context

    fake = FakeCalls()
    monkeypatch.setattr(_lib, "calls", fake)
    monkeypatch.setattr(_gradsync, "_check_device", lambda device: None)
    monkeypatch.setattr(_gradsync, "_stream", lambda: None)
    assert _gradsync._lib is _lib and _pytorch_context._lib is _lib  # both reach `calls` through the module attribute
    return fake
