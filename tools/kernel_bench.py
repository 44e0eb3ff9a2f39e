#!/usr/bin/env python
"""Per-kernel roofline microbenchmark (one GPU): times each libdgs kernel alone with CUDA events on
inputs larger than L2 and prints achieved algorithmic GB/s against MEASURED_PEAKS.json's HBM copy
bandwidth.  Algorithmic bytes per element are the ones DESIGN.md states.

    python tools/kernel_bench.py [--mb 1024] [--iters 20]
"""
import argparse
import ctypes
import json
import pathlib
import sys

import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from determined_clone_b200 import _build, _lib as L  # noqa: E402


WARMUP = None


def peak_gbs() -> float:
    try:
        return float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    except Exception:
        return 6650.0  # B200_PROFILING.md fallback


def ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def time_it(fn, iters, warmup=3):
    warmup = WARMUP if WARMUP is not None else warmup
    st = torch.cuda.current_stream()
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record(st)
        fn()
        b.record(st)
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=1024, help="size of one fp32 array in MiB")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=None)
    args = ap.parse_args()
    global WARMUP
    WARMUP = args.warmup
    _build.build()
    n = args.mb * (1 << 20) // 4
    peak = peak_gbs()
    s = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    results = []

    def report(name, bytes_per_elem, fn):
        med, best = time_it(fn, args.iters)
        gbs = bytes_per_elem * n / (med * 1e-3) / 1e9
        results.append(dict(kernel=name, n=n, bytes_per_elem=bytes_per_elem, ms_median=round(med, 4), ms_best=round(best, 4),
                            achieved_gbs=round(gbs, 1), frac_of_measured_peak=round(gbs / peak, 3)))
        print(json.dumps(results[-1]), flush=True)

    for pdt, gdt, tag in [(torch.float32, torch.float32, "f32"), (torch.bfloat16, torch.bfloat16, "bf16")]:
        ps, gs = pdt.itemsize, gdt.itemsize
        g = (torch.randn(n, device="cuda") * 0.01).to(gdt)
        acc = torch.zeros(n, device="cuda")
        wire = torch.zeros(n, device="cuda", dtype=gdt)
        seg = (L.Seg * 1)()
        seg[0].src, seg[0].dst_off, seg[0].numel = g.data_ptr(), 0, n
        report(f"flatten first (acc=g) {tag}", gs + 4,
               lambda: L.calls.dgs_flatten_accumulate(seg, 1, L.dtype_code(gdt), ptr(acc), None, 0, 0, 1.0, s))
        report(f"flatten add (acc+=g) {tag}", gs + 8,
               lambda: L.calls.dgs_flatten_accumulate(seg, 1, L.dtype_code(gdt), ptr(acc), None, 0, L.READ_ACC, 1.0, s))
        report(f"flatten final (wire=(acc+g)*s) {tag}", gs + 4 + gs,
               lambda: L.calls.dgs_flatten_accumulate(seg, 1, L.dtype_code(gdt), ptr(acc), ptr(wire), L.dtype_code(gdt),
                                                      L.READ_ACC | L.WRITE_WIRE, 0.25, s))
        report(f"flatten A=1 (wire=g*s) {tag}", 2 * gs,
               lambda: L.calls.dgs_flatten_accumulate(seg, 1, L.dtype_code(gdt), None, ptr(wire), L.dtype_code(gdt), L.WRITE_WIRE, 0.25, s))
        partials = torch.zeros(L.NORM_MAX_PARTIALS, device="cuda")
        norm = torch.zeros(1, device="cuda")
        report(f"l2norm (2 passes) {tag}", gs, lambda: L.calls.dgs_l2norm(ptr(wire), L.dtype_code(gdt), n, ptr(partials), ptr(norm), s))
        p = torch.randn(n, device="cuda").to(pdt)
        m = torch.zeros(n, device="cuda", dtype=pdt)
        v = torch.zeros(n, device="cuda", dtype=pdt)
        ar = (L.AdamRange * 1)()
        ar[0].begin, ar[0].end, ar[0].lr, ar[0].beta1, ar[0].beta2, ar[0].eps, ar[0].weight_decay = 0, n, 3e-4, 0.9, 0.95, 1e-8, 0.1
        ar[0].step_size, ar[0].bias_correction2_sqrt, ar[0].flags = 3e-4 / (1 - 0.9), (1 - 0.95) ** 0.5, L.ADAM_DECOUPLED_WD
        report(f"clip+AdamW step {tag}", gs + 6 * ps,
               lambda: L.calls.dgs_step_adam(ptr(p), ptr(m), ptr(v), L.dtype_code(pdt), ptr(wire), L.dtype_code(gdt), n, ar, 1, ptr(norm), 1.0, s))
        sr = (L.SgdRange * 1)()
        sr[0].begin, sr[0].end, sr[0].lr, sr[0].momentum, sr[0].dampening, sr[0].weight_decay, sr[0].flags = 0, n, 0.1, 0.9, 0.0, 1e-4, 0
        report(f"clip+SGD-momentum step {tag}", gs + 4 * ps,
               lambda: L.calls.dgs_step_sgd(ptr(p), ptr(m), L.dtype_code(pdt), ptr(wire), L.dtype_code(gdt), n, sr, 1, ptr(norm), 1.0, s))
        del g, acc, wire, p, m, v
        torch.cuda.empty_cache()
    # torch's own copy as the local yardstick for the peak figure
    a = torch.empty(n, device="cuda")
    b = torch.empty(n, device="cuda")
    med, best = time_it(lambda: b.copy_(a), args.iters)
    print(json.dumps(dict(kernel="torch copy_ (yardstick)", n=n, ms_median=round(med, 4), achieved_gbs=round(8 * n / (med * 1e-3) / 1e9, 1),
                          peak_used_gbs=peak)), flush=True)


if __name__ == "__main__":
    main()
