"""Replay the reference's recorded runs through the C ABI, kernel by kernel, and compare every
intermediate the reference exposes (synced gradients, total norm, parameters, optimizer state) with
the golden fixtures.  Ranks are emulated on one GPU: each rank's wire bucket is produced by the
flatten kernel and the buckets are summed the way the allreduce would."""
import ctypes

import numpy as np
import pytest

from golden_io import GROUP_OF, N_PARAMS, OPT_ORDER, SCENARIOS, Golden, assert_close
from test_gpu_kernels import CODE, TD, adam_ranges, align_up, dev, host, make_segs, ptr, sgd_ranges, stream

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from determined_clone_b200 import _lib as L  # noqa: E402

FUSED = [n for n, sc in SCENARIOS.items() if sc["hp"]["optimizer"] in ("sgd", "adam", "adamw")]


@pytest.mark.parametrize("name", FUSED)
def test_replay(lib, name):
    g = Golden(name)
    dt = g.dtype
    # identical inputs (the recorded gradients): fp32 is held to the purely element-wise rule, 1e-5*|x| + 1e-7
    rule = "elem" if dt == "float32" else None
    init = g.init_params()
    # flat layout in optimizer order, every tensor 128-byte aligned
    offs, off = {}, 0
    for i in OPT_ORDER:
        offs[i] = off
        off = align_up(off + init[i].size)
    total = off
    n_decay = align_up(offs[4] + init[4].size)  # group 0 = [0, n_decay), group 1 = [n_decay, total)

    def flat_of(tensors):
        f = np.zeros(total, dtype=np.float32)
        for i in range(N_PARAMS):
            f[offs[i]:offs[i] + tensors[i].size] = np.asarray(tensors[i], dtype=np.float32).ravel()
        return f

    p = dev(flat_of(init), dt)
    s1 = torch.zeros(total, device="cuda", dtype=TD[dt])  # exp_avg / momentum buffer
    s2 = torch.zeros(total, device="cuda", dtype=TD[dt])  # exp_avg_sq
    acc = torch.zeros(total, device="cuda")
    norm = torch.zeros(1, device="cuda")
    partials = torch.zeros(L.calls.dgs_l2norm_grid(total), device="cuda")
    scale = 1.0 / (g.slots * (g.agg if g.average else 1))
    groups = g.groups()
    kind = g.hp["optimizer"]

    for s in range(g.steps):
        raw = g.raw(s)
        wires = []
        for r in range(g.slots):
            wire = torch.zeros(total, device="cuda", dtype=TD[dt])
            for m in range(g.agg):
                ts = [dev(raw[r][m][i].ravel(), dt) for i in OPT_ORDER]
                segs = make_segs(ts, [offs[i] for i in OPT_ORDER])
                first, last = m == 0, m == g.agg - 1
                mode = (0 if first else L.READ_ACC) | (L.WRITE_WIRE if last else 0)
                L.calls.dgs_flatten_accumulate(segs, N_PARAMS, CODE[dt], ptr(acc), ptr(wire), CODE[dt], mode, scale, stream())
            wires.append(wire)
        wire = wires[0]
        for w in wires[1:]:
            wire = wire + w  # ncclAllReduce(sum) in the wire dtype
        want_avg = flat_of(g.per_step("avg_grads", s))
        assert_close(host(wire), want_avg, dt, f"{name}: synced gradients, step {s}", rule=rule)

        clip_ptr, max_norm = None, 0.0
        if g.clip is not None and g.clip[0] == "norm":
            L.calls.dgs_l2norm(ptr(wire), CODE[dt], total, ptr(partials), ptr(norm), stream())
            clip_ptr, max_norm = ptr(norm), g.clip[1]
        elif g.clip is not None:
            wire.clamp_(-g.clip[1], g.clip[1])  # clip_grad_value_: the user's callable runs on the views

        if kind == "sgd":
            specs = [dict(begin=b, end=e, lr=gr["lr"], momentum=gr.get("momentum", 0.0), dampening=gr.get("dampening", 0.0),
                          weight_decay=gr["weight_decay"], nesterov=gr.get("nesterov", False), first=(s == 0))
                     for (b, e), gr in zip([(0, n_decay), (n_decay, total)], groups)]
            L.calls.dgs_step_sgd(ptr(p), ptr(s1), CODE[dt], ptr(wire), CODE[dt], total, sgd_ranges(specs), 2, clip_ptr, max_norm, stream())
        else:
            specs = [dict(begin=b, end=e, lr=gr["lr"], betas=tuple(gr["betas"]), eps=gr["eps"], weight_decay=gr["weight_decay"],
                          step=s + 1, decoupled=(kind == "adamw"))
                     for (b, e), gr in zip([(0, n_decay), (n_decay, total)], groups)]
            L.calls.dgs_step_adam(ptr(p), ptr(s1), ptr(s2), CODE[dt], ptr(wire), CODE[dt], total, adam_ranges(specs), 2, clip_ptr, max_norm, stream())
        torch.cuda.synchronize()

        if clip_ptr is not None:
            assert_close(float(norm), g.z["total_norm"][s], dt, f"{name}: total norm, step {s}", rule=rule)
        assert_close(host(p), flat_of(g.per_step("params", s)), dt, f"{name}: parameters, step {s}", rule=rule)
        if kind == "sgd":
            assert_close(host(s1), flat_of([g.state(s, i, "momentum_buffer") for i in range(N_PARAMS)]), dt, f"{name}: momentum, step {s}", rule=rule)
        else:
            assert_close(host(s1), flat_of([g.state(s, i, "exp_avg") for i in range(N_PARAMS)]), dt, f"{name}: exp_avg, step {s}", rule=rule)
            assert_close(host(s2), flat_of([g.state(s, i, "exp_avg_sq") for i in range(N_PARAMS)]), dt, f"{name}: exp_avg_sq, step {s}", rule=rule)
