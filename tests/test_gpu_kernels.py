"""GPU parity tests of every kernel, called through the C ABI (ctypes -> libdgs.so), against the
numpy oracle.  torch is used only to own device memory and move data."""
import ctypes

import numpy as np
import pytest

from oracle import gradsync_oracle as O

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from determined_clone_b200 import _lib as L  # noqa: E402

TD = {"float32": torch.float32, "bfloat16": torch.bfloat16, "float16": torch.float16}
CODE = {"float32": L.F32, "bfloat16": L.BF16, "float16": L.F16}


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def dev(a, dtype="float32"):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to("cuda").to(TD[dtype])


def host(t):
    return t.detach().float().cpu().numpy()


def align_up(n, a=64):
    return (n + a - 1) // a * a


def layout(sizes):
    offs, off = [], 0
    for n in sizes:
        offs.append(off)
        off = align_up(off + n)
    return offs, off


def make_segs(tensors, offs):
    segs = (L.Seg * len(tensors))()
    for i, (t, o) in enumerate(zip(tensors, offs)):
        segs[i].src, segs[i].dst_off, segs[i].numel = t.data_ptr(), o, t.numel()
    return segs


SIZES = [1, 7, 64, 1000, 4096, 4097, 70001, 0, 3 * 4096 + 5]


@pytest.mark.parametrize("gdt", ["float32", "bfloat16", "float16"])
@pytest.mark.parametrize("wdt", ["float32", "bfloat16", "float16"])
@pytest.mark.parametrize("misalign", [0, 1])
def test_flatten_all_modes_bit_exact(lib, gdt, wdt, misalign):
    rng = np.random.default_rng(1)
    offs, total = layout(SIZES)
    scale = np.float32(1.0 / 12.0)
    micro = []
    for m in range(3):
        ts = []
        for n in SIZES:
            base = dev(rng.standard_normal(n + misalign) * 0.01, gdt)
            ts.append(base[misalign:])  # misalign=1 -> sources that are not 16-byte aligned
        micro.append(ts)
    acc = torch.full((total,), 7.0, device="cuda")  # stale contents must be overwritten by mode 0
    wire = torch.zeros(total, device="cuda", dtype=TD[wdt])
    L.calls.dgs_flatten_accumulate(make_segs(micro[0], offs), len(SIZES), CODE[gdt], ptr(acc), None, 0, 0, 1.0, stream())
    L.calls.dgs_flatten_accumulate(make_segs(micro[1], offs), len(SIZES), CODE[gdt], ptr(acc), None, 0, L.READ_ACC, 1.0, stream())
    L.calls.dgs_flatten_accumulate(make_segs(micro[2], offs), len(SIZES), CODE[gdt], ptr(acc), ptr(wire), CODE[wdt],
                                   L.READ_ACC | L.WRITE_WIRE, float(scale), stream())
    wire1 = torch.zeros(total, device="cuda", dtype=TD[wdt])
    L.calls.dgs_flatten_accumulate(make_segs(micro[0], offs), len(SIZES), CODE[gdt], None, ptr(wire1), CODE[wdt],
                                   L.WRITE_WIRE, float(scale), stream())
    torch.cuda.synchronize()
    acc_h, wire_h, wire1_h = host(acc), host(wire), host(wire1)
    for i, (n, o) in enumerate(zip(SIZES, offs)):
        g = [host(micro[m][i]) for m in range(3)]
        want_acc = (g[0] + g[1]).astype(np.float32)
        np.testing.assert_array_equal(acc_h[o:o + n], want_acc)
        want_wire = O.rnd((want_acc + g[2]) * scale, wdt)
        np.testing.assert_array_equal(wire_h[o:o + n], want_wire)
        np.testing.assert_array_equal(wire1_h[o:o + n], O.rnd(g[0] * scale, wdt))
        # padding between tensors is never written
        end = offs[i + 1] if i + 1 < len(offs) else total
        assert (acc_h[o + n:end] == 7.0).all() and (wire_h[o + n:end] == 0).all()


def test_flatten_many_segments_and_scale_cast(lib):
    rng = np.random.default_rng(2)
    sizes = [int(s) for s in rng.integers(1, 3000, size=150)]  # > 64 segments: several launches
    offs, total = layout(sizes)
    ts = [dev(rng.standard_normal(n)) for n in sizes]
    acc = torch.zeros(total, device="cuda")
    L.calls.dgs_flatten_accumulate(make_segs(ts, offs), len(sizes), L.F32, ptr(acc), None, 0, 0, 1.0, stream())
    wire = torch.zeros(total, device="cuda", dtype=torch.bfloat16)
    L.calls.dgs_scale_cast(ptr(acc), ptr(wire), L.BF16, total, 0.25, stream())
    torch.cuda.synchronize()
    acc_h = host(acc)
    for t, n, o in zip(ts, sizes, offs):
        np.testing.assert_array_equal(acc_h[o:o + n], host(t))
    np.testing.assert_array_equal(host(wire), O.bf16_round(acc_h * np.float32(0.25)))


@pytest.mark.parametrize("dt", ["float32", "bfloat16", "float16"])
@pytest.mark.parametrize("n", [0, 1, 5, 4096, 1 << 20, (1 << 22) + 3])
def test_l2norm(lib, dt, n):
    rng = np.random.default_rng(3)
    x = dev(rng.standard_normal(n + 8) * 0.1, dt)[:n]
    grid = L.calls.dgs_l2norm_grid(n)
    assert 1 <= grid <= L.NORM_MAX_PARTIALS
    partials = torch.full((grid,), float("nan"), device="cuda")
    out = torch.zeros(1, device="cuda")
    L.calls.dgs_l2norm(ptr(x), CODE[dt], n, ptr(partials), ptr(out), stream())
    torch.cuda.synchronize()
    want = np.sqrt(np.sum(np.square(host(x).astype(np.float64))))
    np.testing.assert_allclose(float(out), want, rtol=2e-6)
    # clip tail: g *= min(max_norm / (norm + 1e-6), 1)
    if n:
        g = x.clone()
        L.calls.dgs_clip_scale(ptr(g), CODE[dt], n, ptr(out), 0.5 * float(want), stream())
        torch.cuda.synchronize()
        coef = np.float32(0.5 * float(want)) / (np.float32(float(out)) + np.float32(1e-6))
        coef = min(coef, np.float32(1.0))
        np.testing.assert_array_equal(host(g), O.rnd(host(x) * coef, dt))


def test_l2norm_two_buffers_share_one_final_pass(lib):
    rng = np.random.default_rng(4)
    a, b = dev(rng.standard_normal(100003)), dev(rng.standard_normal(5000), "bfloat16")
    ga, gb = L.calls.dgs_l2norm_grid(a.numel()), L.calls.dgs_l2norm_grid(b.numel())
    partials = torch.zeros(ga + gb, device="cuda")
    out = torch.zeros(1, device="cuda")
    L.calls.dgs_l2norm_partial(ptr(a), L.F32, a.numel(), ptr(partials), stream())
    L.calls.dgs_l2norm_partial(ptr(b), L.BF16, b.numel(), ctypes.c_void_p(partials.data_ptr() + 4 * ga), stream())
    L.calls.dgs_l2norm_final(ptr(partials), ga + gb, ptr(out), stream())
    torch.cuda.synchronize()
    want = np.sqrt(np.sum(host(a).astype(np.float64) ** 2) + np.sum(host(b).astype(np.float64) ** 2))
    np.testing.assert_allclose(float(out), want, rtol=2e-6)


def adam_ranges(specs):
    r = (L.AdamRange * len(specs))()
    for i, s in enumerate(specs):
        b1, b2 = s["betas"]
        t = s["step"]
        r[i].begin, r[i].end = s["begin"], s["end"]
        r[i].lr, r[i].beta1, r[i].beta2, r[i].eps, r[i].weight_decay = s["lr"], b1, b2, s["eps"], s["weight_decay"]
        r[i].step_size = s["lr"] / (1 - b1 ** t)
        r[i].bias_correction2_sqrt = (1 - b2 ** t) ** 0.5
        r[i].flags = (L.ADAM_DECOUPLED_WD if s.get("decoupled", True) else 0) | (L.OPT_MAXIMIZE if s.get("maximize") else 0)
    return r


def sgd_ranges(specs):
    r = (L.SgdRange * len(specs))()
    for i, s in enumerate(specs):
        r[i].begin, r[i].end = s["begin"], s["end"]
        r[i].lr, r[i].momentum, r[i].dampening, r[i].weight_decay = s["lr"], s["momentum"], s.get("dampening", 0.0), s["weight_decay"]
        r[i].flags = ((L.SGD_NESTEROV if s.get("nesterov") else 0) | (L.SGD_FIRST_STEP if s.get("first") else 0)
                      | (L.OPT_MAXIMIZE if s.get("maximize") else 0))
    return r


STEP_TOL = {"float32": dict(rtol=2e-6, atol=2e-6), "bfloat16": dict(rtol=1e-2, atol=1e-5)}  # values are O(1)


@pytest.mark.parametrize("pdt,gdt", [("float32", "float32"), ("float32", "bfloat16"), ("bfloat16", "bfloat16"),
                                     ("bfloat16", "float32"), ("float32", "float16")])
@pytest.mark.parametrize("clip", [None, 0.3])
def test_step_adam_ranges(lib, pdt, gdt, clip):
    rng = np.random.default_rng(5)
    n = 50000 + 3  # not a multiple of the vector width; last range ends at n
    specs = [
        dict(begin=0, end=8192, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.1, step=3),
        dict(begin=8192, end=8192 + 64, lr=5e-4, betas=(0.8, 0.9), eps=1e-6, weight_decay=0.0, step=1),  # tiny range
        # gap [8256, 16384): parameters without gradients, must stay untouched
        dict(begin=16384, end=40000, lr=2e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01, step=7, decoupled=False),
        dict(begin=40000, end=n, lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, step=2, maximize=True),
    ]
    p0, g0 = rng.standard_normal(n) * 0.5, rng.standard_normal(n) * 0.1
    m0, v0 = rng.standard_normal(n) * 0.01, np.abs(rng.standard_normal(n)) * 1e-3
    p, m, v, g = dev(p0, pdt), dev(m0, pdt), dev(v0, pdt), dev(g0, gdt)
    ph, mh, vh, gh = host(p), host(m), host(v), host(g)
    norm = torch.tensor([float(np.sqrt(np.sum(gh.astype(np.float64) ** 2)))], device="cuda")
    L.calls.dgs_step_adam(ptr(p), ptr(m), ptr(v), CODE[pdt], ptr(g), CODE[gdt], n, adam_ranges(specs), len(specs),
                          ptr(norm) if clip is not None else None, clip or 0.0, stream())
    torch.cuda.synchronize()
    coef = np.float32(1.0)
    if clip is not None:
        coef = min(np.float32(clip) / (np.float32(float(norm)) + np.float32(1e-6)), np.float32(1.0))
    want_p, want_m, want_v = ph.copy(), mh.copy(), vh.copy()
    for s in specs:
        sl = slice(s["begin"], s["end"])
        want_p[sl], want_m[sl], want_v[sl] = O.adam_step(
            ph[sl], (gh[sl] * coef).astype(np.float32), mh[sl], vh[sl], s["step"], lr=s["lr"], betas=s["betas"], eps=s["eps"],
            weight_decay=s["weight_decay"], decoupled=s.get("decoupled", True), maximize=s.get("maximize", False), dtype="float32")
    tol = STEP_TOL[pdt]
    np.testing.assert_allclose(host(p), O.rnd(want_p, pdt), **tol)
    np.testing.assert_allclose(host(m), O.rnd(want_m, pdt), **tol)
    np.testing.assert_allclose(host(v), O.rnd(want_v, pdt), **tol)
    np.testing.assert_array_equal(host(p)[8256:16384], ph[8256:16384])
    np.testing.assert_array_equal(host(m)[8256:16384], mh[8256:16384])


@pytest.mark.parametrize("pdt,gdt", [("float32", "float32"), ("bfloat16", "bfloat16"), ("float32", "bfloat16")])
@pytest.mark.parametrize("first", [True, False])
@pytest.mark.parametrize("nesterov", [True, False])
def test_step_sgd_ranges(lib, pdt, gdt, first, nesterov):
    rng = np.random.default_rng(6)
    n = 30000 + 5
    specs = [
        dict(begin=0, end=10000, lr=0.1, momentum=0.9, weight_decay=5e-4, nesterov=nesterov, first=first),
        dict(begin=12288, end=n, lr=0.05, momentum=0.8, dampening=0.0 if nesterov else 0.1, weight_decay=0.0,
             nesterov=nesterov, first=first, maximize=True),
    ]
    p, b, g = dev(rng.standard_normal(n), pdt), dev(rng.standard_normal(n) * 0.1, pdt), dev(rng.standard_normal(n) * 0.1, gdt)
    ph, bh, gh = host(p), host(b), host(g)
    L.calls.dgs_step_sgd(ptr(p), ptr(b), CODE[pdt], ptr(g), CODE[gdt], n, sgd_ranges(specs), len(specs), None, 0.0, stream())
    torch.cuda.synchronize()
    want_p, want_b = ph.copy(), bh.copy()
    for s in specs:
        sl = slice(s["begin"], s["end"])
        want_p[sl], want_b[sl] = O.sgd_step(ph[sl], gh[sl], None if first else bh[sl], lr=s["lr"], momentum=s["momentum"],
                                            dampening=s.get("dampening", 0.0), weight_decay=s["weight_decay"],
                                            nesterov=nesterov, maximize=s.get("maximize", False))
    tol = STEP_TOL[pdt]
    np.testing.assert_allclose(host(p), O.rnd(want_p, pdt), **tol)
    np.testing.assert_allclose(host(b), O.rnd(want_b, pdt), **tol)
    np.testing.assert_array_equal(host(p)[10000:12288], ph[10000:12288])


@pytest.mark.parametrize("gdt", ["float32", "float16"])
@pytest.mark.parametrize("clip", [None, 0.3])
def test_step_scaled_entries_equal_the_plain_step_on_unscaled_gradients(lib, gdt, clip):
    """dgs_step_*_scaled on gradients that carry a loss scale S, with inv_scale = 1/S and the norm of the SCALED
    buffer, against dgs_step_* on the same gradients without the scale.  S is a power of two (as a GradScaler's
    always is), so scaling commutes with every rounding on the way and the results must be bit-identical."""
    rng = np.random.default_rng(11)
    n, S = 40000 + 3, 1024.0
    g_true = dev(rng.standard_normal(n) * 0.01, gdt)
    g_scaled = (g_true.float() * S).to(TD[gdt])
    assert torch.equal(g_scaled.float() / S, g_true.float())  # nothing overflowed or went subnormal
    inv = torch.tensor([1.0 / S], device="cuda")
    partials = torch.zeros(L.calls.dgs_l2norm_grid(n), device="cuda")
    norms = {}
    for key, g in (("true", g_true), ("scaled", g_scaled)):
        norms[key] = torch.zeros(1, device="cuda")
        L.calls.dgs_l2norm(ptr(g), CODE[gdt], n, ptr(partials), ptr(norms[key]), stream())
    torch.cuda.synchronize()
    assert float(norms["scaled"]) == float(norms["true"]) * S
    nptr = (lambda k: ptr(norms[k])) if clip is not None else (lambda k: None)

    a_specs = [dict(begin=0, end=16384, lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, step=3),
               dict(begin=16384, end=n, lr=2e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, step=5, decoupled=False)]
    s_specs = [dict(begin=0, end=16384, lr=0.1, momentum=0.9, weight_decay=5e-4, nesterov=True),
               dict(begin=16384, end=n, lr=0.05, momentum=0.8, dampening=0.1, weight_decay=0.0)]
    p0, m0, v0 = dev(rng.standard_normal(n) * 0.5), dev(rng.standard_normal(n) * 0.01), dev(np.abs(rng.standard_normal(n)) * 1e-3)
    out = {}
    for key in ("plain", "scaled"):
        p, m, v = p0.clone(), m0.clone(), v0.clone()
        q, b = p0.clone(), m0.clone()
        if key == "plain":
            L.calls.dgs_step_adam(ptr(p), ptr(m), ptr(v), L.F32, ptr(g_true), CODE[gdt], n, adam_ranges(a_specs), 2, nptr("true"), clip or 0.0, stream())
            L.calls.dgs_step_sgd(ptr(q), ptr(b), L.F32, ptr(g_true), CODE[gdt], n, sgd_ranges(s_specs), 2, nptr("true"), clip or 0.0, stream())
        else:
            L.calls.dgs_step_adam_scaled(ptr(p), ptr(m), ptr(v), L.F32, ptr(g_scaled), CODE[gdt], n, adam_ranges(a_specs), 2,
                                         nptr("scaled"), clip or 0.0, ptr(inv), stream())
            L.calls.dgs_step_sgd_scaled(ptr(q), ptr(b), L.F32, ptr(g_scaled), CODE[gdt], n, sgd_ranges(s_specs), 2,
                                        nptr("scaled"), clip or 0.0, ptr(inv), stream())
        torch.cuda.synchronize()
        out[key] = [t.clone() for t in (p, m, v, q, b)]
    for a, b in zip(out["plain"], out["scaled"]):
        assert torch.equal(a, b)
    assert not torch.equal(out["plain"][0], p0) and not torch.equal(out["plain"][3], p0)  # and something did happen
    # inv_scale == NULL is the plain entry
    p, m, v = p0.clone(), m0.clone(), v0.clone()
    L.calls.dgs_step_adam_scaled(ptr(p), ptr(m), ptr(v), L.F32, ptr(g_true), CODE[gdt], n, adam_ranges(a_specs), 2, nptr("true"), clip or 0.0, None, stream())
    torch.cuda.synchronize()
    assert torch.equal(p, out["plain"][0]) and torch.equal(v, out["plain"][2])


def test_step_sgd_without_momentum_needs_no_buffer(lib):
    rng = np.random.default_rng(7)
    n = 9000
    p, g = dev(rng.standard_normal(n)), dev(rng.standard_normal(n))
    ph, gh = host(p), host(g)
    specs = [dict(begin=0, end=n, lr=0.01, momentum=0.0, weight_decay=0.0)]
    L.calls.dgs_step_sgd(ptr(p), None, L.F32, ptr(g), L.F32, n, sgd_ranges(specs), 1, None, 0.0, stream())
    torch.cuda.synchronize()
    want, _ = O.sgd_step(ph, gh, None, lr=0.01)
    np.testing.assert_allclose(host(p), want, rtol=2e-6, atol=1e-9)


def test_argument_errors_are_reported(lib):
    x = torch.zeros(64, device="cuda")
    with pytest.raises(L.DgsError, match="bad dtype"):
        L.calls.dgs_l2norm_partial(ptr(x), 99, 64, ptr(x), stream())
    r = adam_ranges([dict(begin=4, end=64, lr=1e-3, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.0, step=1)])
    with pytest.raises(L.DgsError, match="multiples of 8"):
        L.calls.dgs_step_adam(ptr(x), ptr(x), ptr(x), L.F32, ptr(x), L.F32, 64, r, 1, None, 0.0, stream())
    with pytest.raises(L.DgsError, match="NULL"):
        L.calls.dgs_scale_cast(None, ptr(x), L.F32, 64, 1.0, stream())


def test_metric_kernels(lib):
    rng = np.random.default_rng(8)
    n_ranks, b, m = 4, 37, 5
    vals = rng.standard_normal((n_ranks, b, m))
    valid = rng.random((n_ranks, b, m)) > 0.2
    valid[:, 3, 1] = False  # a batch where no rank reports metric 1
    nb = np.array([b, b, 0, b], dtype=np.int64)  # rank 2 reports nothing and is dropped
    gathered = torch.from_numpy(vals).cuda()
    vmask = torch.from_numpy(valid.astype(np.uint8)).cuda()
    nbd = torch.from_numpy(nb).cuda()
    per_batch = torch.zeros(b, m, dtype=torch.float64, device="cuda")
    pb_valid = torch.zeros(b, m, dtype=torch.uint8, device="cuda")
    avg = torch.zeros(m, dtype=torch.float64, device="cuda")
    count = torch.zeros(m, dtype=torch.int64, device="cuda")
    L.calls.dgs_metric_average(ptr(gathered), ptr(vmask), ptr(nbd), n_ranks, b, m, ptr(per_batch), ptr(pb_valid), ptr(avg),
                               ptr(count), stream())
    torch.cuda.synchronize()
    # oracle: the reference's own data shapes (dict name -> [rank][batch], None for missing)
    names = [f"m{j}" for j in range(m)]
    combined = []
    for r in range(n_ranks):
        if nb[r] == 0:
            combined.append(({}, 0))
            continue
        combined.append(({names[j]: [vals[r, i, j] if valid[r, i, j] else None for i in range(b)] for j in range(m)}, int(nb[r])))
    lists, batches = O.process_combined_metrics_and_batches(combined)
    live = [r for r in range(n_ranks) if nb[r]]
    for j in range(m):
        for i in range(b):
            contributing = [vals[r, i, j] for r in live if valid[r, i, j]]
            if contributing:
                assert pb_valid[i, j] == 1
                np.testing.assert_allclose(float(per_batch[i, j]), np.mean(contributing), rtol=1e-12)
            else:
                assert pb_valid[i, j] == 0
        ok = [float(per_batch[i, j]) for i in range(b) if pb_valid[i, j]]
        assert int(count[j]) == len(ok)  # integer: exact
        np.testing.assert_allclose(float(avg[j]), np.mean(ok), rtol=1e-12)
    # full-validity case against the oracle's average_training_metrics + make_metrics
    vmask.fill_(1)
    nbd.fill_(b)
    L.calls.dgs_metric_average(ptr(gathered), ptr(vmask), ptr(nbd), n_ranks, b, m, ptr(per_batch), ptr(pb_valid), ptr(avg),
                               ptr(count), stream())
    torch.cuda.synchronize()
    series = {names[j]: [[vals[r, i, j] for i in range(b)] for r in range(n_ranks)] for j in range(m)}
    want = O.average_training_metrics(series, [b] * n_ranks)
    mm = O.make_metrics(want)["avg_metrics"]
    for j in range(m):
        np.testing.assert_allclose(host_f64(per_batch[:, j]), [w[names[j]] for w in want], rtol=1e-12)
        np.testing.assert_allclose(float(avg[j]), mm[names[j]], rtol=1e-12)
        assert int(count[j]) == b


def host_f64(t):
    return t.detach().cpu().numpy()


def test_metric_average_packed_records_equal_separate_arrays(lib):
    """One allgather of packed per-rank records must reduce exactly like three separate arrays."""
    rng = np.random.default_rng(9)
    n_ranks, b, m = 5, 13, 3
    vals = rng.standard_normal((n_ranks, b, m))
    valid = (rng.random((n_ranks, b, m)) > 0.3).astype(np.uint8)
    nb = np.array([b, 0, b, b, b], dtype=np.int64)
    rec = (8 * b * m + 8 + b * m + 7) // 8 * 8
    records = np.zeros((n_ranks, rec), dtype=np.uint8)
    for r in range(n_ranks):
        records[r, : 8 * b * m] = vals[r].reshape(-1).view(np.uint8)
        records[r, 8 * b * m: 8 * b * m + 8] = nb[r: r + 1].view(np.uint8)
        records[r, 8 * b * m + 8: 9 * b * m + 8] = valid[r].reshape(-1)
    outs = []
    for packed in (False, True):
        per_batch = torch.zeros(b, m, dtype=torch.float64, device="cuda")
        pb_valid = torch.zeros(b, m, dtype=torch.uint8, device="cuda")
        avg = torch.zeros(m, dtype=torch.float64, device="cuda")
        count = torch.zeros(m, dtype=torch.int64, device="cuda")
        if packed:
            recs = torch.from_numpy(records).cuda()
            L.calls.dgs_metric_average_packed(ptr(recs), rec, n_ranks, b, m, ptr(per_batch), ptr(pb_valid), ptr(avg), ptr(count), stream())
        else:
            g, v, n = torch.from_numpy(vals).cuda(), torch.from_numpy(valid).cuda(), torch.from_numpy(nb).cuda()
            L.calls.dgs_metric_average(ptr(g), ptr(v), ptr(n), n_ranks, b, m, ptr(per_batch), ptr(pb_valid), ptr(avg), ptr(count), stream())
        torch.cuda.synchronize()
        outs.append([t.cpu().numpy() for t in (per_batch, pb_valid, avg, count)])
    for a, c in zip(*outs):
        np.testing.assert_array_equal(a, c)
    with pytest.raises(L.DgsError, match="record_bytes"):
        L.calls.dgs_metric_average_packed(ptr(recs), 16, n_ranks, b, m, ptr(per_batch), ptr(pb_valid), ptr(avg), ptr(count), stream())


def test_metric_reduce_and_pack(lib):
    metrics = np.array([0.25, 0.5, 0.75, 1, 25.5, 1.9])  # reference test_reducer.py:18-29
    weights = np.array([1, 2, 5, 4, 5, 6], dtype=np.int64)
    vals = torch.from_numpy(np.stack([metrics, -metrics], axis=1).copy()).cuda()  # [n=6, m=2]
    w = torch.from_numpy(weights).cuda()
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    for op, oop, ww in [(L.REDUCE_AVG, O.AVG, None), (L.REDUCE_AVG, O.AVG, w), (L.REDUCE_SUM, O.SUM, None),
                        (L.REDUCE_MAX, O.MAX, None), (L.REDUCE_MIN, O.MIN, None)]:
        L.calls.dgs_metric_reduce(ptr(vals), ptr(ww) if ww is not None else None, 6, 2, op, ptr(out), stream())
        torch.cuda.synchronize()
        for j, col in enumerate([metrics, -metrics]):
            want = O.simple_reduce_metrics(oop, col, list(weights) if ww is not None else None)
            np.testing.assert_allclose(float(out[j]), want, rtol=1e-12)
    L.calls.dgs_metric_reduce(ptr(vals), ptr(w), 6, 2, L.REDUCE_AVG, ptr(out), stream())
    torch.cuda.synchronize()
    assert np.around(float(out[0]), decimals=2) == 6.43
    # pack: device scalars of mixed dtypes -> one row of the metric ring, None -> invalid
    scal = [torch.tensor(1.5, device="cuda"), torch.tensor(2.25, device="cuda", dtype=torch.bfloat16),
            torch.tensor(7, device="cuda", dtype=torch.int64), None, torch.tensor(-3.0, device="cuda", dtype=torch.float64)]
    refs = (L.ScalarRef * len(scal))()
    for j, t in enumerate(scal):
        refs[j].ptr = t.data_ptr() if t is not None else None
        refs[j].dtype = L.dtype_code(t.dtype) if t is not None else L.F32
    row = torch.zeros(len(scal), dtype=torch.float64, device="cuda")
    vrow = torch.zeros(len(scal), dtype=torch.uint8, device="cuda")
    L.calls.dgs_metric_pack(refs, len(scal), ptr(row), ptr(vrow), stream())
    torch.cuda.synchronize()
    assert row.tolist() == [1.5, 2.25, 7.0, 0.0, -3.0]
    assert vrow.tolist() == [1, 1, 1, 0, 1]


def test_step_kernel_timer(lib):
    n = 1 << 20
    p, b, g = (torch.zeros(n, device="cuda") for _ in range(3))
    specs = [dict(begin=0, end=n, lr=0.1, momentum=0.9, weight_decay=0.0)]
    out = (ctypes.c_float * 16)()
    cnt = ctypes.c_int(-1)
    L.calls.dgs_timer_enable(0)
    L.calls.dgs_timer_read(out, 16, ctypes.byref(cnt))  # clear
    L.calls.dgs_step_sgd(ptr(p), ptr(b), L.F32, ptr(g), L.F32, n, sgd_ranges(specs), 1, None, 0.0, stream())
    L.calls.dgs_timer_read(out, 16, ctypes.byref(cnt))
    assert cnt.value == 0  # nothing is recorded while timing is off
    L.calls.dgs_timer_enable(1)
    for _ in range(3):
        L.calls.dgs_step_sgd(ptr(p), ptr(b), L.F32, ptr(g), L.F32, n, sgd_ranges(specs), 1, None, 0.0, stream())
    L.calls.dgs_timer_read(out, 16, ctypes.byref(cnt))
    L.calls.dgs_timer_enable(0)
    assert cnt.value == 3
    # 20 MB of traffic: a few microseconds, certainly under a millisecond
    assert all(1.0 < out[i] < 1000.0 for i in range(3)), list(out)[:3]
    L.calls.dgs_timer_read(out, 16, ctypes.byref(cnt))
    assert cnt.value == 0  # reading clears the log


# ---- single-launch step (dgs_direct_step_*): one kernel == flatten -> norm -> step of the entry points above ----------
def _direct_case(dt, sizes, misalign, read_acc, seed):
    rng = np.random.default_rng(seed)
    offs, total = layout(sizes)
    holders = [dev(rng.standard_normal(n + misalign) * 0.1, dt) for n in sizes]
    grads = [h[misalign:] for h in holders]  # misalign=1: sources that are not 16-byte aligned
    acc = dev(rng.standard_normal(total) * 0.1) if read_acc else None
    p0 = dev(rng.standard_normal(total) * 0.5, dt)
    s1 = dev(rng.standard_normal(total) * 0.01, dt)
    s2 = dev(np.abs(rng.standard_normal(total)) * 1e-3, dt)
    seg_range = (ctypes.c_int32 * len(sizes))(*[i % 3 for i in range(len(sizes))])
    return offs, total, grads, acc, p0, s1, s2, seg_range


@pytest.mark.parametrize("dt", ["float32", "bfloat16"])
@pytest.mark.parametrize("read_acc", [False, True])
@pytest.mark.parametrize("clip", [None, 0.05])
@pytest.mark.parametrize("misalign", [0, 1])
def test_direct_step_adam_equals_flatten_norm_step(lib, dt, read_acc, clip, misalign):
    sizes = SIZES + [300000]
    offs, total, grads, acc, p0, m0, v0, seg_range = _direct_case(dt, sizes, misalign, read_acc, 11)
    scale = 0.25
    hp = [dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.1, step=3),
          dict(lr=5e-4, betas=(0.8, 0.9), eps=1e-6, weight_decay=0.0, step=1),
          dict(lr=2e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01, step=7, decoupled=False, maximize=True)]
    segs = make_segs(grads, offs)
    # the three-kernel path on copies
    p, m, v = p0.clone(), m0.clone(), v0.clone()
    wire = torch.zeros(total, dtype=TD[dt], device="cuda")
    mode = L.WRITE_WIRE | (L.READ_ACC if read_acc else 0)
    L.calls.dgs_flatten_accumulate(segs, len(sizes), CODE[dt], ptr(acc) if read_acc else None, ptr(wire), CODE[dt], mode, scale, stream())
    partials = torch.zeros(max(L.calls.dgs_l2norm_grid(total), 2048), device="cuda")
    norm = torch.zeros(1, device="cuda")
    L.calls.dgs_l2norm(ptr(wire), CODE[dt], total, ptr(partials), ptr(norm), stream())
    specs = [dict(hp[i % 3], begin=o, end=align_up(o + n)) for i, (o, n) in enumerate(zip(offs, sizes)) if n]
    for k in range(0, len(specs), L.MAX_RANGES):
        chunk = specs[k:k + L.MAX_RANGES]
        L.calls.dgs_step_adam(ptr(p), ptr(m), ptr(v), CODE[dt], ptr(wire), CODE[dt], total, adam_ranges(chunk), len(chunk),
                              ptr(norm) if clip else None, clip or 0.0, stream())
    # one launch
    q, qm, qv = p0.clone(), m0.clone(), v0.clone()
    out_norm = torch.full((1,), -1.0, device="cuda")
    rows = adam_ranges([dict(h, begin=0, end=0) for h in hp])
    L.calls.dgs_direct_step_adam(segs, seg_range, len(sizes), CODE[dt], ptr(acc) if read_acc else None, scale, ptr(q), ptr(qm), ptr(qv),
                                 CODE[dt], rows, 3, 1 if clip else 0, clip or 0.0, ptr(partials), partials.numel(), ptr(out_norm), stream())
    torch.cuda.synchronize()
    tol = STEP_TOL[dt]
    if clip:
        np.testing.assert_allclose(float(out_norm), float(norm), rtol=2e-6)
    else:
        assert float(out_norm) == -1.0  # no clipping asked: the norm is not computed
    # (the ranged step also walks the alignment gaps behind each tensor, where this test put random values and the
    # engine keeps zeros; the single launch never touches them)
    covered = np.zeros(total, dtype=bool)
    for o, n in zip(offs, sizes):
        covered[o:o + n] = True
    for a, b, what in ((q, p, "param"), (qm, m, "exp_avg"), (qv, v, "exp_avg_sq")):
        np.testing.assert_allclose(host(a)[covered], host(b)[covered], err_msg=what, **tol)
    np.testing.assert_array_equal(host(q)[~covered], host(p0)[~covered])
    np.testing.assert_array_equal(host(qv)[~covered], host(v0)[~covered])


@pytest.mark.parametrize("dt", ["float32", "bfloat16"])
@pytest.mark.parametrize("variant", ["momentum", "first", "nesterov", "plain"])
@pytest.mark.parametrize("clip", [None, 0.05])
def test_direct_step_sgd_equals_flatten_norm_step(lib, dt, variant, clip):
    sizes = SIZES + [200000]
    offs, total, grads, acc, p0, b0, _, seg_range = _direct_case(dt, sizes, 0, variant == "first", 12)
    read_acc = acc is not None
    scale = 0.5
    mom = 0.0 if variant == "plain" else 0.9
    hp = [dict(lr=0.1, momentum=mom, weight_decay=5e-4, nesterov=variant == "nesterov", first=variant == "first"),
          dict(lr=0.05, momentum=mom and 0.8, dampening=0.1 if variant == "momentum" else 0.0, weight_decay=0.0,
               nesterov=variant == "nesterov", first=variant == "first", maximize=True),
          dict(lr=0.01, momentum=mom, weight_decay=1e-2, nesterov=variant == "nesterov", first=variant == "first")]
    segs = make_segs(grads, offs)
    p, b = p0.clone(), b0.clone()
    wire = torch.zeros(total, dtype=TD[dt], device="cuda")
    L.calls.dgs_flatten_accumulate(segs, len(sizes), CODE[dt], ptr(acc) if read_acc else None, ptr(wire), CODE[dt],
                                   L.WRITE_WIRE | (L.READ_ACC if read_acc else 0), scale, stream())
    partials = torch.zeros(max(L.calls.dgs_l2norm_grid(total), 2048), device="cuda")
    norm = torch.zeros(1, device="cuda")
    L.calls.dgs_l2norm(ptr(wire), CODE[dt], total, ptr(partials), ptr(norm), stream())
    specs = [dict(hp[i % 3], begin=o, end=align_up(o + n)) for i, (o, n) in enumerate(zip(offs, sizes)) if n]
    L.calls.dgs_step_sgd(ptr(p), ptr(b) if mom else None, CODE[dt], ptr(wire), CODE[dt], total, sgd_ranges(specs), len(specs),
                         ptr(norm) if clip else None, clip or 0.0, stream())
    q, qb = p0.clone(), b0.clone()
    out_norm = torch.zeros(1, device="cuda")
    L.calls.dgs_direct_step_sgd(segs, seg_range, len(sizes), CODE[dt], ptr(acc) if read_acc else None, scale, ptr(q),
                                ptr(qb) if mom else None, CODE[dt], sgd_ranges([dict(h, begin=0, end=0) for h in hp]), 3,
                                1 if clip else 0, clip or 0.0, ptr(partials), partials.numel(), ptr(out_norm), stream())
    torch.cuda.synchronize()
    tol = STEP_TOL[dt]
    covered = np.zeros(total, dtype=bool)
    for o, n in zip(offs, sizes):
        covered[o:o + n] = True
    np.testing.assert_allclose(host(q)[covered], host(p)[covered], err_msg="param", **tol)
    np.testing.assert_allclose(host(qb)[covered], host(b)[covered], err_msg="momentum_buffer", **tol)
    np.testing.assert_array_equal(host(q)[~covered], host(p0)[~covered])
    if clip:
        np.testing.assert_allclose(float(out_norm), float(norm), rtol=2e-6)


def test_direct_step_large_set_reloads_gradients_and_small_grid_of_partials(lib):
    """More tiles than resident CTAs (the gradients are read a second time after the grid barrier instead of staying in
    registers), and a partials scratch smaller than the grid the GPU could hold (the grid shrinks to fit it)."""
    n = (1 << 24) + 40  # 64 MiB of fp32: 8 Ki tiles
    g = torch.randn(n, device="cuda") * 0.01
    p0, m0, v0 = torch.randn(n, device="cuda"), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    segs = make_segs([g], [0])
    seg_range = (ctypes.c_int32 * 1)(0)
    rows = adam_ranges([dict(begin=0, end=0, lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, step=1)])
    want = None
    for cap in (4096, 7):
        q, qm, qv = p0.clone(), m0.clone(), v0.clone()
        partials = torch.zeros(cap, device="cuda")
        out_norm = torch.zeros(1, device="cuda")
        L.calls.dgs_direct_step_adam(segs, seg_range, 1, L.F32, None, 1.0, ptr(q), ptr(qm), ptr(qv), L.F32, rows, 1, 1, 1.0,
                                     ptr(partials), cap, ptr(out_norm), stream())
        torch.cuda.synchronize()
        np.testing.assert_allclose(float(out_norm), float(g.double().norm()), rtol=2e-6)
        if want is None:
            want = (q, qm, qv)
        else:
            for a, b in zip((q, qm, qv), want):
                np.testing.assert_allclose(host(a), host(b), rtol=2e-6, atol=1e-7)
    coef = min(1.0 / (float(g.double().norm()) + 1e-6), 1.0)
    np.testing.assert_allclose(host(want[1]), host(g) * np.float32(coef) * np.float32(0.1), rtol=3e-6, atol=1e-9)


def test_adam_step_with_tiny_second_moment(lib):
    """exp_avg_sq near 1e-16 with eps 1e-8: the denominator is eps-dominated and sqrt(v) sits near 1e-8, where an
    approximate square root / division (the kernels use sqrt.approx and a fast division) could part from torch's
    IEEE sequence.  Against the float64 formula at the tolerance of the other step tests, relative to the update."""
    rng = np.random.default_rng(21)
    n = 8192
    p0 = rng.standard_normal(n).astype(np.float32) * 0.02
    g0 = (rng.standard_normal(n) * 1e-8).astype(np.float32)
    m0 = (rng.standard_normal(n) * 1e-8).astype(np.float32)
    v0 = (np.abs(rng.standard_normal(n)) * 1e-16).astype(np.float32)
    v0[:64] = 1e-38  # about the smallest normal fp32
    v0[64:128] = 1e-41  # subnormal
    v0[128:192] = 0.0
    p, m, v, g = dev(p0), dev(m0), dev(v0), dev(g0)
    spec = dict(begin=0, end=n, lr=3e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.0, step=5)
    L.calls.dgs_step_adam(ptr(p), ptr(m), ptr(v), L.F32, ptr(g), L.F32, n, adam_ranges([spec]), 1, None, 0.0, stream())
    torch.cuda.synchronize()
    b1, b2, t = 0.9, 0.95, 5
    m64 = m0.astype(np.float64) + (1 - b1) * (g0.astype(np.float64) - m0)
    v64 = v0.astype(np.float64) * b2 + (1 - b2) * g0.astype(np.float64) ** 2
    upd = (3e-4 / (1 - b1 ** t)) * m64 / (np.sqrt(v64) / np.sqrt(1 - b2 ** t) + 1e-8)
    got_upd = p0.astype(np.float64) - host(p).astype(np.float64)
    # the update itself (not p, whose rounding would hide it): within 1e-5 of its own magnitude, elementwise
    np.testing.assert_allclose(got_upd, upd, rtol=1e-5, atol=8e-9)  # atol: one ulp of the largest p (|p| < 0.125)
    np.testing.assert_allclose(host(v), v64, rtol=2e-6, atol=1e-45)
