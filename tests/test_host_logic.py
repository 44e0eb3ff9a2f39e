"""CPU tests of the host-side mirror: C-ABI surface, DistributedContext, reducers, metric utilities and
the context's API contract.  Modelled on the reference's own tests (harness/tests/test_ipc.py:117-188,
harness/tests/experiment/pytorch/test_reducer.py, test_metric_utils.py, test_pytorch_context.py:38-52)."""
import ctypes
import pathlib
import re
import socket
import sys
import threading
import traceback
from typing import Any, Callable, List

import numpy as np
import pytest
import torch

import determined_clone_b200 as det
from determined_clone_b200 import _lib, core, pytorch
from determined_clone_b200.pytorch import _metric_utils as metric_utils

ROOT = pathlib.Path(__file__).resolve().parent.parent


# ---- C ABI ------------------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol(lib):
    header = (ROOT / "include" / "dgs.h").read_text()
    declared = set(re.findall(r"\b(dgs_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found in include/dgs.h"
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} is declared in include/dgs.h but not exported by libdgs.so"
    assert declared == set(_lib.SIGNATURES), "ctypes signatures and include/dgs.h disagree"
    assert lib.dgs_version() == 100
    # struct layouts the header promises
    assert ctypes.sizeof(_lib.Seg) == 24 and ctypes.sizeof(_lib.AdamRange) == 80 and ctypes.sizeof(_lib.SgdRange) == 56


def test_host_side_queries_work_without_a_gpu(lib):
    assert lib.dgs_l2norm_grid(0) == 1
    assert lib.dgs_l2norm_grid(1 << 40) == _lib.NORM_MAX_PARTIALS
    assert lib.dgs_launch_count() >= 0
    # argument validation happens before any CUDA call
    with pytest.raises(_lib.DgsError, match="n_ranges"):
        _lib.calls.dgs_step_adam(None, None, None, 0, None, 0, 64, None, 99, None, 0.0, None)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "libdgs.so")
    with pytest.raises(_lib.DgsError, match="no CPU or torch fallback"):
        _lib.load()


# ---- threads as ranks ---------------------------------------------------------------------------
def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_ranks(cross_size: int, local_size: int, fn: Callable[[core.DistributedContext], Any], force_tcp: bool = False) -> List[Any]:
    """One thread per rank, each with a real DistributedContext over localhost ZeroMQ
    (the reference's harness/tests/parallel.py:15-153 pattern)."""
    size = cross_size * local_size
    ports = (_free_port(), _free_port())
    results: List[Any] = [None] * size
    errors: List[str] = []

    def body(rank: int) -> None:
        ctx = None
        try:
            ctx = core.DistributedContext(
                rank=rank, size=size, local_rank=rank % local_size, local_size=local_size,
                cross_rank=rank // local_size, cross_size=cross_size, chief_ip="127.0.0.1",
                pub_port=ports[0], pull_port=ports[1], force_tcp=force_tcp,
            )
            results[rank] = fn(ctx)
        except Exception:
            errors.append(traceback.format_exc())
        finally:
            if ctx is not None:
                ctx.close()

    threads = [threading.Thread(target=body, args=(r,), daemon=True) for r in range(size)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=60)
    assert not errors, "\n".join(errors)
    assert all(not t.is_alive() for t in threads), "a rank hung"
    return results


@pytest.mark.parametrize("cross_size", [1, 3])
@pytest.mark.parametrize("local_size", [1, 3])
@pytest.mark.parametrize("force_tcp", [False, True])
def test_distributed_context_collectives(cross_size, local_size, force_tcp):
    size = cross_size * local_size

    def fn(d: core.DistributedContext):
        out = {}
        out["bcast"] = d.broadcast("hello" if d.rank == 0 else None)
        out["gather"] = d.gather(d.rank * 10)
        out["allgather"] = d.allgather(("r", d.rank))
        out["bcast_local"] = d.broadcast_local(d.cross_rank if d.local_rank == 0 else None)
        out["gather_local"] = d.gather_local(d.rank)
        out["allgather_local"] = d.allgather_local(d.local_rank)
        out["queries"] = (d.get_rank(), d.get_local_rank(), d.get_size(), d.get_num_agents())
        return out

    res = run_ranks(cross_size, local_size, fn, force_tcp)
    for rank, out in enumerate(res):
        assert out["bcast"] == "hello"
        assert out["gather"] == ([r * 10 for r in range(size)] if rank == 0 else None)
        assert out["allgather"] == [("r", r) for r in range(size)]  # ordered by rank
        assert out["bcast_local"] == rank // local_size
        base = rank // local_size * local_size
        assert out["gather_local"] == (list(range(base, base + local_size)) if rank % local_size == 0 else None)
        assert out["allgather_local"] == list(range(local_size))
        assert out["queries"] == (rank, rank % local_size, size, cross_size)


def test_distributed_context_argument_checks():
    with pytest.raises(ValueError, match="must all be provided"):
        core.DistributedContext(rank=0, size=1, local_rank=None, local_size=1, cross_rank=0, cross_size=1)
    with pytest.raises(AssertionError, match="chief_ip"):
        core.DistributedContext(rank=0, size=4, local_rank=0, local_size=2, cross_rank=0, cross_size=2)
    d = core.DummyDistributedContext()
    assert d.gather(3) == [3] and d.allgather(3) == [3] and d.broadcast(3) == 3
    assert d.gather_local(3) == [3] and d.allgather_local(3) == [3] and d.broadcast_local(3) == 3
    d.close()


def test_out_of_order_messages_are_detected():
    from determined_clone_b200 import ipc

    srv = ipc.ZMQBroadcastServer(1)
    cli = ipc.ZMQBroadcastClient(f"tcp://127.0.0.1:{srv.get_pub_port()}", f"tcp://127.0.0.1:{srv.get_pull_port()}")
    t = threading.Thread(target=cli.safe_start)
    t.start()
    srv.safe_start()
    t.join()
    cli.send("a")
    assert srv.gather() == ["a"]
    cli._sent += 1  # skip a serial number, as a rank that missed a collective would
    cli.send("b")
    with pytest.raises(RuntimeError, match="Out-of-order client message"):
        srv.gather()
    srv.broadcast("x")
    cli._received = 5
    with pytest.raises(RuntimeError, match="Out-of-order server message"):
        cli.recv()
    srv.close()
    cli.close()


# ---- reducers -----------------------------------------------------------------------------------
@pytest.mark.parametrize("cross_size", [1, 3])
@pytest.mark.parametrize("local_size", [1, 3])
def test_custom_reducer_slot_order(cross_size, local_size):
    """47 values dealt round-robin in batches of 2 come back in the original order, sum 1081
    (reference test_reducer.py:43-138)."""
    size = cross_size * local_size
    dataset, batch = list(range(47)), 2
    batches = [dataset[i:i + batch] for i in range(0, len(dataset), batch)]

    def reducer_fn(values):
        flat = [v for sub in values for v in sub]
        return {"values": flat, "sum": sum(flat)}

    def fn(d: core.DistributedContext):
        rc = pytorch._PyTorchReducerContext(d.allgather)
        wrapped = rc.wrap_reducer(reducer_fn, for_training=False)
        for i, b in enumerate(batches):
            if i % size == d.rank:
                wrapped.update(b)
        assert rc.reduce_metrics(for_training=True) == {}
        return rc.reduce_metrics(for_training=False)

    for out in run_ranks(cross_size, local_size, fn):
        assert out["sum"] == 1081
        assert out["values"] == dataset


def test_reducer_contract_errors():
    rc = pytorch._PyTorchReducerContext()

    def f(values):
        return 1

    w = rc.wrap_reducer(f, name=None)
    with pytest.raises(AssertionError, match="same reducer"):
        rc.wrap_reducer(w)
    with pytest.raises(AssertionError, match="invalid reducer"):
        rc.wrap_reducer(3)  # type: ignore
    w.update(1)
    with pytest.raises(AssertionError, match=r"function f\(\) was wrapped by a call to wrap_reducer\(\) with name=None but it did not return a dict"):
        rc.reduce_metrics(for_training=True)

    rc2 = pytorch._PyTorchReducerContext()

    def g(values):
        return {"a": 1}

    rc2.wrap_reducer(g, name="x")
    with pytest.raises(AssertionError, match="with name set but it returned a dict anyway"):
        rc2.reduce_metrics(for_training=True)

    class Tri(pytorch.MetricReducer):
        def __init__(self):
            self.reset()

        def reset(self):
            self.s = 0

        def update(self, v):
            self.s += v

        def per_slot_reduce(self):
            return self.s

        def cross_slot_reduce(self, per_slot):
            return sum(per_slot)

    rc3 = pytorch._PyTorchReducerContext()
    t = rc3.wrap_reducer(Tri(), name="tri")
    for i in range(10):
        t.update(i)
    assert rc3.reduce_metrics(for_training=True) == {"tri": 45}  # integer: exact
    rc3.reset_reducers()
    assert rc3.reduce_metrics(for_training=True) == {"tri": 0}


def test_simple_reduce_metrics_reference_vectors():
    metrics = np.array([0.25, 0.5, 0.75, 1, 25.5, 1.9])
    R = pytorch.Reducer
    assert np.around(pytorch._simple_reduce_metrics(R.AVG, metrics), decimals=2) == 4.98
    assert pytorch._simple_reduce_metrics(R.SUM, metrics) == 29.9
    assert pytorch._simple_reduce_metrics(R.MIN, metrics) == 0.25
    assert pytorch._simple_reduce_metrics(R.MAX, metrics) == 25.5
    assert np.around(pytorch._simple_reduce_metrics(R.AVG, metrics, [1, 2, 5, 4, 5, 6]), decimals=2) == 6.43
    with pytest.raises(RuntimeError, match="Lengths of metrics and num_batches"):
        pytorch._simple_reduce_metrics(R.AVG, metrics, [1])


# ---- metric utilities ---------------------------------------------------------------------------
def test_metric_utils_reference_vectors():
    assert metric_utils._process_combined_metrics_and_batches([({}, 2), ({}, 2)]) == ({}, [2, 2])
    combined = [({}, 0), ({}, 0), ({"loss1": [1, 2], "loss2": [-1, -2]}, 2), ({"loss1": [3, 4], "loss2": [-3, -4]}, 2)]
    lists, batches = metric_utils._process_combined_metrics_and_batches(combined)
    assert lists == {"loss1": [[1, 2], [3, 4]], "loss2": [[-1, -2], [-3, -4]]} and batches == [2, 2]
    assert metric_utils._average_training_metrics(lists, batches) == [{"loss1": 2, "loss2": -2}, {"loss1": 3, "loss2": -3}]
    arr = {"loss1": [[1, 2], [3, 4]], "loss2": [[np.array(-1), np.array(-2)], [np.array(-3), np.array(-4)]]}
    out = metric_utils._average_training_metrics(arr, [2, 2])
    assert out == [{"loss1": 2, "loss2": np.array(-2)}, {"loss1": 3, "loss2": np.array(-3)}]
    assert isinstance(out[0]["loss2"], np.ndarray)
    R = pytorch.Reducer
    assert metric_utils._prepare_metrics_reducers(R.AVG, ["a", "b"]) == {"a": R.AVG, "b": R.AVG}
    assert metric_utils._prepare_metrics_reducers({"a": R.SUM, "b": R.MAX}, ["a", "b"]) == {"a": R.SUM, "b": R.MAX}
    # reference test_metric_utils.py:88-115: both misuses are InvalidExperimentException
    with pytest.raises(det.errors.InvalidExperimentException, match="reducer for every validation metric"):
        metric_utils._prepare_metrics_reducers({"a": R.SUM}, ["a", "b"])
    with pytest.raises(det.errors.InvalidExperimentException, match="reducer for every validation metric"):
        metric_utils._prepare_metrics_reducers({}, {"loss1": 1, "loss2": 2}.keys())
    with pytest.raises(det.errors.InvalidExperimentException, match="Please select `determined.pytorch.Reducer`"):
        metric_utils._prepare_metrics_reducers({"loss1": R.AVG, "loss2": 2}, {"loss1": 1, "loss2": 2}.keys())
    with pytest.raises(det.errors.InvalidExperimentException, match="Please select `determined.pytorch.Reducer`"):
        metric_utils._prepare_metrics_reducers("avg", ["a"])
    converted = metric_utils._convert_metrics_to_numpy({"loss1": 1, "loss2": torch.tensor(2)})
    assert converted == {"loss1": 1, "loss2": np.array(2)} and isinstance(converted["loss2"], np.ndarray)
    scaler = torch.amp.GradScaler("cpu")
    ctx = pytorch.PyTorchTrialContext()
    assert ctx.wrap_scaler(scaler) is scaler and ctx._scaler is scaler  # reference test_pytorch_context.py:54-57


@pytest.mark.parametrize("size", [2, 3])
def test_training_metric_averaging_across_ranks(size):
    """The object route end to end: every rank contributes per-batch metrics, the chief averages."""
    def fn(d: core.DistributedContext):
        per_batch = [{"loss": float(d.rank + b), "acc": np.array(0.5 * d.rank)} for b in range(4)]
        out = metric_utils._combine_and_average_training_metrics(d, per_batch)
        val = metric_utils._reduce_metrics(
            d, [{"v": np.float64(d.rank + 1)}] * (d.rank + 1), ["v"], {"v": pytorch.Reducer.AVG}
        )
        return out, val

    res = run_ranks(1, size, fn)
    out0, val0 = res[0]
    mean_rank = (size - 1) / 2
    assert [m["loss"] for m in out0] == [mean_rank + b for b in range(4)]
    assert all(isinstance(m["acc"], np.ndarray) and m["acc"] == 0.5 * mean_rank for m in out0)
    # validation: AVG weighted by batches per rank: sum((r+1)^2) / sum(r+1)
    want = sum((r + 1) ** 2 for r in range(size)) / sum(r + 1 for r in range(size))
    assert val0 == {"v": pytest.approx(want)}
    for out, val in res[1:]:
        assert val == {}


# ---- context API contract ---------------------------------------------------------------------------
def test_context_api_contract():
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, aggregation_frequency=2)
    # reference test_pytorch_context.py:38-52
    pytorch.PyTorchTrialContext._average_gradients(None, 1)
    with pytest.raises(det.errors.InternalException, match="Training hasn't started."):
        ctx._should_communicate_and_update()
    ctx._current_batch_idx = 0
    assert ctx._should_communicate_and_update() is False
    ctx._current_batch_idx = 1
    assert ctx._should_communicate_and_update() is True
    model = torch.nn.Linear(2, 2)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    with pytest.raises(det.errors.InvalidExperimentException, match="backward_passes_per_step"):
        ctx.wrap_optimizer(opt, backward_passes_per_step=0)
    with pytest.raises(det.errors.InvalidExperimentException, match="auto_zero_grads must be set to true"):
        ctx.step_optimizer(opt, auto_zero_grads=False)
    ctx._current_batch_idx = 0
    ctx.step_optimizer(opt)  # not a boundary: silent no-op, even for an optimizer that was never wrapped
    if not torch.cuda.is_available():
        # no GPU: the grad-sync path must refuse rather than fall back to torch on the CPU
        with pytest.raises(_lib.DgsError, match="no CPU implementation"):
            ctx.wrap_optimizer(opt)
    unmanaged = pytorch.PyTorchTrialContext(managed_training=False)
    assert unmanaged._should_communicate_and_update() is True
    p = torch.nn.Parameter(torch.ones(3))
    p.grad = torch.full((3,), 4.0)
    pytorch.PyTorchTrialContext._average_gradients([p, torch.nn.Parameter(torch.ones(1))], 4)
    assert p.grad.tolist() == [1.0, 1.0, 1.0]
    assert ctx.get_per_slot_batch_size() == 8 and ctx.get_global_batch_size() == 8
    with pytest.raises(ValueError):
        ctx.get_hparam("nope")


def test_batch_size_split_and_data_sharding():
    assert det.util.calculate_batch_sizes({"global_batch_size": 10}, 4, "PyTorchTrial") == (2, 8)  # floor + warning
    with pytest.raises(det.errors.InvalidExperimentException):
        det.util.calculate_batch_sizes({"global_batch_size": 2}, 4, "PyTorchTrial")
    ds = torch.utils.data.TensorDataset(torch.arange(20))
    for size in (1, 2, 3):
        seen = []
        for rank in range(size):
            dl = pytorch.DataLoader(ds, batch_size=2, shuffle=False).get_data_loader(repeat=False, num_replicas=size, rank=rank)
            mine = [b[0].tolist() for b in dl]
            # rank r's k-th batch is global batch k*N + r (reference samplers.py:113-120)
            assert mine == [[2 * (k * size + rank), 2 * (k * size + rank) + 1] for k in range(len(mine))]
            seen += [x for b in mine for x in b]
        assert sorted(seen) == list(range(20))
    dl = pytorch.DataLoader(ds, batch_size=5).get_data_loader(repeat=True, skip=3, num_replicas=1, rank=0)
    it = iter(dl)
    assert next(it)[0].tolist() == [15, 16, 17, 18, 19] and next(it)[0].tolist() == [0, 1, 2, 3, 4]  # resume + repeat


def test_lr_scheduler_step_modes():
    """Never step on non-boundary batches; STEP_EVERY_BATCH catches up A steps at the boundary
    (reference _pytorch_trial.py:845-870)."""
    SM = pytorch.LRScheduler.StepMode

    class Count:
        def __init__(self):
            self.n = 0

        def step(self):
            self.n += 1

    ctx = pytorch.PyTorchTrialContext(aggregation_frequency=4)
    ctx._epoch_len = 6
    ctl = pytorch._PyTorchTrialController.__new__(pytorch._PyTorchTrialController)
    ctl.context = ctx
    counts = {}
    for mode in (SM.STEP_EVERY_BATCH, SM.STEP_EVERY_OPTIMIZER_STEP, SM.STEP_EVERY_EPOCH, SM.MANUAL_STEP):
        c = Count()
        sched = pytorch.LRScheduler(c, mode)
        for b in range(24):
            ctx._current_batch_idx = b
            ctl._auto_step_lr_scheduler_per_batch(b, sched)
        counts[mode] = c.n
    assert counts == {SM.STEP_EVERY_BATCH: 24, SM.STEP_EVERY_OPTIMIZER_STEP: 6, SM.STEP_EVERY_EPOCH: 4, SM.MANUAL_STEP: 0}
    with pytest.raises(TypeError):
        pytorch.LRScheduler(Count(), "every batch")  # type: ignore


# ---- flat layout and bucket planning (pure host logic of the engine) --------------------------------------------
def test_layout_and_bucket_planning():
    from determined_clone_b200.pytorch import _gradsync as G

    rng = np.random.default_rng(0)
    for trial in range(20):
        numels = [int(x) for x in rng.integers(1, 3_000_000, size=int(rng.integers(1, 60)))]
        offs, total = G.plan_layout(numels)
        assert offs[0] == 0 and total % G.ALIGN_ELEMS == 0
        for o, n, nxt in zip(offs, numels, offs[1:] + [total]):
            assert o % G.ALIGN_ELEMS == 0 and o + n <= nxt < o + n + G.ALIGN_ELEMS  # aligned, no overlap, minimal padding
        for itemsize, first, cap in ((4, 1 << 20, 25 << 20), (2, 1 << 20, 128 << 20), (4, 1, 1)):
            buckets = G.plan_buckets(offs, numels, total, itemsize, first, cap)
            # the buckets tile [0, total) from the end, each a contiguous run of whole tensors
            assert buckets[0][1] == total and buckets[-1][0] == 0
            for (b0, e0, _), (b1, e1, _) in zip(buckets, buckets[1:]):
                assert e1 == b0 and b1 < e1
            seen = [i for _, _, idxs in buckets for i in idxs]
            assert sorted(seen) == list(range(len(numels))) and seen == sorted(seen, reverse=True)
            for k, (b, e, idxs) in enumerate(buckets):
                assert b == offs[min(idxs)] and (e - b) % G.ALIGN_ELEMS == 0  # divisible by 8 * world for world in {2,4,8}
                nbytes = sum(numels[i] for i in idxs) * itemsize
                limit = first if k == 0 else cap
                if k < len(buckets) - 1:
                    assert nbytes >= limit  # closed because it reached its limit ...
                    assert nbytes - numels[min(idxs)] * itemsize < limit  # ... with the tensor that crossed it
    assert G.adaptive_bucket_cap(44 << 20) == 25 << 20       # small model: torch DDP's size
    assert G.adaptive_bucket_cap(710 << 20) == (710 << 20) // 8
    assert G.adaptive_bucket_cap(14 << 30) == 128 << 20      # capped where NCCL's bandwidth has levelled off


def test_step_plan_cuts_ranges_to_this_ranks_slices():
    """Sharded step: the per-group element ranges of one optimizer step are intersected with this rank's 1/N
    slice of every bucket; together the ranks cover every range exactly once, each launch holds at most
    DGS_MAX_RANGES ranges and its window [lo, hi) encloses them."""
    from determined_clone_b200.pytorch import _gradsync as G

    class FakeComm:
        def __init__(self, rank):
            self.rank = rank

    numels = [640 * 64, 64, 128, 1000 * 64, 192, 64, 300 * 64]
    offs, total = G.plan_layout(numels)
    buckets = [G._Bucket(b, e, idxs) for b, e, idxs in G.plan_buckets(offs, numels, total, 4, 4096, 64 * 400 * 4)]
    fs = G._FlatSet(torch.float32, torch.device("cpu"))
    fs.total, fs.buckets = total, buckets
    # ranges: one per "param group run" -- here simply one per tensor, with a key
    items = [(o, G._align(o + n), ("g", i)) for i, (o, n) in enumerate(zip(offs, numels))]
    eng = G.GradSyncEngine.__new__(G.GradSyncEngine)
    eng.sharded, eng.world, eng.comm = False, 1, None
    plan = eng._step_plan(fs, items)
    assert plan == [(0, total, items)]
    for world in (2, 4, 8):
        covered = np.zeros(total, dtype=np.int32)
        for rank in range(world):
            eng.sharded, eng.world, eng.comm = True, world, FakeComm(rank)
            for lo, hi, chunk in eng._step_plan(fs, items):
                assert 0 < len(chunk) <= _lib.MAX_RANGES and lo == chunk[0][0] and hi == chunk[-1][1]
                for b, e, key in chunk:
                    assert lo <= b < e <= hi and b % 8 == 0 and e % 8 == 0
                    src = [it for it in items if it[2] == key][0]
                    assert src[0] <= b and e <= src[1]
                    covered[b:e] += 1
        assert (covered == 1).all()  # every element of every range is stepped by exactly one rank


def test_clip_interception_routes_only_the_exact_pattern():
    """_run_clip: torch's clip_grad_norm_ over exactly this optimizer's gradient views is answered by the engine
    (norm kernels, clip deferred into the step); anything else falls through to torch untouched.  Exercised on
    CPU with a stand-in engine that computes its norm with torch."""
    from torch.nn.utils import clip_grad_norm_ as imported_at_module_level

    from determined_clone_b200.pytorch import _gradsync as G
    from determined_clone_b200.pytorch import _pytorch_context as C

    params = [torch.nn.Parameter(torch.zeros(5)), torch.nn.Parameter(torch.zeros(3, 2))]
    wire = torch.arange(11, dtype=torch.float32) * 0.5

    class FakeSet:
        tensors = [G._Tensor(params[0], 0, 0, 0), G._Tensor(params[1], 0, 5, 1)]
        grad_views = [wire[0:5], wire[5:11].view(3, 2)]
        wire_dtype = dtype = torch.float32

    class FakeEngine:
        sets = [FakeSet()]
        calls = 0

        def total_norm(self):
            self.calls += 1
            return torch.linalg.vector_norm(wire)

    eng = FakeEngine()
    for p, v in zip(params, FakeSet.grad_views):
        p.grad = v
    want = float(torch.linalg.vector_norm(wire))
    seen = {}

    # (1) the exact pattern, through a name bound at import time: intercepted, clip deferred, gradients untouched
    req = C._run_clip(eng, lambda ps: seen.__setitem__("n", imported_at_module_level(ps, 0.25)), params, defer=True)
    assert req.max_norm == 0.25 and eng.calls == 1 and float(seen["n"]) == pytest.approx(want)
    assert torch.equal(wire, torch.arange(11, dtype=torch.float32) * 0.5)

    # (2) a subset of the parameters, another norm type, or error_if_nonfinite: torch's own code runs
    before = wire.clone()
    req = C._run_clip(eng, lambda ps: torch.nn.utils.clip_grad_norm_(ps[:1], 0.1), params, defer=True)
    assert req.max_norm is None and eng.calls == 1
    assert float(torch.linalg.vector_norm(wire[:5])) == pytest.approx(0.1, rel=1e-4) and torch.equal(wire[5:], before[5:])
    req = C._run_clip(eng, lambda ps: torch.nn.utils.clip_grad_norm_(ps, 1.0, norm_type=float("inf")), params, defer=True)
    assert req.max_norm is None and eng.calls == 1 and float(wire.abs().max()) == pytest.approx(1.0, rel=1e-4)

    # (3) a function that never asks for a norm: nothing requested, the patch is removed afterwards
    from torch.nn.utils import clip_grad as tcg

    real = tcg._get_total_norm
    req = C._run_clip(eng, lambda ps: None, params, defer=True)
    assert req.max_norm is None and tcg._get_total_norm is real
    with pytest.raises(ZeroDivisionError):  # ... also when the user's function raises
        C._run_clip(eng, lambda ps: 1 / 0, params, defer=True)
    assert tcg._get_total_norm is real


def test_controller_reporting_cadence_and_reducers_on_host_values():
    """The train/report loop with host-valued metrics (the object route): a report every reporting_period batches
    and at the end, per-batch metrics preserved, averages per period, custom reducers reset per period, one batch
    in test_mode; delivered one period late but complete and in order."""

    class Trial(pytorch.PyTorchTrial):
        def __init__(self, context):
            self.context = context
            self.total = context.wrap_reducer(lambda vals: sum(vals), name="seen_sum", for_validation=False)

        def build_training_data_loader(self):
            ds = torch.utils.data.TensorDataset(torch.arange(40).float())
            return pytorch.DataLoader(ds, batch_size=5)

        def train_batch(self, batch, epoch_idx, batch_idx):
            self.total.update(batch_idx)
            return {"loss": float(batch_idx), "first": float(batch[0][0]), "maybe": None if batch_idx % 2 else 1.0}

    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 5})
    got = []
    ctl = pytorch._PyTorchTrialController(Trial(ctx), ctx, max_length=8, reporting_period=3, on_report=lambda n, r: got.append((n, r)))
    ctl.run()
    assert [n for n, _ in got] == [3, 6, 8] and ctl.batches_trained == 8
    assert [r["avg_metrics"]["loss"] for _, r in got] == [1.0, 4.0, 6.5]
    assert [r["avg_metrics"]["seen_sum"] for _, r in got] == [3, 12, 13]          # reducers are reset every period
    assert [b["first"] for b in got[0][1]["batch_metrics"]] == [0.0, 5.0, 10.0]    # batches arrive in order
    assert [r["avg_metrics"]["maybe"] for _, r in got] == [1.0, 1.0, 1.0]          # None values are ignored
    assert all(r["steps_completed"] == n and r["samples_per_second"] > 0 for n, r in got)
    assert ctx.is_epoch_start() is False and ctx.current_train_epoch() == 0 and ctx._epoch_len == 8

    once = pytorch._PyTorchTrialController(Trial(pytorch.PyTorchTrialContext(hparams={"global_batch_size": 5})),
                                           ctx, max_length=100, test_mode=True)
    once.run()
    assert once.batches_trained == 1 and len(once.reports) == 1


def _model_only_controller(model, trial_id=1, epoch_len=None, val_from_previous_run=0):
    class Trial(pytorch.PyTorchTrial):
        def __init__(self, context):
            self.context = context
            self.model = context.wrap_model(model)

        def build_training_data_loader(self):
            raise AssertionError("not used")

        def train_batch(self, batch, epoch_idx, batch_idx):
            raise AssertionError("not used")

    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 16})
    ctx._epoch_len = epoch_len
    return pytorch._PyTorchTrialController(Trial(ctx), ctx, max_length=1, trial_id=trial_id, val_from_previous_run=val_from_previous_run)


def test_checkpoint_load_accepts_every_layout_the_reference_accepts(tmp_path):
    """_load (reference _pytorch_trial.py:1117-1313) on host objects only: older file locations, the singular
    pre-multi-model keys and their error texts, DDP's "module." prefix, trial_state.pkl vs the older
    workload_sequencer.pkl, and the trial_id rule."""
    import pickle

    src = torch.nn.Linear(3, 2)
    rng = {"cpu_rng_state": torch.random.get_rng_state(), "np_rng_state": np.random.get_state(), "random_rng_state": __import__("random").getstate()}

    def write(d, payload, where=("state_dict.pth",), **state_files):
        d.joinpath(*where).parent.mkdir(parents=True, exist_ok=True)
        torch.save(payload, d.joinpath(*where))
        for name, state in state_files.items():
            with (d / f"{name}.pkl").open("wb") as f:
                pickle.dump(state, f)
        return d

    def loaded(ctl):
        return {k: v.clone() for k, v in ctl.context.models[0].state_dict().items()}

    def same(ctl):
        return all(torch.equal(v, src.state_dict()[k]) for k, v in loaded(ctl).items())

    # current layout, keys as a DDP-wrapped model writes them, state of the same trial
    ddp_keys = {"module." + k: v for k, v in src.state_dict().items()}
    d = write(tmp_path / "a", {"models_state_dict": [ddp_keys], "optimizers_state_dict": [], "lr_schedulers_state_dict": [], "rng_state": rng},
              trial_state=dict(trial_id=1, last_ckpt=40, step_id=4, last_val=30, batches_trained=40, epochs_trained=2))
    ctl = _model_only_controller(torch.nn.Linear(3, 2), val_from_previous_run=40)
    ctl._load(d)
    assert same(ctl) and ctl.batches_trained == 40
    # the validation made against exactly this checkpoint is known to the master only (:1287-1293)
    assert vars(ctl.state) == dict(trial_id=1, last_ckpt=40, step_id=4, last_val=40, batches_trained=40, epochs_trained=2)

    # oldest layout: pedl/state_dict.pth with singular keys, and the pre-Trainer workload_sequencer.pkl
    d = write(tmp_path / "b", {"model_state_dict": src.state_dict(), "optimizer_state_dict": {}, "rng_state": rng}, where=("pedl", "state_dict.pth"),
              workload_sequencer=dict(trial_id=1, last_ckpt=1, steps_completed=25, step_id=1, last_val=0))
    ctl = _model_only_controller(torch.nn.Linear(3, 2), epoch_len=10)
    ctl._load(d)
    assert same(ctl)
    assert vars(ctl.state) == dict(trial_id=1, last_ckpt=1, step_id=1, last_val=0, batches_trained=25, epochs_trained=2)

    # another trial's checkpoint warm-starts the weights, the counters start over
    for files in (dict(trial_state=dict(trial_id=9, batches_trained=40)), dict(workload_sequencer=dict(trial_id=9, steps_completed=40))):
        d = write(tmp_path / ("c" + next(iter(files))), {"models_state_dict": [src.state_dict()], "optimizers_state_dict": [], "lr_schedulers_state_dict": []}, **files)
        ctl = _model_only_controller(torch.nn.Linear(3, 2), trial_id=3)
        ctl._load(d)
        assert same(ctl) and ctl.batches_trained == 0
        assert vars(ctl.state) == dict(trial_id=3, last_ckpt=0, step_id=0, last_val=0, batches_trained=0, epochs_trained=0)

    # both spellings at once is an error, with the reference's message
    d = write(tmp_path / "d", {"model_state_dict": src.state_dict(), "models_state_dict": [src.state_dict()]})
    with pytest.raises(RuntimeError, match="Both model_state_dict and models_state_dict in checkpoint."):
        _model_only_controller(torch.nn.Linear(3, 2))._load(d)
    # a singular-key checkpoint cannot feed two models
    d = write(tmp_path / "e", {"model_state_dict": src.state_dict(), "optimizer_state_dict": {}})
    ctl = _model_only_controller(torch.nn.Linear(3, 2))
    ctl.context.wrap_model(torch.nn.Linear(3, 2))
    with pytest.raises(RuntimeError, match="Old-format checkpoint cannot be loaded into a context with more than one model."):
        ctl._load(d)
    # no checkpoint file at all, or a file that is not a dict: nothing happens
    ctl = _model_only_controller(torch.nn.Linear(3, 2))
    before = loaded(ctl)
    ctl._load(tmp_path / "nothing-here")
    ctl._load(write(tmp_path / "f", [1, 2, 3], where=("checkpoint.pt",)))
    assert all(torch.equal(v, before[k]) for k, v in loaded(ctl).items()) and ctl.batches_trained == 0


ANCIENT = pathlib.Path("/root/reference/harness/tests/experiment/fixtures/ancient-checkpoints")


@pytest.mark.skipif(not ANCIENT.exists(), reason="the reference tree is only present in the build container")
@pytest.mark.parametrize("ckpt,trial_id,want", [
    # reference test_legacy_checkpoint_loading (test_pytorch_trial.py:1262-1303) asserts exactly these six values
    ("0.20.0-pytorch", 1, dict(trial_id=1, last_ckpt=1, step_id=1, last_val=0, batches_trained=1, epochs_trained=0)),
    ("0.21.0-pytorch-main", 60, dict(trial_id=60, last_ckpt=1, step_id=0, last_val=1, batches_trained=1, epochs_trained=0)),
    ("0.21.0-pytorch", 62, dict(trial_id=62, last_ckpt=1, step_id=0, last_val=1, batches_trained=1, epochs_trained=0)),
])
def test_checkpoints_from_old_reference_releases_load(ckpt, trial_id, want):
    """The reference keeps checkpoints written by its old releases to pin this path; they are read in place (data
    files only: state_dict.pth and the .pkl beside it), the one-weight model they hold is rebuilt here."""
    saved = torch.load(ANCIENT / ckpt / "state_dict.pth", map_location="cpu", weights_only=False)["models_state_dict"][0]
    ctl = _model_only_controller(torch.nn.Linear(1, 1, False), trial_id=trial_id)
    ctl._load(ANCIENT / ckpt)
    assert vars(ctl.state) == want and ctl.batches_trained == 1
    assert torch.equal(ctl.context.models[0].weight, saved["weight"])


def test_autocast_forward_wrapper_keeps_the_module_and_widens_fp16_outputs():
    """experimental.use_amp() before wrap_model (reference _pytorch_context.py:235-282, :307-308): the model the user
    gets back is their own module (attributes, state_dict keys), forward runs under autocast, and fp16 tensors in
    the output -- bare, or inside dicts / lists / tuples / named tuples -- come back as fp32."""
    import collections

    Pair = collections.namedtuple("Pair", ["half", "full"])

    class Out(dict):
        def __init__(self, **kw):
            super().__init__(**kw)

    class M(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = torch.nn.Linear(3, 2)
            self.tag = "mine"
            self.autocast_seen = []

        def forward(self, x, kind):
            self.autocast_seen.append(torch.is_autocast_enabled("cpu"))
            h, f = torch.ones(2, dtype=torch.float16), torch.ones(2, dtype=torch.float64)
            return {"bare": h, "dict": {"h": h, "f": f}, "list": [h, f], "tuple": (h, 3), "named": Pair(h, f), "sub": Out(h=h)}[kind]

    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4})
    ctx.experimental.use_amp()
    assert ctx.experimental._auto_amp and ctx._scaler is not None
    m = M()
    wrapped = ctx.wrap_model(m)
    assert wrapped is m and wrapped.tag == "mine" and list(wrapped.state_dict()) == ["lin.weight", "lin.bias"]
    x = torch.zeros(1, 3)
    assert wrapped(x, "bare").dtype == torch.float32
    d = wrapped(x, "dict")
    assert d["h"].dtype == torch.float32 and d["f"].dtype == torch.float64  # only fp16 is widened
    lst, tup, named, sub = wrapped(x, "list"), wrapped(x, "tuple"), wrapped(x, "named"), wrapped(x, "sub")
    assert isinstance(lst, list) and lst[0].dtype == torch.float32
    assert isinstance(tup, tuple) and tup[0].dtype == torch.float32 and tup[1] == 3
    assert isinstance(named, Pair) and named.half.dtype == torch.float32 and named.full.dtype == torch.float64
    assert isinstance(sub, Out) and sub["h"].dtype == torch.float32
    assert all(m.autocast_seen)
    # a model wrapped BEFORE use_amp() is left alone (the reference checks the flag inside wrap_model)
    ctx2 = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4})
    m2 = ctx2.wrap_model(M())
    ctx2.experimental.use_amp()
    assert m2(x, "bare").dtype == torch.float16 and not any(m2.autocast_seen)


def test_context_host_side_surface(tmp_path, monkeypatch):
    """The host-only members of PyTorchTrialContext that trial code calls (reference _pytorch_context.py:200-229,
    :424-479, :981-1030): argument checks with the reference's texts, stop requests ending the loop after the
    current batch with everything trained so far reported, the torch profiler entered and stepped by the loop."""
    monkeypatch.setenv("DET_TENSORBOARD_PATH", str(tmp_path))
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4})
    assert ctx.get_trial_id() == 0 and ctx.get_experiment_id() == 0  # no master, no launcher
    assert ctx.get_stop_requested() is False and ctx.get_enable_tensorboard_logging() is False
    with pytest.raises(AssertionError, match="stop_requested must be a boolean"):
        ctx.set_stop_requested(1)
    with pytest.raises(AssertionError, match="enable_tensorboard_logging must be a boolean"):
        ctx.set_enable_tensorboard_logging("yes")
    ctx.set_enable_tensorboard_logging(True)
    assert ctx.get_enable_tensorboard_logging() is True
    tb = ctx.get_tensorboard_path()
    assert tb.is_dir() and str(tb).startswith(str(tmp_path)) and ctx.get_tensorboard_path() == tb
    with ctx._record_timing("anything", accumulate=True):
        pass
    ctx.experimental.disable_dataset_reproducibility_checks()
    assert ctx.experimental._data_repro_checks_disabled
    ctx._set_is_pre_trainer()
    assert ctx._is_pre_trainer

    m = ctx.wrap_model(torch.nn.Linear(2, 2))
    other = torch.nn.Linear(2, 2)
    opt = torch.optim.SGD([m.bias, other.weight], lr=0.1)  # never wrapped: only the name filter is exercised
    assert [n for n, _ in ctx._filter_named_parameters(opt)] == ["model_0.bias"]

    class Prof:
        entered = exited = steps = 0

        def __enter__(self):
            Prof.entered += 1
            return self

        def __exit__(self, *a):
            Prof.exited += 1

        def step(self):
            Prof.steps += 1

    class Trial(pytorch.PyTorchTrial):
        def __init__(self, context):
            self.context = context

        def build_training_data_loader(self):
            return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(40).float()), batch_size=4)

        def train_batch(self, batch, epoch_idx, batch_idx):
            if batch_idx == 4:
                self.context.set_stop_requested(True)
            return {"loss": float(batch_idx)}

    ctx.profiler = Prof()
    got = []
    ctl = pytorch._PyTorchTrialController(Trial(ctx), ctx, max_length=100, reporting_period=3, on_report=lambda n, r: got.append((n, r)))
    ctl.run()
    assert ctl.batches_trained == 5 and [n for n, _ in got] == [3, 5]
    assert [r["avg_metrics"]["loss"] for _, r in got] == [1.0, 3.5]  # batches 3 and 4 are reported although the period is 3
    assert (Prof.entered, Prof.exited, Prof.steps) == (1, 1, 5)

    # set_profiler builds a torch.profiler.profile that writes its traces into the TensorBoard directory
    ctx.set_profiler(activities=[torch.profiler.ProfilerActivity.CPU], schedule=torch.profiler.schedule(wait=0, warmup=0, active=1))
    assert isinstance(ctx.profiler, torch.profiler.profile)
    with ctx.profiler as prof:
        torch.ones(4).sum()
        prof.step()
    assert any(f.name.endswith(".pt.trace.json") for f in tb.iterdir())


class _CountingCallback(pytorch.PyTorchCallback):
    """Counts hook calls like the reference's test fixture does (harness/tests/experiment/fixtures/
    pytorch_counter_callback.py), and keeps a picklable state."""

    def __init__(self):
        self.counts = {k: 0 for k in (
            "trial_startups", "validation_steps_started", "validation_steps_ended", "checkpoints_written",
            "checkpoints_uploaded", "training_started_times", "training_epochs_started", "training_epochs_ended",
            "training_workloads_ended", "trial_shutdowns")}
        self.workloads, self.val_outputs, self.loaded_from = [], [], None

    def on_trial_startup(self, first_batch_idx, checkpoint_uuid):
        self.counts["trial_startups"] += 1
        self.started_at = (first_batch_idx, checkpoint_uuid)

    def on_trial_shutdown(self):
        self.counts["trial_shutdowns"] += 1

    def on_training_start(self):
        self.counts["training_started_times"] += 1

    def on_training_epoch_start(self, epoch_idx):
        self.counts["training_epochs_started"] += 1

    def on_training_epoch_end(self, epoch_idx):
        self.counts["training_epochs_ended"] += 1

    def on_training_workload_end(self, avg_metrics, batch_metrics):
        self.counts["training_workloads_ended"] += 1
        self.workloads.append((dict(avg_metrics), len(batch_metrics)))

    def on_validation_start(self):
        self.counts["validation_steps_started"] += 1

    def on_validation_epoch_end(self, outputs):
        self.val_outputs.append(len(outputs))

    def on_validation_end(self, metrics):
        self.counts["validation_steps_ended"] += 1

    def on_checkpoint_write_end(self, checkpoint_dir):
        self.counts["checkpoints_written"] += 1
        assert (pathlib.Path(checkpoint_dir) / "state_dict.pth").exists() and (pathlib.Path(checkpoint_dir) / "trial_state.pkl").exists()

    def on_checkpoint_upload_end(self, uuid):
        self.counts["checkpoints_uploaded"] += 1

    def state_dict(self):
        return {"workloads_seen": self.counts["training_workloads_ended"]}

    def load_state_dict(self, state):
        self.loaded_from = state


class _LegacyEpochCallback(pytorch.PyTorchCallback):
    calls = 0

    def on_training_epoch_start(self):  # the signature from before epoch_idx existed; still accepted
        _LegacyEpochCallback.calls += 1


class _HostTrial(pytorch.PyTorchTrial):
    """No optimizer (so it runs without a GPU): train_batch nudges the one weight by hand."""

    def __init__(self, context):
        self.context = context
        self.model = context.wrap_model(torch.nn.Linear(1, 1, bias=False))
        with torch.no_grad():
            self.model.weight.fill_(0.0)
        self.counter = _CountingCallback()
        self.seen = []

    def build_callbacks(self):
        return {"counter": self.counter, "legacy": _LegacyEpochCallback()}

    def build_training_data_loader(self):
        return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(16).float()), batch_size=2)  # epoch = 8 batches

    def build_validation_data_loader(self):
        return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(6).float()), batch_size=2)

    def train_batch(self, batch, epoch_idx, batch_idx):
        self.seen.append((epoch_idx, batch_idx, float(batch[0][0])))
        with torch.no_grad():
            self.model.weight.add_(1.0)
        return {"loss": float(batch_idx)}

    def evaluate_batch(self, batch):  # no batch_idx parameter: the older signature
        return {"val_loss": float(self.model.weight) + float(batch[0].sum())}


def test_trainer_loop_periods_callbacks_and_resume(tmp_path):
    """Trainer.fit as the reference's local training loop (_pytorch_trial.py:626-820): hook counts for the three
    period settings of the reference's test_callbacks (test_pytorch_trial.py:401-497), checkpoint contents,
    and continuing from a checkpoint at the right batch with the weights, counters and callback state restored."""
    _LegacyEpochCallback.calls = 0
    base = dict(trial_startups=1, training_started_times=1, training_epochs_started=2, training_epochs_ended=2, trial_shutdowns=1)
    cases = [
        # (validation_period, checkpoint_period) -> validations, checkpoints, workloads
        ((None, None), dict(validation_steps_started=1, validation_steps_ended=1, checkpoints_written=1, checkpoints_uploaded=1, training_workloads_ended=1)),
        ((pytorch.Batch(8), None), dict(validation_steps_started=2, validation_steps_ended=2, checkpoints_written=2, checkpoints_uploaded=2, training_workloads_ended=2)),
        ((None, pytorch.Epoch(1)), dict(validation_steps_started=1, validation_steps_ended=1, checkpoints_written=2, checkpoints_uploaded=2, training_workloads_ended=2)),
    ]
    for k, ((val_p, ckpt_p), want) in enumerate(cases):
        ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
        trial = _HostTrial(ctx)
        trainer = pytorch.Trainer(trial, ctx)
        reports = trainer.fit(max_length=pytorch.Epoch(2), validation_period=val_p, checkpoint_period=ckpt_p,
                              reporting_period=pytorch.Batch(10 ** 9), checkpoint_policy="all", checkpoint_dir=str(tmp_path / f"case{k}"))
        assert trial.counter.counts == dict(base, **want), (k, trial.counter.counts)
        assert _LegacyEpochCallback.calls == 2 * (k + 1)
        assert [r["steps_completed"] for r in reports] == ([16] if k == 0 else [8, 16])
        assert len(trainer.checkpoints) == want["checkpoints_written"]
        assert [v["steps_completed"] for v in trainer.validation_reports] == ([16] if k != 1 else [8, 16])
        assert trial.counter.val_outputs == [3] * want["validation_steps_started"]  # 3 validation batches, as host values
        # avg over 3 batches of (weight + x0 + x1): weight = batches trained
        last = trainer.validation_reports[-1]["validation_metrics"]["val_loss"]
        assert last == 16.0 + (1 + 5 + 9) / 3.0
        assert trial.counter.workloads[-1][0]["loss"] == (np.mean(range(8, 16)) if k else np.mean(range(16)))
        assert [s[1] for s in trial.seen] == list(range(16)) and trial.seen[8][0] == 1 and trial.seen[8][2] == 0.0  # epoch 1 restarts the data

    # checkpoint_policy "best" falls back to "all" off-cluster; "none" leaves only the end-of-training checkpoint
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
    trial = _HostTrial(ctx)
    trainer = pytorch.Trainer(trial, ctx)
    trainer.fit(max_length=pytorch.Batch(6), validation_period=pytorch.Batch(2), checkpoint_policy="none", checkpoint_dir=str(tmp_path / "none"))
    assert trial.counter.counts["validation_steps_ended"] == 3 and trial.counter.counts["checkpoints_written"] == 1
    uuid = trainer.checkpoints[-1]
    ckpt = tmp_path / "none" / uuid
    assert sorted(p.name for p in ckpt.iterdir()) == ["load_data.json", "metadata.json", "state_dict.pth", "trial_state.pkl"]
    saved = torch.load(ckpt / "state_dict.pth", weights_only=False)
    assert saved["callbacks"]["counter"] == {"workloads_seen": 3} and float(saved["models_state_dict"][0]["weight"]) == 6.0

    # continue from it: same trial id -> counters continue, the loader resumes at batch 6, callback state comes back
    ctx2 = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
    trial2 = _HostTrial(ctx2)
    trainer2 = pytorch.Trainer(trial2, ctx2)
    reports = trainer2.fit(max_length=pytorch.Batch(10), reporting_period=pytorch.Batch(2), latest_checkpoint=uuid,
                           checkpoint_policy="none", checkpoint_dir=str(tmp_path / "none"))
    assert [s[1] for s in trial2.seen] == [6, 7, 8, 9] and trial2.seen[0][2] == 12.0 and trial2.seen[2][2] == 0.0
    assert float(trial2.model.weight) == 10.0 and trial2.counter.loaded_from == {"workloads_seen": 3}
    assert [r["steps_completed"] for r in reports] == [8, 10]
    assert trial2.counter.counts["training_epochs_started"] == 1 and trial2.counter.counts["training_epochs_ended"] == 1  # epoch 0 ends at batch 8

    # misuse
    with pytest.raises(ValueError, match="max_length must be defined in local training mode"):
        pytorch.Trainer(trial2, ctx2).fit()
    with pytest.raises(ValueError, match="checkpoint_policy"):
        pytorch._PyTorchTrialController(trial2, ctx2, max_length=1, checkpoint_policy="sometimes")

    class BothEval(_HostTrial):
        def evaluate_full_dataset(self, data_loader):
            return {"val_loss": 0.0}

    ctx3 = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
    with pytest.raises(det.errors.InvalidExperimentException, match="exactly one of"):
        pytorch.Trainer(BothEval(ctx3), ctx3).fit(max_length=pytorch.Batch(1), checkpoint_dir=str(tmp_path / "x"))

    class FullEval(pytorch.PyTorchTrial):
        def __init__(self, context):
            self.context = context
            context.wrap_model(torch.nn.Linear(1, 1))

        def build_training_data_loader(self):
            return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(4).float()), batch_size=2)

        def build_validation_data_loader(self):
            return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(4).float()), batch_size=2)

        def train_batch(self, batch, epoch_idx, batch_idx):
            return {"loss": 1.0}

        def evaluate_full_dataset(self, data_loader):
            return {"n": sum(len(b[0]) for b in data_loader)}

    ctx4 = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
    t4 = pytorch.Trainer(FullEval(ctx4), ctx4)
    t4.fit(max_length=pytorch.Batch(1), checkpoint_dir=str(tmp_path / "full"))
    assert t4.validation_reports == [{"steps_completed": 1, "validation_metrics": {"n": 4}}]


def test_load_trial_from_checkpoint_path_round_trip(tmp_path):
    """A checkpoint written by this harness comes back as a trial object with the trained weights, without a GPU
    and without a training context (reference _load.py:68-237)."""
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
    trainer = pytorch.Trainer(_HostTrial(ctx), ctx)
    trainer.fit(max_length=pytorch.Batch(5), checkpoint_policy="none", checkpoint_dir=str(tmp_path))
    ckpt = tmp_path / trainer.checkpoints[-1]
    trial = pytorch.load_trial_from_checkpoint_path(str(ckpt), trial_class=_HostTrial, torch_load_kwargs={"map_location": "cpu"})
    assert isinstance(trial, _HostTrial) and isinstance(trial.context, pytorch.CheckpointLoadContext)
    assert float(trial.model.weight.detach()) == 5.0
    with pytest.warns(FutureWarning, match="torch_load_kwargs"):
        pytorch.load_trial_from_checkpoint_path(str(ckpt), trial_class=_HostTrial, map_location="cpu")
    with pytest.raises(ValueError, match="kwargs may not be set"):
        pytorch.load_trial_from_checkpoint_path(str(ckpt), trial_class=_HostTrial, torch_load_kwargs={"map_location": "cpu"}, map_location="cpu")
    # without trial_class the class named in load_data.json is imported (here: this very test module) ...
    import json

    again = pytorch.load_trial_from_checkpoint_path(str(ckpt))
    assert type(again).__name__ == "_HostTrial" and float(again.model.weight.detach()) == 5.0
    # ... and a name that cannot be imported fails loudly
    info = json.loads((ckpt / "load_data.json").read_text())
    (ckpt / "load_data.json").write_text(json.dumps(dict(info, trial_cls_spec="no_such_module_anywhere:Trial")))
    with pytest.raises(ValueError, match="Automatic import logic failed to import Trial class"):
        pytorch.load_trial_from_checkpoint_path(str(ckpt))
    (ckpt / "load_data.json").unlink()
    (ckpt / "metadata.json").unlink()
    with pytest.raises(AssertionError, match="does not have either load_data.json or metadata.json"):
        pytorch.load_trial_from_checkpoint_path(str(ckpt), trial_class=_HostTrial)
    # a load-only context registers optimizers but cannot train
    lctx = pytorch.CheckpointLoadContext()
    m = lctx.wrap_model(torch.nn.Linear(1, 1))
    opt = lctx.wrap_optimizer(torch.optim.SGD(m.parameters(), lr=0.1))
    assert lctx.optimizers == [opt] and not lctx._engines
    with pytest.raises(det.errors.InvalidExperimentException, match="only loads trials"):
        lctx.step_optimizer(opt)


@pytest.mark.skipif(not ANCIENT.exists(), reason="the reference tree is only present in the build container")
@pytest.mark.parametrize("ckpt,istrial,trial_spec,trial_kwargs", [
    # the cases of the reference's test_trial_checkpoint_loading (test_pytorch_trial.py:1367-1405)
    ("0.13.13-pytorch-old", False, None, {}),
    ("0.13.13-pytorch-flex", True, None, {}),
    ("0.17.6-pytorch", True, None, {}),
    ("0.17.7-pytorch", True, None, {}),
    ("0.20.0-pytorch", True, None, {}),
    ("0.21.0-pytorch", True, None, {"lr": 0.001}),
    ("0.21.0-pytorch-main", True, "train:OneVarPytorchTrial", {"lr": 0.001}),
    ("0.21.0-pytorch", True, "model_def:OneVarPytorchTrial", {"lr": 0.001}),
])
def test_trials_load_from_checkpoints_of_old_reference_releases(monkeypatch, ckpt, istrial, trial_spec, trial_kwargs):
    """The reference's own old checkpoints, read in place, with the user code stored inside them (`code/`, written
    against `import determined`) importing this repo through the compat package."""
    import importlib

    monkeypatch.syspath_prepend(str(pathlib.Path(det.__file__).parent / "compat"))
    for name in [n for n in sys.modules if n == "determined" or n.startswith("determined.")]:
        monkeypatch.delitem(sys.modules, name)
    directory = ANCIENT / ckpt
    trial_class = None
    if trial_spec:
        with det.import_from_path(directory / "code"):
            module, cls = trial_spec.split(":")
            trial_class = getattr(importlib.import_module(module), cls)
    loaded = pytorch.load_trial_from_checkpoint_path(str(directory), trial_class=trial_class, trial_kwargs=trial_kwargs,
                                                      torch_load_kwargs={"map_location": "cpu"})
    saved = torch.load(directory / "state_dict.pth", map_location="cpu", weights_only=False)["models_state_dict"][0]
    if istrial:
        assert isinstance(loaded, pytorch.PyTorchTrial), type(loaded)
        assert torch.equal(loaded.context.models[0].weight, saved["weight"])
    else:
        assert isinstance(loaded, torch.nn.Module) and not isinstance(loaded, pytorch.PyTorchTrial)
        assert torch.equal(loaded.weight, saved["weight"])
    assert "model_def" not in sys.modules and "train" not in sys.modules  # nothing of the checkpoint's code stays imported


def test_samplers_reference_vectors():
    """Known answers of the reference's sampler tests (harness/tests/experiment/pytorch/test_pytorch_data.py:27-135):
    which records / batches each stage lets through, and the two permutations RandomState(777) produces."""
    from torch.utils import data as tud

    S = pytorch.samplers
    seq = tud.SequentialSampler(range(15))
    assert len(S.SkipSampler(seq, 2)) == 15 and list(S.SkipSampler(seq, 2)) == list(range(2, 15))
    batches = tud.BatchSampler(seq, batch_size=2, drop_last=False)
    skipped = S.SkipBatchSampler(batches, 2)
    assert len(skipped) == 8 and list(skipped) == list(batches)[2:]
    assert list(S.SkipBatchSampler(batches, 100)) == []

    ten = tud.SequentialSampler(range(10))
    rep = iter(S.RepeatSampler(ten))
    assert len(S.RepeatSampler(ten)) == 10 and [next(rep) for _ in range(30)] == list(range(10)) * 3
    threes = tud.BatchSampler(ten, 3, False)
    rep = iter(S.RepeatBatchSampler(threes))
    assert len(S.RepeatBatchSampler(threes)) == 4 and [next(rep) for _ in range(12)] == list(threes) * 3

    nineteen = tud.SequentialSampler(range(19))
    want = [[0, 4, 8, 12, 16], [1, 5, 9, 13, 17], [2, 6, 10, 14, 18], [3, 7, 11, 15]]
    for rank in range(4):
        d = S.DistributedSampler(nineteen, 4, rank)
        assert list(d) == want[rank] and len(d) == len(want[rank])
    pairs = tud.BatchSampler(nineteen, batch_size=2, drop_last=False)
    want = [[[0, 1], [8, 9], [16, 17]], [[2, 3], [10, 11], [18]], [[4, 5], [12, 13]], [[6, 7], [14, 15]]]
    for rank in range(4):
        d = S.DistributedBatchSampler(pairs, 4, rank)
        assert list(d) == want[rank] and len(d) == len(want[rank])
    assert list(S.DistributedBatchSampler(pairs, 1, 0)) == list(pairs)
    for bad in ((4, -1, "non-negative"), (0, 0, "greater than zero"), (2, 2, "less than num_workers")):
        with pytest.raises(ValueError, match=bad[2]):
            S.DistributedBatchSampler(pairs, bad[0], bad[1])

    shuffled = S.ReproducibleShuffleSampler(tud.SequentialSampler(range(5)), 777)
    assert list(shuffled) == [0, 4, 1, 2, 3] and list(shuffled) == [2, 0, 1, 3, 4] and len(shuffled) == 5
    fives = tud.BatchSampler(ten, batch_size=2, drop_last=False)
    for cls in (S.ReproducibleShuffleSampler, S.ReproducibleShuffleBatchSampler):
        sb = cls(fives, 777)
        assert list(sb) == [[0, 1], [8, 9], [2, 3], [4, 5], [6, 7]] and list(sb) == [[4, 5], [0, 1], [2, 3], [6, 7], [8, 9]]


def test_data_loader_repeats_shards_and_skips_like_the_reference():
    """determined.pytorch.DataLoader -> get_data_loader (reference _data.py:244-313): endless, resumable, and dealt to
    the ranks along the ENDLESS sequence -- with 5 batches and 2 ranks the ranks swap batches every epoch."""
    from torch.utils import data as tud

    ds = tud.TensorDataset(torch.arange(10).float())
    dl = pytorch.DataLoader(ds, batch_size=2)
    assert len(dl) == 5 and [b[0].tolist() for b in dl] == [[0.0, 1.0], [2.0, 3.0], [4.0, 5.0], [6.0, 7.0], [8.0, 9.0]]

    def first(loader, n):
        it = iter(loader)
        return [int(next(it)[0][0]) // 2 for _ in range(n)]  # batch numbers

    assert first(dl.get_data_loader(repeat=True, skip=2), 9) == [2, 3, 4, 0, 1, 2, 3, 4, 0]
    r0 = dl.get_data_loader(repeat=True, num_replicas=2, rank=0)
    r1 = dl.get_data_loader(repeat=True, num_replicas=2, rank=1)
    assert (len(r0), len(r1)) == (3, 2)
    assert first(r0, 7) == [0, 2, 4, 1, 3, 0, 2] and first(r1, 7) == [1, 3, 0, 2, 4, 1, 3]
    # continuing after 3 batches per rank: exactly the tail of the sequences above
    assert first(dl.get_data_loader(repeat=True, skip=3, num_replicas=2, rank=0), 4) == [1, 3, 0, 2]
    assert first(dl.get_data_loader(repeat=True, skip=3, num_replicas=2, rank=1), 4) == [2, 4, 1, 3]
    # validation: one pass, sharded, ranks may get different counts
    assert [int(b[0][0]) // 2 for b in dl.get_data_loader(num_replicas=2, rank=1)] == [1, 3]
    # every rank shuffles alike under the same seed
    torch.manual_seed(3)
    a = [b[0].tolist() for b in pytorch.DataLoader(ds, batch_size=2, shuffle=True)]
    torch.manual_seed(3)
    b = [b[0].tolist() for b in pytorch.DataLoader(ds, batch_size=2, shuffle=True)]
    assert a == b and sorted(x for bb in a for x in bb) == list(map(float, range(10)))
    assert pytorch.DataLoader(ds, drop_last=True, shuffle=True, batch_size=2).get_data_loader() is not None

    # the argument checks of torch's DataLoader, made at construction
    class Stream(tud.IterableDataset):
        def __iter__(self):
            return iter(range(3))

    for kwargs, text in (
        (dict(dataset=Stream()), "IterableDatasets are not supported"),
        (dict(batch_size=None), "batch_sampler=None and batch_size=None"),
        (dict(num_workers=-1), "num_workers option should be non-negative"),
        (dict(timeout=-1), "timeout option should be non-negative"),
        (dict(prefetch_factor=2), "prefetch_factor option could only be specified in multiprocessing"),
        (dict(persistent_workers=True), "persistent_workers option needs num_workers > 0"),
        (dict(sampler=tud.SequentialSampler(ds), shuffle=True), "sampler option is mutually exclusive with shuffle"),
        (dict(batch_sampler=tud.BatchSampler(tud.SequentialSampler(ds), 2, False), batch_size=2), "batch_sampler option is mutually exclusive"),
    ):
        with pytest.raises(ValueError, match=text):
            pytorch.DataLoader(**dict(dict(dataset=ds), **kwargs))
    custom = pytorch.DataLoader(ds, batch_sampler=tud.BatchSampler(tud.SequentialSampler(ds), 5, False))
    assert [len(b[0]) for b in custom] == [5, 5] and custom.batch_size is None


def test_data_length_and_to_device_reference_cases(caplog):
    """reference test_pytorch_data.py:176-236."""
    n, t = np.array([0]), torch.Tensor(np.array([0]))
    for data in ([n], [t], [[n]], (n,), (t,), {"a": t}, {"a": n}, {"a": [[t]]}):
        assert pytorch.data_length(data) == 1
    assert pytorch.data_length(torch.zeros(7, 3)) == 7
    with pytest.raises(ValueError, match="at least one"):
        pytorch.data_length({})
    with pytest.raises(ValueError, match="list of inputs"):
        pytorch.data_length([])
    with pytest.raises(TypeError, match="Data of incorrect type"):
        pytorch.data_length(0)

    nested = {"input_1": torch.Tensor(1), "input_3": "str", "input_4": 1}
    assert pytorch.to_device(nested, "cpu") == nested
    moved = pytorch.to_device(np.array([0, 1, 2]), "cpu")
    assert isinstance(moved, torch.Tensor) and moved.tolist() == [0, 1, 2]
    assert pytorch.to_device((np.array(["a"]), [torch.ones(1)]), "cpu")[0].dtype.kind == "U"  # strings stay numpy
    import logging

    for keep, want in ((True, 1), (False, 2)):
        caplog.clear()
        seen = set() if keep else None
        with caplog.at_level(logging.WARNING, logger="determined.pytorch"):
            pytorch.to_device(["string_data", "string_data"], "cpu", seen)
            pytorch.to_device(["string_data", "string_data"], "cpu", seen)
        msgs = [r.getMessage() for r in caplog.records]
        assert len(msgs) == want and all("not able to move data" in m for m in msgs)
    assert "not a subclass of det.pytorch.DataLoader" in pytorch._dataset_repro_warning("build_training_data_loader", object())


def test_rng_state_travels_in_the_checkpoint(tmp_path):
    """The reference's test_rng_restore (test_pytorch_trial.py:1305-1365): a trial whose metrics are random numbers
    checkpoints after every batch; a second run restored from the first checkpoint -- started under a different
    seed -- reproduces every later batch of the first run, for all three generators the checkpoint carries."""
    import random

    class NoiseTrial(pytorch.PyTorchTrial):
        def __init__(self, context):
            self.context = context

        def build_training_data_loader(self):
            return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.zeros(8)), batch_size=1, shuffle=True)

        def build_validation_data_loader(self):
            return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.zeros(2)), batch_size=1)

        def train_batch(self, batch, epoch_idx, batch_idx):
            return {"t": float(torch.rand(())), "n": float(np.random.rand()), "p": random.random()}

        def evaluate_batch(self, batch):
            return {"validation_error": float(torch.rand(())) + float(np.random.rand()) + random.random()}

    def run(seed, **fit):
        pytorch._PyTorchTrialController._set_random_seeds(seed)
        ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 1})
        trainer = pytorch.Trainer(NoiseTrial(ctx), ctx)
        reports = trainer.fit(max_length=pytorch.Batch(5), reporting_period=pytorch.Batch(1), validation_period=pytorch.Batch(1),
                              checkpoint_period=pytorch.Batch(1), checkpoint_policy="none", checkpoint_dir=str(tmp_path), **fit)
        train = [(r["steps_completed"], r["batch_metrics"][0]) for r in reports]
        val = [(v["steps_completed"], v["validation_metrics"]["validation_error"]) for v in trainer.validation_reports]
        return train, val, trainer.checkpoints

    train_a, val_a, uuids = run(seed=17)
    assert len(uuids) == 5, "one checkpoint per batch"
    train_b, val_b, _ = run(seed=0, latest_checkpoint=uuids[0])
    assert [n for n, _ in train_b] == [2, 3, 4, 5]
    assert train_b == train_a[1:] and val_b == val_a[1:]
    different, _, _ = run(seed=0)
    assert different != train_a  # the comparison above is not vacuous


@pytest.mark.parametrize("dataset_len", [2, 3])
def test_epoch_bookkeeping_agrees_across_ranks(dataset_len):
    """The reference's test_epoch_sync (test_pytorch_trial.py:1027-1051): global batch 2 over 2 ranks, a dataset
    of 2 or 3 records; every rank reports batch k in epoch k // ceil(len / 2), using the chief's epoch length."""

    def rank_main(dctx):
        seen = []

        class T(pytorch.PyTorchTrial):
            def __init__(self, context):
                self.context = context

            def build_training_data_loader(self):
                ds = torch.utils.data.TensorDataset(torch.arange(dataset_len).float())
                return pytorch.DataLoader(ds, batch_size=self.context.get_per_slot_batch_size())

            def train_batch(self, batch, epoch_idx, batch_idx):
                seen.append((batch_idx, epoch_idx, self.context.current_train_epoch(), self.context.is_epoch_start(), self.context.is_epoch_end()))
                return {"loss": 0.0}

        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 2}, slots_per_trial=2)
        pytorch._PyTorchTrialController(T(ctx), ctx, max_length=10, reporting_period=10).run()
        return seen

    per_epoch = (dataset_len + 1) // 2
    want = [(k, k // per_epoch, k // per_epoch, k % per_epoch == 0, k % per_epoch == per_epoch - 1) for k in range(10)]
    assert run_ranks(1, 2, rank_main) == [want, want]


def test_lr_scheduler_coupling_matches_rates_recorded_from_the_reference():
    """tests/golden/lr_schedule.json holds the learning rates the unmodified reference controller produced for every
    step mode x aggregation_frequency x scheduler frequency (made by tests/golden/make_golden_lr.py).  The same
    StepLR(gamma=0.5) driven by this repo's _auto_step_lr_scheduler_per_batch must show the same rate at every
    batch -- powers of two, compared exactly."""
    import json

    golden = json.loads((pathlib.Path(__file__).parent / "golden" / "lr_schedule.json").read_text())
    assert len(golden["cases"]) == 36
    for case in golden["cases"]:
        agg, freq = case["aggregation_frequency"], case["frequency"]
        ctx = pytorch.PyTorchTrialContext(aggregation_frequency=agg)
        ctx._epoch_len = golden["epoch_len"]
        ctl = pytorch._PyTorchTrialController.__new__(pytorch._PyTorchTrialController)
        ctl.context = ctx
        opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=1.0)
        opt._opt_called = True  # no optimizer.step() happens in this replay; silences torch's ordering warning
        sched = pytorch.LRScheduler(torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5),
                                    getattr(pytorch.LRScheduler.StepMode, case["mode"]), freq)
        got = []
        for b in range(golden["batches"]):
            got.append(opt.param_groups[0]["lr"])
            ctx._current_batch_idx = b
            ctl._auto_step_lr_scheduler_per_batch(b, sched)
        got.append(opt.param_groups[0]["lr"])
        assert got == case["lr"], (case["mode"], agg, freq)


def test_native_host_loop_equals_the_python_loop(monkeypatch):
    """csrc/dgs_host.cpp (used whenever it is built; DGS_HOST_EXT=0 switches it off) against the Python loops it replaces, on CPU tensors: the same
    dgs_seg tables reach dgs_flatten_accumulate, p.grad ends up None, and cases the fast path does not cover (a
    missing gradient, a dtype or layout the kernel cannot read directly) leave everything untouched."""
    import importlib.util

    from determined_clone_b200 import _build
    from determined_clone_b200.pytorch import _gradsync as G

    path = _build.build_host()
    spec = importlib.util.spec_from_file_location("_dgs_host", str(path))
    host = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(host)

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3), torch.nn.BatchNorm2d(4), torch.nn.Flatten(), torch.nn.Linear(4 * 36, 7))
    params = list(model.parameters())
    launches = []

    def fake_flatten(segs, n, gdt, acc, wire, wdt, mode, scale, stream):
        launches.append(([(segs[k].src, segs[k].dst_off, segs[k].numel) for k in range(n)], n, gdt, mode, scale))
        return 0

    monkeypatch.setattr(G, "_stream", lambda: None)
    monkeypatch.setattr(G._lib.calls, "dgs_flatten_accumulate", fake_flatten, raising=False)

    def engine():
        eng = G.GradSyncEngine.__new__(G.GradSyncEngine)
        eng.comm, eng.device, eng.scale, eng.unscale, eng.direct, eng.peer = None, torch.device("cpu"), 0.5, 1.0, False, False
        eng.bucket_bytes, eng.first_bucket_bytes = 0, 64
        eng._final_pass, eng._next_launch, eng._launch_order, eng._eager_norm = True, 0, [], False
        fs = G._FlatSet(torch.float32, torch.device("cpu"))
        offs, fs.total = G.plan_layout([p.numel() for p in params])
        fs.tensors = [G._Tensor(p, 0, off, i) for i, (p, off) in enumerate(zip(params, offs))]
        fs.in_window, fs.in_wire = [False] * len(params), [False] * len(params)
        eng._cut_buckets(fs)
        assert len(fs.buckets) > 1 and [p for b in fs.buckets for p in b.params] == [fs.tensors[ti].param for b in fs.buckets for ti in b.tensors]
        fs.dtype_code, fs.wire_code, fs.acc_ptr, fs.wire_ptr = 0, 0, None, None
        fs.segs_all = (G._lib.Seg * len(params))()
        fs.params = list(params)
        for k, t in enumerate(fs.tensors):
            fs.segs_all[k].dst_off, fs.segs_all[k].numel = t.off, t.numel
        eng.sets = [fs]
        return eng, fs

    def run(native, spoil=None):
        launches.clear()
        monkeypatch.setattr(G, "_host", host if native else None)
        grads = [torch.randn_like(p) for p in params]
        for p, g in zip(params, grads):
            p.grad = g
        if spoil == "missing":
            params[2].grad = None
        elif spoil == "dtype":
            params[2].grad = None
            params[2].data = params[2].data.double()
            params[2].grad = grads[2].double()
        elif spoil == "strided":
            params[0].grad = grads[0].permute(0, 1, 3, 2).contiguous().permute(0, 1, 3, 2)
        eng, fs = engine()
        done = eng._flush_whole_set(fs)
        state = [p.grad is None for p in params]
        if spoil == "dtype":
            params[2].grad = None
            params[2].data = params[2].data.float()
        return done, state, [(l[0], l[1:]) for l in launches], [g.data_ptr() for g in grads]

    for native in (False, True):
        done, cleared, calls, ptrs = run(native)
        assert done and all(cleared) and len(calls) == 1
        table, rest = calls[0]
        assert [t[0] for t in table] == ptrs and rest == (len(params), 0, G._lib.WRITE_WIRE, 0.5)
        for spoil in ("missing", "dtype", "strided"):
            done, cleared, calls, _ = run(native, spoil)
            assert not done and not calls and cleared == [spoil == "missing" and i == 2 for i in range(len(params))], (native, spoil)

    # the per-bucket entry used by the sentinel hooks
    for native in (False, True):
        monkeypatch.setattr(G, "_host", host if native else None)
        launches.clear()
        grads = [torch.randn_like(p) for p in params]
        for p, g in zip(params, grads):
            p.grad = g
        eng, fs = engine()
        eng._reduce_ready = lambda fs=None: None
        eng._sentinels = {id(fs.tensors[b.tensors[-1]].param): (0, bi) for bi, b in enumerate(fs.buckets)}
        for bi, b in enumerate(fs.buckets):
            eng._on_sentinel(fs.tensors[b.tensors[-1]].param)
            assert b.flushed and all(fs.in_wire[ti] for ti in b.tensors)
        assert all(p.grad is None for p in params) and len(launches) == len(fs.buckets)
        got = {src for table, *_ in launches for src, _, _ in table}
        assert got == {g.data_ptr() for g in grads}
    assert host.all_have_grads(params) is False


def test_controller_with_a_core_context_and_train_units(tmp_path):
    """The way the reference's own test scaffolding builds a run (harness/tests/experiment/pytorch_utils.py:16-120):
    core._dummy_init() -> PyTorchTrialContext(core_context=...) -> _PyTorchTrialController with Batch / Epoch units.
    Reports go through core.train, checkpoints through core.checkpoint (a uuid directory with metadata.json), and a
    second run continues from that uuid."""
    from determined_clone_b200 import core as dcore

    cc = dcore._dummy_init(checkpoint_storage=str(tmp_path / "store"), tensorboard_path=tmp_path / "tb")
    ctx = pytorch.PyTorchTrialContext(core_context=cc, trial_seed=3, hparams={"global_batch_size": 2}, slots_per_trial=1, num_gpus=0,
                                      exp_conf={"hyperparameters": {"global_batch_size": 2}}, aggregation_frequency=1,
                                      steps_completed=0, managed_training=True, debug_enabled=False)
    assert ctx.get_trial_id() == 0 and ctx.get_tensorboard_path() == tmp_path / "tb"
    trial = _HostTrial(ctx)
    ctl = pytorch._PyTorchTrialController(
        trial_inst=trial, context=ctx, max_length=pytorch.Epoch(1), checkpoint_period=pytorch.Batch(sys.maxsize),
        validation_period=pytorch.Batch(4), reporting_period=pytorch.Batch(3), searcher_metric_name="val_loss",
        smaller_is_better=True, steps_completed=0, latest_checkpoint=None, local_training=True, test_mode=False,
        checkpoint_policy="best", step_zero_validation=False, det_profiler=None, global_batch_size=None,
        final_validation=True, final_checkpoint=True, trial_id=0,
    )
    ctl.run()
    assert (ctl.max_length, ctl.reporting_period, ctl.validation_period) == (8, 3, 4)
    assert [n for n, _, _ in cc.train.training_reports] == [3, 4, 6, 8] and cc.train.training_reports[0][1]["loss"] == 1.0
    assert len(cc.train.training_reports[1][2]) == 1  # batch 3 alone: the validation at 4 closed the period early
    assert [n for n, _ in cc.train.validation_reports] == [4, 8]
    # val_loss grows with the weight, so only the first validation is "best"; the end of training adds one checkpoint
    assert len(ctl.checkpoints) == 2
    first, last = (tmp_path / "store" / u for u in ctl.checkpoints)
    assert cc.checkpoint.get_metadata(ctl.checkpoints[0])["steps_completed"] == 4 and (first / "state_dict.pth").exists()
    assert cc.checkpoint.get_metadata(ctl.checkpoints[1])["steps_completed"] == 8 and (last / "trial_state.pkl").exists()

    cc2 = dcore._dummy_init(checkpoint_storage=str(tmp_path / "store"))
    ctx2 = pytorch.PyTorchTrialContext(core_context=cc2, hparams={"global_batch_size": 2})
    trial2 = _HostTrial(ctx2)
    ctl2 = pytorch._PyTorchTrialController(trial_inst=trial2, context=ctx2, max_length=pytorch.Batch(6), reporting_period=pytorch.Batch(100),
                                           latest_checkpoint=ctl.checkpoints[0], trial_id=0)
    ctl2.run()
    assert [s[1] for s in trial2.seen] == [4, 5] and float(trial2.model.weight.detach()) == 6.0
    assert [n for n, _, _ in cc2.train.training_reports] == [6]
    with pytest.raises(FileNotFoundError):
        pytorch._PyTorchTrialController(trial_inst=trial2, context=ctx2, max_length=1, latest_checkpoint="no-such-uuid").run()
    with cc2 as entered:
        assert entered is cc2 and cc2.preempt.should_preempt() is False


def test_plain_norm_clip_detector_is_conservative():
    """_is_plain_norm_clip decides whether a clip_grads function can be served without ever showing it a gradient (the
    single-launch step on one GPU, the sliced routes on several): only code that provably does nothing but call
    torch.nn.utils.clip_grad_norm_ qualifies -- the reference's docstring example among it."""
    import functools

    from torch.nn.utils import clip_grad_norm_

    from determined_clone_b200.pytorch._pytorch_context import _is_plain_norm_clip as plain

    class T:
        limit = 1.0

        @staticmethod
        def a(params):
            torch.nn.utils.clip_grad_norm_(params, 1.0)

        def b(self, params):
            clip_grad_norm_(params, 0.5)

        def reads(self, params):
            self.seen = [p.grad for p in params]
            clip_grad_norm_(params, 1.0)

        def attr(self, params):
            clip_grad_norm_(params, self.limit)  # `limit` could be anything: not provable

    def closure(value):
        return lambda params: torch.nn.utils.clip_grad_norm_(params, value)

    yes = [lambda params: torch.nn.utils.clip_grad_norm_(params, 0.0001),  # reference _pytorch_context.py:851-853
           T.a, T().a, T().b, functools.partial(clip_grad_norm_, max_norm=1.0), closure(2.0), closure(3)]
    no = [T().reads, T().attr, closure([1.0]), closure(torch.ones(())), print, functools.partial(print, end=""),
          lambda params: torch.nn.utils.clip_grad_value_(params, 1.0), lambda params: None,
          lambda params: [torch.nn.utils.clip_grad_norm_(p, 1.0) for p in params]]
    assert [plain(f) for f in yes] == [True] * len(yes)
    assert [plain(f) for f in no] == [False] * len(no)
    shadow = {"torch": type("fake", (), {"nn": None})}
    impostor = eval("lambda params: torch.nn.utils.clip_grad_norm_(params, 1.0)", shadow)  # some other `torch`
    assert plain(impostor) is False
