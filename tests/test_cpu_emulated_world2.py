"""Two ranks (and one case each with four and eight), the N>1 Python logic of the engine -- one hook per bucket, buckets flushed and "all-reduced" while
backward runs, the learned launch order, the parameter broadcast from rank 0, norm partials per bucket, the opt-in
sharded step -- in the CPU suite: two processes, torch.distributed over gloo standing in for NCCL through the
emulated communicator entry points of tests/fake_dgs.py (test infrastructure), the control plane being the real
DistributedContext over ZeroMQ.  Compared with the fixtures recorded from the reference on 2-rank torch DDP."""
import os
import pathlib
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
OPT_ORDER_TO_MODEL = [0, 4, 1, 2, 3, 5]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _np(t):
    return t.detach().float().cpu().numpy().copy()


def _worker(rank: int, size: int, port: int, zmq_ports, name: str, sharded, small_buckets: bool, out_dir: str) -> None:
    route = {False: "allreduce", True: "sharded"}.get(sharded, sharded)  # False / True (round 1's spelling) or a route name
    sharded = route != "allreduce"
    for p in (str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    import fake_dgs
    from golden_io import N_PARAMS, Golden, assert_close
    from golden_model import GoldenNet, make_data, make_optimizer

    from determined_clone_b200 import core, pytorch

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    fake = fake_dgs.install()
    if small_buckets:
        # the fixture model has ~1500 parameters: shrink the bucket sizes so that it is cut into several buckets and
        # the per-bucket hooks, the learned launch order and the per-bucket norm slots have something to do
        from determined_clone_b200.pytorch import _gradsync

        class SmallBuckets(_gradsync.GradSyncEngine):
            def __init__(self, *args, **kwargs):
                kwargs.update(bucket_bytes=1024, first_bucket_bytes=256)
                super().__init__(*args, **kwargs)

        real_engine = _gradsync.GradSyncEngine
        _gradsync.GradSyncEngine = SmallBuckets
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        g = Golden(name)
        rec = {"avg": [], "norm": [], "params": [], "state": []}
        reports = []

        class Trial(pytorch.PyTorchTrial):
            def __init__(self, context):
                self.context = context
                torch.manual_seed(1000 + rank)  # different initial weights per rank: rank 0's must win
                model = GoldenNet()
                if rank == 0:
                    with torch.no_grad():
                        for p, init in zip(model.parameters(), g.init_params()):
                            p.copy_(torch.from_numpy(init))
                self.model = context.wrap_model(model)
                self.plist = list(self.model.parameters())
                self.opt = context.wrap_optimizer(make_optimizer(self.model, g.hp), average_aggregated_gradients=g.average)
                x, y = make_data(g.steps * g.agg, g.hp["global_batch_size"])
                self.ds = torch.utils.data.TensorDataset(x, y)

            def build_training_data_loader(self):
                return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

            def _clip(self, params):
                params = list(params)
                if not sharded:
                    rec["avg"].append([_np(p.grad) for p in params])
                if g.clip is not None:
                    rec["norm"].append(float(torch.nn.utils.clip_grad_norm_(params, g.clip[1])))

            def train_batch(self, batch, epoch_idx, batch_idx):
                x, y = batch
                loss = torch.nn.functional.mse_loss(self.model(x), y)
                self.context.backward(loss)
                self.context.step_optimizer(self.opt, clip_grads=self._clip if (g.clip is not None or not sharded) else None)
                if (batch_idx + 1) % g.agg == 0:
                    rec["params"].append([_np(p) for p in self.plist])
                return {"loss": loss}

        pytorch._PyTorchTrialController.pre_execute_hook(17)
        conf = {"optimizations": {"sharded_step": True}} if route == "sharded" else {"optimizations": {"grad_sync_route": route}}
        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams=g.hp, slots_per_trial=size, aggregation_frequency=g.agg,
                                          trial_seed=17, exp_conf=conf)
        trial = Trial(ctx)
        eng = ctx._engines[trial.opt]
        assert eng.sharded == sharded and eng.comm is not None and eng.peer == (route == "peer")
        ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=g.steps * g.agg, reporting_period=g.agg,
                                              on_report=lambda n, r: reports.append((n, r)))
        ctl.run()
        for s in range(g.steps):
            if not sharded:
                want_avg = g.per_step("avg_grads", s)
                for j, i in enumerate(OPT_ORDER_TO_MODEL):
                    assert_close(rec["avg"][s][j], want_avg[i], "float32", f"{name}: synced grad, step {s}, param {i}")
            for i in range(N_PARAMS):
                assert_close(rec["params"][s][i], g.per_step("params", s)[i], "float32", f"{name}: parameter {i} after step {s}")
            if g.clip is not None:
                assert_close(rec["norm"][s], g.z["total_norm"][s], "float32", f"{name}: total norm step {s}")
        key = "momentum_buffer" if g.hp["optimizer"] == "sgd" else "exp_avg"
        state = trial.opt.state_dict()["state"]  # sharded: gathers the other rank's slices (a collective: every rank asks)
        for j, i in enumerate(OPT_ORDER_TO_MODEL):
            assert_close(_np(state[j][key]), g.state(g.steps - 1, i, key), "float32", f"{name}: {key} of param {i}")
        kinds = [e[0] for e in fake.log]
        assert ("reduce_scatter" in kinds) == (route == "sharded")
        # peer route: the flatten kernel's stores are the transfer, the owner reduces, barriers frame the step
        assert ("flatten_push" in kinds) == ("peer_reduce" in kinds) == ("peer_barrier" in kinds) == (route == "peer")
        if route == "peer":
            assert ("peer_norm_exchange" in kinds) == (g.clip is not None)
        assert any(e[0] == "allreduce" and e[1] > 1 for e in fake.log) != sharded  # sharded: only the norm's one float is all-reduced
        assert "broadcast" in kinds  # rank 0's parameters at wrap_optimizer; the learned bucket order
        if rank == 0:
            assert [n for n, _ in reports] == g.z["reported_steps"].tolist()
            for (n, r), want in zip(reports, g.z["reported_avg_loss"]):
                assert_close(r["avg_metrics"]["loss"], want, "float32", f"{name}: reported avg loss at {n}")
        else:
            assert reports == []
        # one allreduce (or reduce-scatter) per bucket per optimizer step, issued in the same order on both ranks
        n_buckets = sum(len(fs.buckets) for fs in eng.sets)
        assert (n_buckets >= 3) == small_buckets
        per_step = [e for e in fake.log if e[0] in ("allreduce", "reduce_scatter", "peer_reduce") and e[1] > 1]
        assert len(per_step) == n_buckets * g.steps
        everyone = dctx.allgather([e[1] for e in per_step])
        assert everyone[0] == everyone[1]
        ctx.close()
        open(os.path.join(out_dir, f"ok{rank}-{name}-{route}-{small_buckets}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()
        if small_buckets:
            _gradsync.GradSyncEngine = real_engine


SCENARIOS_2 = [("sgd_n2_a2", False, True), ("adamw_n2_a4_clip", False, False), ("adamw_n2_a4_clip", True, True),
               ("sgd_n2_a2", "peer", True), ("adamw_n2_a4_clip", "peer", True)]


def _scenarios_worker(rank: int, size: int, ports, out_dir: str) -> None:
    """All scenario / route / bucket combinations in ONE pair of processes (each spawn costs ~12 s of interpreter and
    torch start-up, the scenarios themselves a second or two); every combination gets its own process group, control
    plane and ports, and leaves its own marker file, so a failure names the combination."""
    for (name, sharded, small_buckets), (port, zmq_ports) in zip(SCENARIOS_2, ports):
        _worker(rank, size, port, zmq_ports, name, sharded, small_buckets, out_dir)


def test_two_rank_scenarios_with_emulated_kernels_and_gloo(tmp_path):
    import torch.multiprocessing as mp

    ports = [(_free_port(), (_free_port(), _free_port())) for _ in SCENARIOS_2]
    mp.spawn(_scenarios_worker, args=(2, ports, str(tmp_path)), nprocs=2, join=True)
    for name, sharded, small_buckets in SCENARIOS_2:
        route = {False: "allreduce", True: "sharded"}.get(sharded, sharded)
        for rank in range(2):
            assert (tmp_path / f"ok{rank}-{name}-{route}-{small_buckets}").exists(), (name, route, small_buckets)


def _auto_worker(rank: int, size: int, port: int, zmq_ports, name: str, out_dir: str) -> None:
    """Route auto (the default) with a clip function that is provably nothing but clip_grad_norm_: two optimizer steps on
    the allreduce route, then the peer route -- on buffers that the allreduce route has used, which is what matters:
    the rows of a bucket's alignment gaps must not carry last step's gradients into the norm."""
    for p in (str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    import fake_dgs
    from golden_io import N_PARAMS, Golden, assert_close
    from golden_model import GoldenNet, make_data, make_optimizer

    from determined_clone_b200 import core, pytorch
    from determined_clone_b200.pytorch import _gradsync

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    fake = fake_dgs.install()

    class SmallBuckets(_gradsync.GradSyncEngine):
        def __init__(self, *args, **kwargs):
            kwargs.update(bucket_bytes=1024, first_bucket_bytes=256)
            super().__init__(*args, **kwargs)

    _gradsync.GradSyncEngine = SmallBuckets
    _gradsync._PEER_OVERLAP_MIN = 1  # every flat set takes the per-bucket mode: barrier + owner reduction as each bucket is pushed
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        g = Golden(name)
        params_seen = []
        max_norm = float(g.clip[1])

        class Trial(pytorch.PyTorchTrial):
            def __init__(self, context):
                self.context = context
                model = GoldenNet()
                with torch.no_grad():
                    for p, init in zip(model.parameters(), g.init_params()):
                        p.copy_(torch.from_numpy(init))
                self.model = context.wrap_model(model)
                self.plist = list(self.model.parameters())
                self.opt = context.wrap_optimizer(make_optimizer(self.model, g.hp), average_aggregated_gradients=g.average)
                x, y = make_data(g.steps * g.agg, g.hp["global_batch_size"])
                self.ds = torch.utils.data.TensorDataset(x, y)

            def build_training_data_loader(self):
                return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

            def train_batch(self, batch, epoch_idx, batch_idx):
                x, y = batch
                loss = torch.nn.functional.mse_loss(self.model(x), y)
                self.context.backward(loss)
                self.context.step_optimizer(self.opt, clip_grads=lambda params: torch.nn.utils.clip_grad_norm_(params, max_norm))
                if (batch_idx + 1) % g.agg == 0:
                    params_seen.append([_np(p) for p in self.plist])
                return {"loss": loss}

        pytorch._PyTorchTrialController.pre_execute_hook(17)
        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams=g.hp, slots_per_trial=size, aggregation_frequency=g.agg, trial_seed=17)
        trial = Trial(ctx)
        eng = ctx._engines[trial.opt]
        assert eng.route == "auto" and not eng.sharded and eng.peers is not None
        pytorch._PyTorchTrialController(trial, ctx, max_length=g.steps * g.agg, reporting_period=g.agg).run()
        assert eng.peer and eng.sharded, "route auto did not move to the peer kernels"
        for s in range(g.steps):
            for i in range(N_PARAMS):
                assert_close(params_seen[s][i], g.per_step("params", s)[i], "float32", f"{name}: parameter {i} after step {s}")
        kinds = [e[0] for e in fake.log]
        n_buckets = sum(len(fs.buckets) for fs in eng.sets)
        assert n_buckets >= 3
        assert sum(1 for e in fake.log if e[0] == "allreduce" and e[1] > 1) == 2 * n_buckets  # the first two steps
        assert kinds.count("peer_norm_exchange") == g.steps - 2
        # per-bucket mode: one barrier and one owner reduction per bucket per step (+ the parameter barrier per step)
        assert all(fs.peer_overlap for fs in eng.sets)
        assert kinds.count("peer_reduce") == (g.steps - 2) * n_buckets
        assert kinds.count("peer_barrier") == (g.steps - 2) * (n_buckets + 1)
        ctx.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_route_auto_moves_to_the_peer_kernels_and_keeps_matching_the_reference(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_auto_worker, args=(2, _free_port(), (_free_port(), _free_port()), "adamw_n2_a4_clip", str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def _demote_worker(rank: int, size: int, port: int, zmq_ports, out_dir: str) -> None:
    """Route auto end to end against live torch DDP over gloo: AdamW + plain clip for a few steps (allreduce, then the
    peer kernels), then ONE step that the sliced routes cannot honour (auto_zero_grads=False: the caller keeps
    p.grad, so every rank needs the whole averaged gradient) -- the engine must bring the reduced slices and the
    other ranks' optimizer state back (demote) in the middle of that step, finish it on the allreduce route and stay
    there; parameters and Adam state must follow torch through all of it."""
    for p in (str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import copy

    import torch.distributed as dist

    import fake_dgs
    from golden_io import assert_close

    from determined_clone_b200 import core, pytorch

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    fake = fake_dgs.install()
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        torch.manual_seed(11)
        base = torch.nn.Sequential(torch.nn.Linear(10, 48), torch.nn.LayerNorm(48), torch.nn.Tanh(), torch.nn.Linear(48, 3))
        gen = torch.Generator().manual_seed(500 + rank)
        data = [(torch.randn(8, 10, generator=gen), torch.randn(8, 3, generator=gen)) for _ in range(8)]
        keep_grads_at, max_norm = 4, 0.3

        ref_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
        ref_opt = torch.optim.AdamW(ref_model.parameters(), lr=1e-2, betas=(0.9, 0.95), weight_decay=0.05)
        want, want_kept = [], None
        for i, (x, y) in enumerate(data):
            torch.nn.functional.mse_loss(ref_model(x), y).backward()
            torch.nn.utils.clip_grad_norm_(ref_model.parameters(), max_norm)
            ref_opt.step()
            if i == keep_grads_at:
                want_kept = [_np(p.grad) for p in ref_model.parameters()]
            ref_opt.zero_grad()
            want.append(([_np(p) for p in ref_model.parameters()], [_np(ref_opt.state[p]["exp_avg_sq"]) for p in ref_model.parameters()]))

        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 16}, slots_per_trial=size)
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, betas=(0.9, 0.95), weight_decay=0.05))
        eng = ctx._engines[opt]
        routes = []
        for i, (x, y) in enumerate(data):
            ctx._current_batch_idx = i
            ctx.backward(torch.nn.functional.mse_loss(model(x), y))
            routes.append((eng.sharded, eng.peer))
            ctx.step_optimizer(opt, clip_grads=lambda ps: torch.nn.utils.clip_grad_norm_(ps, max_norm), auto_zero_grads=i != keep_grads_at)
            if i == keep_grads_at:
                # the caller asked to keep the gradients: whole, averaged over the ranks and clipped, as torch leaves them
                for p, g in zip(model.parameters(), want_kept):
                    assert_close(_np(p.grad), g, "float32", "kept gradient")
                opt.zero_grad()
            state = opt.state_dict()["state"]  # (a collective while the engine is sliced: every rank asks)
            for j, (p, (wp, wv)) in enumerate(zip(model.parameters(), zip(*want[i]))):
                assert_close(_np(p), wp, "float32", f"parameter {j} after step {i}")
                assert_close(_np(state[j]["exp_avg_sq"]), wv, "float32", f"exp_avg_sq {j} after step {i}")
        # two steps on NCCL allreduce, then the peer kernels, then back for good from the step that kept its gradients
        assert routes == [(False, False)] * 2 + [(True, True)] * (keep_grads_at - 1) + [(False, False)] * (len(data) - keep_grads_at - 1), routes
        assert eng.route == "allreduce"
        ctx.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_route_auto_returns_to_allreduce_for_a_step_the_sliced_routes_cannot_serve(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_demote_worker, args=(2, _free_port(), (_free_port(), _free_port()), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def _amp_worker(rank: int, size: int, port: int, zmq_ports, out_dir: str) -> None:
    """tests/test_gpu_amp.py's two-slot case on the CPU: auto-AMP with accumulation and clipping; an inf in ONE rank's
    data must make EVERY rank skip that step (the check runs on the all-reduced gradients), against torch DDP over
    gloo with torch's own GradScaler protocol."""
    for p in (str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import contextlib
    import copy

    import torch.distributed as dist

    import fake_dgs
    from golden_io import assert_close

    from determined_clone_b200 import core, pytorch

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    fake_dgs.install()
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        agg, clip = 2, 0.5
        torch.manual_seed(5)
        base = torch.nn.Sequential(torch.nn.Linear(12, 32), torch.nn.LayerNorm(32), torch.nn.Tanh(), torch.nn.Linear(32, 5))
        g = torch.Generator().manual_seed(300 + rank)
        data = [(torch.randn(8, 12, generator=g), torch.randn(8, 5, generator=g)) for _ in range(8 * agg)]
        if rank == 1:
            data[4][0][0, 0] = float("inf")  # first micro-batch of the third window, on rank 1 only

        def new_scaler():
            return torch.amp.GradScaler("cpu", init_scale=1024.0, growth_interval=2)

        def snap(model, opt, scaler, out):
            out["params"].append([_np(p) for p in model.parameters()])
            out["scale"].append(scaler.get_scale())
            out["steps"].append([float(opt.state[p]["step"]) for p in model.parameters() if "step" in opt.state.get(p, {})])

        ref_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
        ref_opt, ref_scaler = torch.optim.AdamW(ref_model.parameters(), lr=1e-2, betas=(0.9, 0.95)), new_scaler()
        ref = {"params": [], "scale": [], "steps": [], "norms": []}
        params = list(ref_model.parameters())
        for i, (x, y) in enumerate(data):
            boundary = (i + 1) % agg == 0
            with contextlib.nullcontext() if boundary else ref_model.no_sync():
                ref_scaler.scale(torch.nn.functional.mse_loss(ref_model(x), y)).backward()
            if not boundary:
                continue
            for p in params:
                p.grad.div_(float(agg))
            ref_scaler.unscale_(ref_opt)
            ref["norms"].append(float(torch.nn.utils.clip_grad_norm_(params, clip)))
            ref_scaler.step(ref_opt)
            ref_opt.zero_grad()
            ref_scaler.update()
            snap(ref_model, ref_opt, ref_scaler, ref)

        out = {"params": [], "scale": [], "steps": [], "norms": []}

        class Trial(pytorch.PyTorchTrial):
            def __init__(self, context):
                self.context = context
                self.model = context.wrap_model(copy.deepcopy(base))
                context.experimental.use_amp()
                context.wrap_scaler(new_scaler())
                self.opt = context.wrap_optimizer(torch.optim.AdamW(self.model.parameters(), lr=1e-2, betas=(0.9, 0.95)))

            def build_training_data_loader(self):
                raise AssertionError("batches are fed by the test")

            def train_batch(self, batch, epoch_idx, batch_idx):
                loss = torch.nn.functional.mse_loss(self.model(batch[0]), batch[1])
                self.context.backward(loss)
                self.context.step_optimizer(self.opt, clip_grads=lambda ps: out["norms"].append(float(torch.nn.utils.clip_grad_norm_(list(ps), clip))))
                return {"loss": loss}

        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 16}, slots_per_trial=size, aggregation_frequency=agg)
        trial = Trial(ctx)
        ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=len(data), reporting_period=len(data))
        for i, batch in enumerate(data):
            ctx._current_batch_idx = i
            ctl._train_batch(batch, 0, i)
            if (i + 1) % agg == 0:
                snap(trial.model, trial.opt, ctx._scaler, out)
        assert ref["scale"][:3] == [1024.0, 2048.0, 1024.0] and ref["steps"][2] == ref["steps"][1]  # the third step was skipped
        assert out["scale"] == ref["scale"] and out["steps"] == ref["steps"]
        for a, b in zip(out["norms"], ref["norms"]):
            assert np.isfinite(a) == np.isfinite(b)
            if np.isfinite(b):
                np.testing.assert_allclose(a, b, rtol=1e-5)
        for s, (pa, pb) in enumerate(zip(out["params"], ref["params"])):
            for a, b in zip(pa, pb):
                assert_close(a, b, "float32", f"after optimizer step {s}")
        ctx.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_two_rank_auto_amp_skips_on_every_rank(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_amp_worker, args=(2, _free_port(), (_free_port(), _free_port()), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def _bn_worker(rank: int, size: int, port: int, zmq_ports, out_dir: str) -> None:
    """Module buffers under 2 ranks (tests/test_gpu_trial_parity.py's batch-norm case on the CPU): rank 0's initial
    weights AND buffers win, rank 0's running statistics overwrite the others' before every training forward
    (DDP's broadcast_buffers), num_batches_tracked stays an exact integer -- against torch DDP over gloo."""
    for p in (str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import copy

    import torch.distributed as dist

    import fake_dgs
    from oracle import torch_port

    from determined_clone_b200 import core, pytorch

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    fake = fake_dgs.install()
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        torch.manual_seed(100 + rank)  # different initial weights and buffers per rank
        base = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.BatchNorm2d(8), torch.nn.ReLU(),
                                   torch.nn.AdaptiveAvgPool2d(1), torch.nn.Flatten(), torch.nn.Linear(8, 4))
        with torch.no_grad():
            base[1].running_mean.add_(rank + 1.0)
        g = torch.Generator().manual_seed(7 + rank)
        data = [(torch.randn(4, 3, 8, 8, generator=g), torch.randn(4, 4, generator=g)) for _ in range(5)]

        ref = torch_port.RefContext()
        ref_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
        ref.models.append(ref_model)
        ref_opt = ref.wrap_optimizer(torch.optim.SGD(ref_model.parameters(), lr=0.1, momentum=0.9))
        ours = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 8}, slots_per_trial=size)
        model = ours.wrap_model(copy.deepcopy(base))
        opt = ours.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9))
        assert len(ours._buffer_syncs) == 1 and ours._buffer_syncs[0].flats, "the model's buffers were not picked up"
        for i, (x, y) in enumerate(data):
            for ctx, mdl, o in ((ref, ref_model, ref_opt), (ours, model, opt)):
                ctx._current_batch_idx = i
                ctx.backward(torch.nn.functional.mse_loss(mdl(x), y))
                ctx.step_optimizer(o)
        for (name, a), b in zip(model.state_dict().items(), ref_model.module.state_dict().values()):
            if a.dtype.is_floating_point:
                np.testing.assert_allclose(a.numpy(), b.numpy(), rtol=1e-5, atol=1e-6, err_msg=name)
            else:
                assert torch.equal(a, b), name
        flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
        gathered = [torch.empty_like(flat) for _ in range(size)]
        dist.all_gather(gathered, flat)
        assert all(torch.equal(gathered[0], t) for t in gathered)
        # one buffer broadcast at wrap_model plus one per training forward; evaluation does not broadcast
        n_before = sum(1 for e in fake.log if e[0] == "broadcast")
        model.eval()
        model(data[0][0])
        assert sum(1 for e in fake.log if e[0] == "broadcast") == n_before
        ours.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_two_rank_module_buffers_follow_rank_zero(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_bn_worker, args=(2, _free_port(), (_free_port(), _free_port()), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def _wide_worker(rank: int, size: int, port: int, zmq_ports, per_bucket: bool, out_dir: str) -> None:
    """More than two ranks (the fixtures recorded from the reference stop at two slots): the peer route pinned, several
    buckets, aggregation over two micro-batches, two parameter groups, clipping -- against live torch DDP over gloo on
    the same ranks and data.  What changes with the world size is the slicing: a bucket is `size` rows of span =
    bucket / size elements, owner r reduces row block r and steps it, the ranges of a parameter group are cut at the
    slice borders; with `per_bucket` every bucket is reduced as soon as its pushes have landed."""
    for p in (str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import copy

    import torch.distributed as dist

    import fake_dgs
    from golden_io import assert_close

    from determined_clone_b200 import core, pytorch
    from determined_clone_b200.pytorch import _gradsync

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    fake = fake_dgs.install()

    class SmallBuckets(_gradsync.GradSyncEngine):
        def __init__(self, *args, **kwargs):
            kwargs.update(bucket_bytes=2048, first_bucket_bytes=512)
            super().__init__(*args, **kwargs)

    _gradsync.GradSyncEngine = SmallBuckets
    _gradsync._PEER_OVERLAP_MIN = 1 if per_bucket else 1 << 40
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        torch.manual_seed(23)
        base = torch.nn.Sequential(torch.nn.Linear(12, 40), torch.nn.LayerNorm(40), torch.nn.Tanh(), torch.nn.Linear(40, 33),
                                   torch.nn.Tanh(), torch.nn.Linear(33, 5))
        gen = torch.Generator().manual_seed(700 + rank)
        agg, steps, max_norm = 2, 5, 0.25
        data = [(torch.randn(6, 12, generator=gen), torch.randn(6, 5, generator=gen)) for _ in range(agg * steps)]

        def groups(model):
            decay = [p for p in model.parameters() if p.ndim > 1]
            rest = [p for p in model.parameters() if p.ndim <= 1]
            return [{"params": decay, "weight_decay": 0.1}, {"params": rest, "weight_decay": 0.0, "lr": 2e-2}]

        ref_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base))
        ref_opt = torch.optim.AdamW(groups(ref_model), lr=1e-2, betas=(0.9, 0.95))
        want = []
        for i, (x, y) in enumerate(data):  # the reference's lines: no_sync off the boundary, /A, clip, step (_pytorch_context.py:802-924)
            if (i + 1) % agg:
                with ref_model.no_sync():
                    torch.nn.functional.mse_loss(ref_model(x), y).backward()
                continue
            torch.nn.functional.mse_loss(ref_model(x), y).backward()
            for p in ref_model.parameters():
                p.grad.div_(agg)
            norm = torch.nn.utils.clip_grad_norm_(ref_model.parameters(), max_norm)
            ref_opt.step()
            ref_opt.zero_grad()
            want.append((float(norm), [_np(p) for p in ref_model.parameters()], [_np(ref_opt.state[p]["exp_avg"]) for p in ref_model.parameters()]))

        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 6 * size}, slots_per_trial=size, aggregation_frequency=agg,
                                          exp_conf={"optimizations": {"grad_sync_route": "peer"}})
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(torch.optim.AdamW(groups(model), lr=1e-2, betas=(0.9, 0.95)))
        eng = ctx._engines[opt]
        assert eng.peer and eng.sharded and eng.world == size
        n_buckets = sum(len(fs.buckets) for fs in eng.sets)
        assert n_buckets >= 3 and all(fs.peer_overlap == per_bucket for fs in eng.sets)
        norms = []
        index_of = {p: j for j, p in enumerate(p for g in opt.param_groups for p in g["params"])}
        for i, (x, y) in enumerate(data):
            ctx._current_batch_idx = i
            ctx.backward(torch.nn.functional.mse_loss(model(x), y))
            ctx.step_optimizer(opt, clip_grads=lambda ps: norms.append(torch.nn.utils.clip_grad_norm_(ps, max_norm)))
            if (i + 1) % agg:
                continue
            s = i // agg
            state = opt.state_dict()["state"]  # (a collective on the sliced routes: every rank asks)
            assert_close(float(norms[-1]), want[s][0], "float32", f"total norm, step {s}")
            for j, p in enumerate(model.parameters()):
                assert_close(_np(p), want[s][1][j], "float32", f"parameter {j} after step {s}")
                assert_close(_np(state[index_of[p]]["exp_avg"]), want[s][2][j], "float32", f"exp_avg {j} after step {s}")
        kinds = [e[0] for e in fake.log]
        assert kinds.count("flatten_push") >= steps * n_buckets and kinds.count("peer_norm_exchange") == steps
        # per bucket: one barrier + one reduction per bucket; otherwise one barrier and one (multi-bucket) reduction per step
        assert kinds.count("peer_barrier") == steps * ((n_buckets if per_bucket else 1) + 1)
        if per_bucket:
            assert kinds.count("peer_reduce") == steps * n_buckets
        ctx.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()


@pytest.mark.parametrize("size,per_bucket", [(4, True), (8, False)])
def test_peer_route_on_four_and_eight_ranks_matches_torch_ddp(tmp_path, size, per_bucket):
    import torch.multiprocessing as mp

    mp.spawn(_wide_worker, args=(size, _free_port(), (_free_port(), _free_port()), per_bucket, str(tmp_path)), nprocs=size, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(size))
