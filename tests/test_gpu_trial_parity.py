"""End-to-end parity: the same trial code that produced the golden fixtures under the REFERENCE
harness (tests/golden/make_golden.py) runs here under this repo's harness on the GPU -- same API
calls (wrap_model, wrap_optimizer, backward, step_optimizer(clip_grads), aggregation_frequency) -- and
must reproduce the reference's gradients, norms, parameters, optimizer state, losses and reported
metrics over 10 optimizer steps (rel 1e-5 fp32 / 1e-2 bf16; integer counters exactly)."""
import os
import socket
from typing import Any, Dict, List

import numpy as np
import pytest

from golden_io import N_PARAMS, SCENARIOS, Golden, assert_close
from golden_model import GoldenNet, make_data, make_optimizer

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import determined_clone_b200 as det  # noqa: E402
from determined_clone_b200 import core, pytorch  # noqa: E402


def _np(t):
    return t.detach().float().cpu().numpy().copy()


class GoldenTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext, g: Golden, rec: Dict[str, List[Any]], lazy_import: bool = False) -> None:
        self.context, self.g, self.rec = context, g, rec
        self.lazy_import = lazy_import
        dtype = getattr(torch, g.dtype)
        model = GoldenNet().to(dtype)
        with torch.no_grad():
            for p, init in zip(model.parameters(), g.init_params()):
                p.copy_(torch.from_numpy(init).to(dtype))
        self.model = context.wrap_model(model)
        self.plist = list(self.model.parameters())
        self.opt = context.wrap_optimizer(make_optimizer(self.model, g.hp), average_aggregated_gradients=g.average)
        x, y = make_data(g.steps * g.agg, g.hp["global_batch_size"])
        self.ds = torch.utils.data.TensorDataset(x.to(dtype), y.to(dtype))

    def build_training_data_loader(self):
        return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

    def _clip(self, params):
        params = list(params)
        self.rec["avg"].append([_np(p.grad) for p in params])
        if self.g.clip is not None:
            kind, val = self.g.clip
            if kind == "norm":
                self.rec["norm"].append(torch.nn.utils.clip_grad_norm_(params, val))
            else:
                torch.nn.utils.clip_grad_value_(params, val)

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x).float(), y.float())
        self.context.backward(loss)
        self.context.step_optimizer(self.opt, clip_grads=self._clip)
        if (batch_idx + 1) % self.g.agg == 0:
            self.rec["params"].append([_np(p) for p in self.plist])
            self.rec["state"].append([{k: (_np(v) if torch.is_tensor(v) else v) for k, v in self.opt.state[p].items()} for p in self.plist])
        return {"loss": loss}


def run_scenario(name: str, distributed=None) -> None:
    g = Golden(name)
    rec: Dict[str, List[Any]] = {"avg": [], "norm": [], "params": [], "state": []}
    reports = []
    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(
        distributed=distributed, hparams=g.hp, slots_per_trial=g.slots, aggregation_frequency=g.agg, trial_seed=17,
    )
    trial = GoldenTrial(ctx, g, rec)
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=g.steps * g.agg, reporting_period=g.agg,
                                          on_report=lambda n, r: reports.append((n, r)))
    launches0 = det._lib.calls.dgs_launch_count()
    ctl.run()
    torch.cuda.synchronize()
    assert det._lib.calls.dgs_launch_count() > launches0, "the CUDA extension did not run"
    dt = g.dtype
    # bf16 end to end: the raw gradients themselves come from a different device here (GPU matmuls vs the
    # CPU run that produced the fixtures) and already differ by ~1e-2; Adam divides by sqrt(v), so that
    # input noise is amplified a few times in the state and in zero-initialised parameters.  The path's own
    # bf16 parity at 1e-2 is asserted on identical inputs in test_gpu_golden_replay.py.
    # Measured on B200 (profiles/r2_tolerance_report_gpu.txt): the 1-slot bf16 scenario stays within 1.15x of the plain
    # rule, so 3x; the 2-slot scenarios only run where two GPUs are (inside bench.py's parity pass) and keep the rule
    # and the factor they were last measured with.
    two_slot = g.slots > 1
    wide = 1.0 if dt == "float32" else (5.0 if two_slot else 3.0)
    rule = "tensor" if two_slot else None  # None: golden_io's default (fp32: 1e-5*(|x| + rms); bf16: tensor-relative)
    opt_order_to_model = [0, 4, 1, 2, 3, 5]
    for s in range(g.steps):
        want_avg = g.per_step("avg_grads", s)
        for j, i in enumerate(opt_order_to_model):
            assert_close(rec["avg"][s][j], want_avg[i], dt, f"{name}: synced grad, step {s}, param {i}", rule=rule)
        for i in range(N_PARAMS):
            assert_close(rec["params"][s][i], g.per_step("params", s)[i], dt, f"{name}: parameter {i} after step {s}", factor=wide, rule=rule)
            st = rec["state"][s][i]
            if g.hp["optimizer"] == "sgd":
                assert_close(st["momentum_buffer"], g.state(s, i, "momentum_buffer"), dt, f"{name}: momentum {i} step {s}", rule=rule)
            elif g.hp["optimizer"] in ("adam", "adamw"):
                assert_close(st["exp_avg"], g.state(s, i, "exp_avg"), dt, f"{name}: exp_avg {i} step {s}", factor=wide, rule=rule)
                assert_close(st["exp_avg_sq"], g.state(s, i, "exp_avg_sq"), dt, f"{name}: exp_avg_sq {i} step {s}", factor=wide, rule=rule)
                assert float(st["step"]) == float(g.state(s, i, "step")) == s + 1  # integer counter: exact
        if g.clip is not None and g.clip[0] == "norm":
            assert_close(float(rec["norm"][s]), g.z["total_norm"][s], dt, f"{name}: total norm step {s}", rule=rule)
    if ctx.distributed.rank == 0:
        # chief's reports: batch counters exactly, averaged loss within tolerance
        assert [n for n, _ in reports] == g.z["reported_steps"].tolist()
        for (n, r), want_avg, want_batches in zip(reports, g.z["reported_avg_loss"], g.z["reported_batch_loss"]):
            assert_close(r["avg_metrics"]["loss"], want_avg, dt, f"{name}: reported avg loss at {n}", rule=rule)
            assert_close([float(b["loss"]) for b in r["batch_metrics"]], want_batches, dt, f"{name}: per-batch loss at {n}", rule=rule)
    ctx.close()


@pytest.mark.parametrize("name", [n for n, sc in SCENARIOS.items() if sc["slots"] == 1])
def test_single_slot_trial_matches_reference(lib, name):
    run_scenario(name)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank: int, size: int, name: str, ports) -> None:
    torch.cuda.set_device(rank)
    dist = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=ports[0], pull_port=ports[1])
    try:
        run_scenario(name, distributed=dist)
    finally:
        dist.close()


@pytest.mark.parametrize("name", [n for n, sc in SCENARIOS.items() if sc["slots"] == 2])
def test_two_slot_trial_matches_reference(lib, name):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp

    ports = (_free_port(), _free_port())
    mp.spawn(_rank_main, args=(2, name, ports), nprocs=2, join=True)


# ---- module buffers (batch norm) under 2 slots: ours vs torch DDP on the same GPUs -------------------------
class _BnNet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = torch.nn.Conv2d(3, 8, 3, padding=1)
        self.bn = torch.nn.BatchNorm2d(8)
        self.fc = torch.nn.Linear(8, 4)

    def forward(self, x):
        return self.fc(torch.relu(self.bn(self.conv(x))).mean(dim=(2, 3)))


def _bn_rank_main(rank: int, size: int, ports, init_port: int) -> None:
    import copy

    import torch.distributed as dist

    from oracle import torch_port

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{init_port}", rank=rank, world_size=size,
                            device_id=torch.device("cuda", rank))
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=ports[0], pull_port=ports[1])
    try:
        torch.manual_seed(100 + rank)  # DIFFERENT initial weights and buffers per rank: rank 0's must win
        base = _BnNet()
        with torch.no_grad():
            base.bn.running_mean.add_(rank + 1.0)
        g = torch.Generator().manual_seed(7 + rank)
        data = [(torch.randn(4, 3, 8, 8, generator=g).cuda(), torch.randn(4, 4, generator=g).cuda()) for _ in range(5)]

        ref = torch_port.RefContext(device=f"cuda:{rank}")
        ref_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base).cuda())  # reference: _pytorch_context.py:295-296
        ref.models.append(ref_model)
        ref_opt = ref.wrap_optimizer(torch.optim.SGD(ref_model.parameters(), lr=0.1, momentum=0.9))
        ours = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 8}, slots_per_trial=size)
        model = ours.wrap_model(copy.deepcopy(base))
        opt = ours.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9))
        for i, (x, y) in enumerate(data):
            for ctx, mdl, o in ((ref, ref_model, ref_opt), (ours, model, opt)):
                ctx._current_batch_idx = i
                ctx.backward(torch.nn.functional.mse_loss(mdl(x), y))
                ctx.step_optimizer(o)
        torch.cuda.synchronize()
        for (name, a), b in zip(model.state_dict().items(), ref_model.module.state_dict().values()):
            if a.dtype.is_floating_point:
                np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=1e-5, atol=1e-6, err_msg=name)
            else:
                assert torch.equal(a.cpu(), b.cpu()), name  # num_batches_tracked: integer, exact
        # every rank ends with identical parameters (rank 0's initial weights were broadcast)
        flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
        gathered = [torch.empty_like(flat) for _ in range(size)]
        dist.all_gather(gathered, flat)
        assert all(torch.equal(gathered[0], t) for t in gathered)
        ours.close()
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_two_slot_buffers_and_initial_broadcast_match_torch_ddp(lib):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp

    mp.spawn(_bn_rank_main, args=(2, (_free_port(), _free_port()), _free_port()), nprocs=2, join=True)


# ---- opt-in sharded step (reduce-scatter, step on 1/N of each bucket, all-gather parameters) ----------------
class ShardedGoldenTrial(GoldenTrial):
    """Same trial, but the clip function only does what sharded mode allows: clip_grad_norm_ on its argument."""

    def _clip(self, params):
        self.rec["norm"].append(torch.nn.utils.clip_grad_norm_(list(params), self.g.clip[1]))

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x).float(), y.float())
        self.context.backward(loss)
        self.context.step_optimizer(self.opt, clip_grads=self._clip if self.g.clip is not None else None)
        if (batch_idx + 1) % self.g.agg == 0:
            self.rec["params"].append([_np(p) for p in self.plist])
        return {"loss": loss}


def _sharded_rank_main(rank: int, size: int, name: str, ports, route: str = "sharded") -> None:
    torch.cuda.set_device(rank)
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=ports[0], pull_port=ports[1])
    try:
        g = Golden(name)
        rec = {"avg": [], "norm": [], "params": [], "state": []}
        pytorch._PyTorchTrialController.pre_execute_hook(17)
        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams=g.hp, slots_per_trial=g.slots, aggregation_frequency=g.agg,
                                          exp_conf={"optimizations": {"sharded_step": True} if route == "sharded" else {"grad_sync_route": route}})
        trial = ShardedGoldenTrial(ctx, g, rec)
        if route == "auto" and g.clip is not None:
            # route auto only leaves full gradients behind for a clip function that provably never reads one: the
            # reference's own example, a lambda around clip_grad_norm_ (its norm is not recorded then)
            max_norm = float(g.clip[1])
            trial._clip = lambda params: torch.nn.utils.clip_grad_norm_(params, max_norm)
        eng = ctx._engines[trial.opt]
        # "auto" starts on NCCL allreduce and moves to the peer kernels after two optimizer steps (the clip function
        # of this trial is provably nothing but clip_grad_norm_); "peer" runs them from the first step
        assert eng.sharded == (route != "auto") and eng.peer == (route == "peer")
        pytorch._PyTorchTrialController(trial, ctx, max_length=g.steps * g.agg, reporting_period=g.agg).run()
        torch.cuda.synchronize()
        if route == "auto":
            assert eng.peer and eng.sharded, f"route auto did not reach the peer kernels: {eng.peer_unavailable}"
        dt = g.dtype
        wide = 1.0 if dt == "float32" else 5.0
        rule = "tensor"  # 2 slots: the rule these comparisons were last measured with on two GPUs
        for s in range(g.steps):
            for i in range(N_PARAMS):  # every rank ends every step with the full, identical parameters
                assert_close(rec["params"][s][i], g.per_step("params", s)[i], dt, f"{name}: parameter {i} after step {s}", factor=wide, rule=rule)
            if g.clip is not None and rec["norm"]:
                assert_close(float(rec["norm"][s]), g.z["total_norm"][s], dt, f"{name}: total norm step {s}", rule=rule)
        # optimizer.state_dict() gathers the other ranks' slices: complete and in the reference layout
        sd = trial.opt.state_dict()["state"]
        key = "momentum_buffer" if g.hp["optimizer"] == "sgd" else "exp_avg"
        for j, i in enumerate([0, 4, 1, 2, 3, 5]):  # optimizer order -> model order
            assert_close(_np(sd[j][key]), g.state(g.steps - 1, i, key), dt, f"{name}: gathered {key} of param {i}", factor=wide, rule=rule)
        with pytest.raises(det.errors.InvalidExperimentException, match="sharded_step"):
            ctx._current_batch_idx = g.agg - 1  # a boundary batch
            ctx.backward(trial.model(torch.randn(4, 24, device="cuda").to(trial.plist[0].dtype)).float().sum())
            ctx.step_optimizer(trial.opt, clip_grads=lambda ps: torch.nn.utils.clip_grad_value_(ps, 0.1))
        ctx.close()
    finally:
        dctx.close()


@pytest.mark.parametrize("route", ["sharded", "peer", "auto"])
@pytest.mark.parametrize("name", [n for n, sc in SCENARIOS.items() if sc["slots"] == 2])
def test_two_slot_sharded_step_matches_reference(lib, name, route):
    """The sliced routes against the fixtures recorded from the reference on 2-rank torch DDP: NCCL reduce-scatter /
    all-gather ("sharded"), this library's peer kernels over NVLink ("peer"), and the default ("auto")."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp

    mp.spawn(_sharded_rank_main, args=(2, name, (_free_port(), _free_port()), route), nprocs=2, join=True)
