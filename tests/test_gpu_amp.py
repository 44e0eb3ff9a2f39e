"""GradScaler / AMP on the fused path.

The reference's protocol (harness/determined/pytorch/_pytorch_context.py:798-799 scale the loss, :818-825 average
over the window, :899-911 unscale_ before the user's clip function, :906-921 scaler.step, _pytorch_trial.py:903-910
scaler.update at every aggregation boundary) is written out below with plain torch objects and compared with this
repo's context, which folds 1/S into its kernels instead of running GradScaler.unscale_ as a pass of its own.
Scale values and Adam's step counters are integers / powers of two: exact.  Parameters: fp32 tolerance of the
other parity tests."""
import contextlib
import copy
import math

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import determined_clone_b200 as det  # noqa: E402
from determined_clone_b200 import pytorch  # noqa: E402


class Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(12, 32)
        self.norm = torch.nn.LayerNorm(32)
        self.b = torch.nn.Linear(32, 5)

    def forward(self, x):
        return self.b(torch.tanh(self.norm(self.a(x))))


def make_opt(kind, model):
    decay = [p for p in model.parameters() if p.ndim >= 2]
    rest = [p for p in model.parameters() if p.ndim < 2]
    groups = [{"params": decay, "weight_decay": 0.05}, {"params": rest, "weight_decay": 0.0}]
    if kind == "adamw":
        return torch.optim.AdamW(groups, lr=1e-2, betas=(0.9, 0.95))
    return torch.optim.SGD(groups, lr=0.05, momentum=0.9, nesterov=True)


def make_data(n, poison):
    g = torch.Generator().manual_seed(77)
    data = [(torch.randn(8, 12, generator=g).cuda(), torch.randn(8, 5, generator=g).cuda()) for _ in range(n)]
    for i in poison:
        data[i][0][0, 0] = float("inf")  # -> non-finite loss -> non-finite gradients -> the step must be skipped
    return data


def new_scaler():
    return torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=2)


def snapshot(model, opt, scaler, out):
    out["params"].append([p.detach().cpu().numpy().copy() for p in model.parameters()])
    out["scale"].append(scaler.get_scale())
    steps = [float(opt.state[p]["step"]) for p in model.parameters() if "step" in opt.state.get(p, {})]
    out["adam_steps"].append(steps)


def reference_auto_amp(base, kind, data, agg, clip):
    """What the reference harness does with experimental.use_amp() (fp32 forward: use_amp after wrap_model)."""
    model = copy.deepcopy(base).cuda()
    opt = make_opt(kind, model)
    scaler = new_scaler()
    out = {"params": [], "scale": [], "adam_steps": [], "norms": []}
    params = [p for g in opt.param_groups for p in g["params"]]
    for i, (x, y) in enumerate(data):
        loss = torch.nn.functional.mse_loss(model(x), y)
        scaler.scale(loss).backward()
        if (i + 1) % agg:
            continue
        for p in params:
            p.grad.div_(float(agg))
        if clip is not None:
            scaler.unscale_(opt)
            out["norms"].append(float(torch.nn.utils.clip_grad_norm_(params, clip)))
        scaler.step(opt)
        opt.zero_grad()
        scaler.update()
        snapshot(model, opt, scaler, out)
    return out


class AmpTrial(pytorch.PyTorchTrial):
    def __init__(self, context, base, kind, clip, out, autocast=False):
        self.context, self.clip, self.out = context, clip, out
        if autocast:
            context.experimental.use_amp()
        self.model = context.wrap_model(copy.deepcopy(base))
        if not autocast:
            context.experimental.use_amp()  # after wrap_model: loss scaling without the autocast wrapper
        context._scaler = context.wrap_scaler(new_scaler())
        self.opt = context.wrap_optimizer(make_opt(kind, self.model))

    def build_training_data_loader(self):
        raise AssertionError("batches are fed by the test")

    def _clip(self, params):
        self.out["norms"].append(torch.nn.utils.clip_grad_norm_(list(params), self.clip))

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.opt, clip_grads=self._clip if self.clip is not None else None)
        return {"loss": loss}


def ours_auto_amp(base, kind, data, agg, clip, autocast=False):
    out = {"params": [], "scale": [], "adam_steps": [], "norms": []}
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, aggregation_frequency=agg, device=torch.device("cuda", 0))
    trial = AmpTrial(ctx, base, kind, clip, out, autocast=autocast)
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=len(data), reporting_period=len(data))
    launches0 = det._lib.calls.dgs_launch_count()
    for i, batch in enumerate(data):
        ctx._current_batch_idx = i
        ctl._train_batch(batch, 0, i)
        if (i + 1) % agg == 0:
            snapshot(trial.model, trial.opt, ctx._scaler, out)
    assert det._lib.calls.dgs_launch_count() > launches0, "the CUDA extension did not run"
    out["norms"] = [float(v) for v in out["norms"]]
    ctx.close()
    return out, trial


def assert_same_run(ours, ref, rtol=1e-5, atol=1e-7):
    assert ours["scale"] == ref["scale"]            # powers of two: exact
    assert ours["adam_steps"] == ref["adam_steps"]  # a skipped step must not advance Adam's counters
    assert len(ours["norms"]) == len(ref["norms"])
    for a, b in zip(ours["norms"], ref["norms"]):
        assert (math.isfinite(a), math.isfinite(b)) in ((True, True), (False, False))
        if math.isfinite(b):
            np.testing.assert_allclose(a, b, rtol=1e-5)
    for s, (pa, pb) in enumerate(zip(ours["params"], ref["params"])):
        for a, b in zip(pa, pb):
            np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, err_msg=f"after optimizer step {s}")


@pytest.mark.parametrize("kind,agg,clip,poison", [
    ("adamw", 1, 0.5, (2,)),        # inf in the 3rd step: skipped, scale backs off, Adam's step stays
    ("adamw", 2, 0.5, (5,)),        # with accumulation; the poisoned micro-batch is the boundary one
    ("adamw", 2, None, (2,)),       # poisoned micro-batch inside the window, no clipping
    ("sgd", 1, None, (0,)),         # very first step skipped: the momentum buffer is initialised one step later
    ("sgd", 1, 0.5, ()),            # nothing skipped: scale grows every 2 steps
])
def test_auto_amp_matches_the_reference_protocol(lib, kind, agg, clip, poison):
    torch.manual_seed(5)
    base = Net()
    data = make_data(8 * agg, poison)
    ref = reference_auto_amp(base, kind, data, agg, clip)
    ours, _ = ours_auto_amp(base, kind, data, agg, clip)
    assert len(set(ref["scale"])) > 1  # the scenario does exercise growth and/or backoff
    assert_same_run(ours, ref)


def test_use_amp_autocasts_the_forward_pass(lib):
    """experimental.use_amp() before wrap_model: forward under fp16 autocast, outputs handed back as fp32, the
    module object and its attributes untouched; the run tracks the plain-torch autocast run."""
    torch.manual_seed(6)
    base = Net()
    data = make_data(6, ())
    seen = []
    ours, trial = None, None
    out = {"params": [], "scale": [], "adam_steps": [], "norms": []}
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, device=torch.device("cuda", 0))
    trial = AmpTrial(ctx, base, "adamw", 1.0, out, autocast=True)
    assert isinstance(trial.model, Net) and trial.model.a.weight.dtype == torch.float32
    assert list(trial.model.state_dict()) == list(base.state_dict())
    trial.model.a.register_forward_hook(lambda m, i, o: seen.append(o.dtype))
    y = trial.model(data[0][0])
    assert seen == [torch.float16] and y.dtype == torch.float32
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=len(data), reporting_period=len(data))
    for i, batch in enumerate(data):
        ctx._current_batch_idx = i
        ctl._train_batch(batch, 0, i)
        snapshot(trial.model, trial.opt, ctx._scaler, out)
    ctx.close()

    model = copy.deepcopy(base).cuda()
    opt, scaler = make_opt("adamw", model), new_scaler()
    ref = {"params": [], "scale": [], "adam_steps": [], "norms": []}
    for x, yy in data:
        with torch.autocast("cuda"):
            pred = model(x)
        loss = torch.nn.functional.mse_loss(pred.float(), yy)
        scaler.scale(loss).backward()
        scaler.unscale_(opt)
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        scaler.step(opt)
        opt.zero_grad()
        scaler.update()
        snapshot(model, opt, scaler, ref)
    assert out["scale"] == ref["scale"] and out["adam_steps"] == ref["adam_steps"]
    for pa, pb in zip(out["params"], ref["params"]):
        for a, b in zip(pa, pb):
            # fp16 forward: a last-bit difference in a weight can flip an fp16 rounding downstream
            np.testing.assert_allclose(a, b, rtol=2e-2, atol=2e-3)


@pytest.mark.parametrize("kind", ["adamw", "sgd"])
@pytest.mark.parametrize("mode", ["no_clip", "clip_after_own_unscale", "clip_without_unscale"])
def test_scaler_passed_to_step_optimizer(lib, kind, mode):
    """Manual AMP (reference :906-911 `scaler.step(optimizer)` with the scaler the user passes): the user scales the
    loss; gradients reach step_optimizer still scaled."""
    torch.manual_seed(8)
    base = Net()
    data = make_data(8, (3,))
    runs = []
    for which in ("ours", "torch"):
        model = copy.deepcopy(base).cuda()
        scaler = new_scaler()
        out = {"params": [], "scale": [], "adam_steps": [], "norms": []}
        if which == "ours":
            ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, device=torch.device("cuda", 0))
            model = ctx.wrap_model(model)
            opt = ctx.wrap_optimizer(make_opt(kind, model))
        else:
            opt = make_opt(kind, model)

        def clip_fn(params, opt=opt, scaler=scaler, out=out):
            if mode == "clip_after_own_unscale":
                scaler.unscale_(opt)
            out["norms"].append(torch.nn.utils.clip_grad_norm_(list(params), 0.5 if mode == "clip_after_own_unscale" else 300.0))

        clip = None if mode == "no_clip" else clip_fn
        launches0 = det._lib.calls.dgs_launch_count()
        for i, (x, y) in enumerate(data):
            loss = torch.nn.functional.mse_loss(model(x), y)
            if which == "ours":
                ctx._current_batch_idx = i
                ctx.backward(scaler.scale(loss))
                ctx.step_optimizer(opt, clip_grads=clip, scaler=scaler)
            else:
                scaler.scale(loss).backward()
                if clip is not None:
                    clip([p for g in opt.param_groups for p in g["params"]])
                scaler.step(opt)
                opt.zero_grad()
            scaler.update()
            snapshot(model, opt, scaler, out)
        if which == "ours":
            assert det._lib.calls.dgs_launch_count() > launches0
            ctx.close()
        out["norms"] = [float(v) for v in out["norms"]]
        runs.append(out)
    assert_same_run(runs[0], runs[1])
    assert min(runs[1]["scale"]) < 1024.0 or max(runs[1]["scale"]) > 1024.0


# ---- two slots: an inf on ONE rank must make EVERY rank skip (the check runs on the all-reduced gradients) ---------
def _free_port() -> int:
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _amp_rank_main(rank: int, size: int, ports, init_port: int) -> None:
    import torch.distributed as dist

    from determined_clone_b200 import core

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{init_port}", rank=rank, world_size=size,
                            device_id=torch.device("cuda", rank))
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=ports[0], pull_port=ports[1])
    try:
        agg, clip = 2, 0.5
        torch.manual_seed(5)
        base = Net()
        g = torch.Generator().manual_seed(300 + rank)  # every rank trains on its own batches
        data = [(torch.randn(8, 12, generator=g).cuda(), torch.randn(8, 5, generator=g).cuda()) for _ in range(8 * agg)]
        if rank == 1:
            data[4][0][0, 0] = float("inf")  # first micro-batch of the third window, on rank 1 only

        # reference protocol over torch DDP (reference _pytorch_context.py:295-296, :802-816 no_sync off the boundary)
        ref_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base).cuda())
        ref_opt, ref_scaler = make_opt("adamw", ref_model), new_scaler()
        ref = {"params": [], "scale": [], "adam_steps": [], "norms": []}
        params = [p for grp in ref_opt.param_groups for p in grp["params"]]
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
            snapshot(ref_model, ref_opt, ref_scaler, ref)

        out = {"params": [], "scale": [], "adam_steps": [], "norms": []}
        ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 16}, slots_per_trial=size,
                                          aggregation_frequency=agg)
        trial = AmpTrial(ctx, base, "adamw", clip, out)
        ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=len(data), reporting_period=len(data))
        for i, batch in enumerate(data):
            ctx._current_batch_idx = i
            ctl._train_batch(batch, 0, i)
            if (i + 1) % agg == 0:
                snapshot(trial.model, trial.opt, ctx._scaler, out)
        out["norms"] = [float(v) for v in out["norms"]]
        assert ref["scale"][:3] == [1024.0, 2048.0, 1024.0] and ref["adam_steps"][2] == ref["adam_steps"][1]  # 3rd step skipped
        assert_same_run(out, ref)
        ctx.close()
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_two_slot_auto_amp_matches_torch_ddp_with_grad_scaler(lib):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp

    mp.spawn(_amp_rank_main, args=(2, (_free_port(), _free_port()), _free_port()), nprocs=2, join=True)
