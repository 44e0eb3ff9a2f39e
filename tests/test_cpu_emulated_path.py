"""The Python side of the path, end to end, in the CPU suite: the same trial code and fixtures as the GPU parity
tests, with libdgs.so's entry points replaced by the host-memory emulator of tests/fake_dgs.py (test infrastructure;
see its header).  What this checks is everything AROUND the kernels -- window bookkeeping, segment tables, range
tables, clip interception, step counters, the training loop, checkpoint save / restore, the AMP protocol, the
legacy entrypoint -- on every commit, without a GPU.  The kernels' own arithmetic is checked on the GPU."""
import copy
import math
import os
import pathlib
import sys

import numpy as np
import pytest
import torch

import fake_dgs
from golden_io import N_PARAMS, SCENARIOS, Golden, assert_close
from golden_model import GoldenNet, make_data, make_optimizer

import determined_clone_b200 as det
from determined_clone_b200 import pytorch

HERE = pathlib.Path(__file__).resolve().parent
CKPT = HERE / "golden" / "ckpt_adamw_a2"


def _np(t):
    return t.detach().float().cpu().numpy().copy()


class GoldenTrial(pytorch.PyTorchTrial):
    """tests/test_gpu_trial_parity.py's trial: the code that produced the fixtures under the reference harness."""

    def __init__(self, context, g, rec):
        self.context, self.g, self.rec = context, g, rec
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


FP32_ONE_SLOT = [n for n, sc in SCENARIOS.items() if sc["slots"] == 1]  # fp32 and the bf16 scenario


def _native_host_loop():
    import importlib.util

    from determined_clone_b200 import _build

    spec = importlib.util.spec_from_file_location("_dgs_host", str(_build.build_host()))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("name,native", [(n, False) for n in FP32_ONE_SLOT] + [("sgd_n1_a1", True), ("adamw_n1_a4_clip", True)])
def test_single_slot_scenarios_through_the_whole_python_path(monkeypatch, name, native):
    """The reference's recorded runs (10 optimizer steps each) reproduced by this harness with emulated kernels:
    synchronized gradients, clip norm, parameters, optimizer state, step counters, reported metrics.  native: with
    the C++ host loop (csrc/dgs_host.cpp, the default once built) taking the gradients instead of Python."""
    fake = fake_dgs.install(monkeypatch)
    if native:
        from determined_clone_b200.pytorch import _gradsync

        monkeypatch.setattr(_gradsync, "_host", _native_host_loop())
    g = Golden(name)
    rec = {"avg": [], "norm": [], "params": [], "state": []}
    reports = []
    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(hparams=g.hp, slots_per_trial=1, aggregation_frequency=g.agg, trial_seed=17)
    trial = GoldenTrial(ctx, g, rec)
    fused = ctx._engines[trial.opt].can_fuse()
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=g.steps * g.agg, reporting_period=g.agg,
                                          on_report=lambda n, r: reports.append((n, r)))
    ctl.run()
    assert fake.launches > 0
    kinds = {e[0] for e in fake.log}
    assert "flatten" in kinds and (("adam" in kinds or "sgd" in kinds) == fused)  # Adadelta steps in torch, on our gradient views
    opt_order_to_model = [0, 4, 1, 2, 3, 5]
    dt = g.dtype
    for s in range(g.steps):
        want_avg = g.per_step("avg_grads", s)
        for j, i in enumerate(opt_order_to_model):
            assert_close(rec["avg"][s][j], want_avg[i], dt, f"{name}: synced grad, step {s}, param {i}")
        for i in range(N_PARAMS):
            assert_close(rec["params"][s][i], g.per_step("params", s)[i], dt, f"{name}: parameter {i} after step {s}")
            st = rec["state"][s][i]
            if g.hp["optimizer"] == "sgd":
                assert_close(st["momentum_buffer"], g.state(s, i, "momentum_buffer"), dt, f"{name}: momentum {i} step {s}")
            elif g.hp["optimizer"] in ("adam", "adamw"):
                assert_close(st["exp_avg"], g.state(s, i, "exp_avg"), dt, f"{name}: exp_avg {i} step {s}")
                assert_close(st["exp_avg_sq"], g.state(s, i, "exp_avg_sq"), dt, f"{name}: exp_avg_sq {i} step {s}")
                assert float(st["step"]) == float(g.state(s, i, "step")) == s + 1
        if g.clip is not None and g.clip[0] == "norm":
            assert_close(float(rec["norm"][s]), g.z["total_norm"][s], dt, f"{name}: total norm step {s}")
    assert [n for n, _ in reports] == g.z["reported_steps"].tolist()
    for (n, r), want_avg, want_batches in zip(reports, g.z["reported_avg_loss"], g.z["reported_batch_loss"]):
        assert_close(r["avg_metrics"]["loss"], want_avg, dt, f"{name}: reported avg loss at {n}")
        assert_close([float(b["loss"]) for b in r["batch_metrics"]], want_batches, dt, f"{name}: per-batch loss at {n}")
    ctx.close()


# ---- checkpoint written by the reference -> restored here -> the reference's own continuation ----------------------
CK_HP = dict(optimizer="adamw", lr=3e-3, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1, global_batch_size=8)
CK_AGG, CK_SAVE_AT, CK_RESUME, CK_CLIP = 2, 3, 3, 1.0


class CkptTrial(pytorch.PyTorchTrial):
    def __init__(self, context, rec):
        self.context, self.rec = context, rec
        self.model = context.wrap_model(GoldenNet())
        self.plist = list(self.model.parameters())
        self.opt = context.wrap_optimizer(make_optimizer(self.model, CK_HP))
        x, y = make_data((CK_SAVE_AT + CK_RESUME) * CK_AGG, CK_HP["global_batch_size"])
        self.ds = torch.utils.data.TensorDataset(x, y)

    def build_training_data_loader(self):
        return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

    def build_validation_data_loader(self):
        return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

    def _clip(self, params):
        self.rec["norm"].append(float(torch.nn.utils.clip_grad_norm_(list(params), CK_CLIP)))

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x), y)
        self.context.backward(loss)
        self.rec["loss"].append(float(loss))
        self.rec["batch_idx"].append(batch_idx)
        self.context.step_optimizer(self.opt, clip_grads=self._clip)
        if (batch_idx + 1) % CK_AGG == 0:
            self.rec["params"].append([_np(p) for p in self.plist])
        return {"loss": loss}

    def evaluate_batch(self, batch):
        x, y = batch
        return {"val_loss": torch.nn.functional.mse_loss(self.model(x), y)}


def test_resume_from_the_reference_checkpoint_through_trainer_fit(monkeypatch, tmp_path):
    """Trainer.fit(latest_checkpoint=<the directory the reference wrote>): weights, AdamW state, counters and data
    position come back, training continues to the reference's own continuation, validation and the final
    checkpoint run on top (the combination the legacy entrypoint exercises on a GPU)."""
    fake_dgs.install(monkeypatch)
    z = np.load(CKPT / "continuation.npz")
    rec = {"norm": [], "loss": [], "batch_idx": [], "params": []}
    pytorch._PyTorchTrialController.pre_execute_hook(99)  # another seed: everything must come from the files
    ctx = pytorch.PyTorchTrialContext(hparams=CK_HP, slots_per_trial=1, aggregation_frequency=CK_AGG, trial_seed=99)
    trial = CkptTrial(ctx, rec)
    trainer = pytorch.Trainer(trial, ctx)
    assert ctx.get_trial_id() == 0  # == the reference's local-training trial id in trial_state.pkl
    total = (CK_SAVE_AT + CK_RESUME) * CK_AGG
    reports = trainer.fit(max_length=pytorch.Batch(total), reporting_period=pytorch.Batch(CK_AGG), validation_period=pytorch.Batch(4),
                          latest_checkpoint=str(CKPT), checkpoint_policy="all", checkpoint_dir=str(tmp_path))
    assert rec["batch_idx"] == z["batch_idx"].tolist() and [r["steps_completed"] for r in reports] == [8, 10, 12]
    assert_close(rec["loss"], z["loss"], "float32", "per-batch loss after resume")
    assert_close(rec["norm"], z["total_norm"], "float32", "clip norm after resume")
    for s in range(CK_RESUME):
        for i in range(N_PARAMS):
            assert_close(rec["params"][s][i], z[f"params/{s}/{i}"], "float32", f"parameter {i}, step {s} after resume")
    for i, p in enumerate(trial.plist):
        st = trial.opt.state[p]
        assert float(st["step"]) == float(z[f"state/{i}/step"]) == CK_SAVE_AT + CK_RESUME
        assert_close(_np(st["exp_avg"]), z[f"state/{i}/exp_avg"], "float32", f"exp_avg {i} after resume")
        assert_close(_np(st["exp_avg_sq"]), z[f"state/{i}/exp_avg_sq"], "float32", f"exp_avg_sq {i} after resume")
    assert [v["steps_completed"] for v in trainer.validation_reports] == [8, 12] and len(trainer.checkpoints) == 2
    # what we wrote at the end loads back into plain torch objects, as the reference's _load would do it
    saved = torch.load(tmp_path / trainer.checkpoints[-1] / "state_dict.pth", weights_only=False)
    plain = GoldenNet()
    plain.load_state_dict(saved["models_state_dict"][0])
    plain_opt = make_optimizer(plain, CK_HP)
    plain_opt.load_state_dict(saved["optimizers_state_dict"][0])
    for i, p in enumerate(plain.parameters()):
        np.testing.assert_array_equal(_np(p), rec["params"][-1][i])
        assert float(plain_opt.state[p]["step"]) == CK_SAVE_AT + CK_RESUME
    ctx.close()


# ---- AMP protocol (tests/test_gpu_amp.py on the CPU: torch.amp.GradScaler("cpu"), fp32 forward) --------------------
class Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(12, 32)
        self.norm = torch.nn.LayerNorm(32)
        self.b = torch.nn.Linear(32, 5)

    def forward(self, x):
        return self.b(torch.tanh(self.norm(self.a(x))))


def _amp_opt(kind, model):
    decay = [p for p in model.parameters() if p.ndim >= 2]
    rest = [p for p in model.parameters() if p.ndim < 2]
    groups = [{"params": decay, "weight_decay": 0.05}, {"params": rest, "weight_decay": 0.0}]
    if kind == "adamw":
        return torch.optim.AdamW(groups, lr=1e-2, betas=(0.9, 0.95))
    return torch.optim.SGD(groups, lr=0.05, momentum=0.9, nesterov=True)


def _amp_data(n, poison):
    g = torch.Generator().manual_seed(77)
    data = [(torch.randn(8, 12, generator=g), torch.randn(8, 5, generator=g)) for _ in range(n)]
    for i in poison:
        data[i][0][0, 0] = float("inf")
    return data


def _scaler():
    return torch.amp.GradScaler("cpu", init_scale=1024.0, growth_interval=2)


def _snap(model, opt, scaler, out):
    out["params"].append([_np(p) for p in model.parameters()])
    out["scale"].append(scaler.get_scale())
    out["adam_steps"].append([float(opt.state[p]["step"]) for p in model.parameters() if "step" in opt.state.get(p, {})])


def _same_run(ours, ref):
    assert ours["scale"] == ref["scale"] and ours["adam_steps"] == ref["adam_steps"] and len(ours["norms"]) == len(ref["norms"])
    for a, b in zip(ours["norms"], ref["norms"]):
        assert math.isfinite(a) == math.isfinite(b)
        if math.isfinite(b):
            np.testing.assert_allclose(a, b, rtol=1e-5)
    for s, (pa, pb) in enumerate(zip(ours["params"], ref["params"])):
        for a, b in zip(pa, pb):
            np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-7, err_msg=f"after optimizer step {s}")


@pytest.mark.parametrize("kind,agg,clip,poison", [
    ("adamw", 1, 0.5, (2,)), ("adamw", 2, 0.5, (5,)), ("adamw", 2, None, (2,)), ("sgd", 1, None, (0,)), ("sgd", 1, 0.5, ()),
])
def test_auto_amp_protocol(monkeypatch, kind, agg, clip, poison):
    fake = fake_dgs.install(monkeypatch)
    torch.manual_seed(5)
    base = Net()
    data = _amp_data(8 * agg, poison)

    model = copy.deepcopy(base)
    opt, scaler = _amp_opt(kind, model), _scaler()
    ref = {"params": [], "scale": [], "adam_steps": [], "norms": []}
    params = [p for grp in opt.param_groups for p in grp["params"]]
    for i, (x, y) in enumerate(data):
        scaler.scale(torch.nn.functional.mse_loss(model(x), y)).backward()
        if (i + 1) % agg:
            continue
        for p in params:
            p.grad.div_(float(agg))
        if clip is not None:
            scaler.unscale_(opt)
            ref["norms"].append(float(torch.nn.utils.clip_grad_norm_(params, clip)))
        scaler.step(opt)
        opt.zero_grad()
        scaler.update()
        _snap(model, opt, scaler, ref)

    out = {"params": [], "scale": [], "adam_steps": [], "norms": []}

    class AmpTrial(pytorch.PyTorchTrial):
        def __init__(self, context):
            self.context = context
            self.model = context.wrap_model(copy.deepcopy(base))
            context.experimental.use_amp()  # after wrap_model: loss scaling without the autocast wrapper
            context.wrap_scaler(_scaler())
            self.opt = context.wrap_optimizer(_amp_opt(kind, self.model))

        def build_training_data_loader(self):
            raise AssertionError("batches are fed by the test")

        def train_batch(self, batch, epoch_idx, batch_idx):
            loss = torch.nn.functional.mse_loss(self.model(batch[0]), batch[1])
            self.context.backward(loss)
            fn = (lambda ps: out["norms"].append(float(torch.nn.utils.clip_grad_norm_(list(ps), clip)))) if clip is not None else None
            self.context.step_optimizer(self.opt, clip_grads=fn)
            return {"loss": loss}

    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, aggregation_frequency=agg)
    trial = AmpTrial(ctx)
    assert ctx._prefolds(ctx._engines[trial.opt])
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=len(data), reporting_period=len(data))
    for i, batch in enumerate(data):
        ctx._current_batch_idx = i
        ctl._train_batch(batch, 0, i)
        if (i + 1) % agg == 0:
            _snap(trial.model, trial.opt, ctx._scaler, out)
    assert len(set(ref["scale"])) > 1
    _same_run(out, ref)
    # the loss scale went into the flatten call's factor: 1/(A*S) with the S in force for that window
    factors = [e[3] for e in fake.log if e[0] == "flatten" and e[2] & det._lib.WRITE_WIRE]
    scales_in_force = [1024.0] + ref["scale"][:-1]
    assert factors == [1.0 / (agg * s) for s in scales_in_force]
    ctx.close()


@pytest.mark.parametrize("kind", ["adamw", "sgd"])
@pytest.mark.parametrize("mode", ["no_clip", "clip_after_own_unscale", "clip_without_unscale"])
def test_scaler_passed_to_step_optimizer(monkeypatch, kind, mode):
    fake_dgs.install(monkeypatch)
    torch.manual_seed(8)
    base = Net()
    data = _amp_data(8, (3,))
    runs = []
    for which in ("ours", "torch"):
        model, scaler = copy.deepcopy(base), _scaler()
        out = {"params": [], "scale": [], "adam_steps": [], "norms": []}
        if which == "ours":
            ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8})
            model = ctx.wrap_model(model)
            opt = ctx.wrap_optimizer(_amp_opt(kind, model))
        else:
            opt = _amp_opt(kind, model)

        def clip_fn(params, opt=opt, scaler=scaler, out=out):
            if mode == "clip_after_own_unscale":
                scaler.unscale_(opt)
            out["norms"].append(float(torch.nn.utils.clip_grad_norm_(list(params), 0.5 if mode == "clip_after_own_unscale" else 300.0)))

        clip = None if mode == "no_clip" else clip_fn
        for i, (x, y) in enumerate(data):
            loss = torch.nn.functional.mse_loss(model(x), y)
            if which == "ours":
                ctx._current_batch_idx = i
                ctx.backward(scaler.scale(loss))
                ctx.step_optimizer(opt, clip_grads=clip, scaler=scaler)
            else:
                scaler.scale(loss).backward()
                if clip is not None:
                    clip([p for grp in opt.param_groups for p in grp["params"]])
                scaler.step(opt)
                opt.zero_grad()
            scaler.update()
            _snap(model, opt, scaler, out)
        if which == "ours":
            ctx.close()
        runs.append(out)
    _same_run(runs[0], runs[1])


# ---- the reference's closed-form experiment through the legacy entrypoint, in this process ---------------------------
def test_identity_experiment_from_its_yaml(monkeypatch, tmp_path):
    """tests/fixtures/pytorch_identity/const.yaml (16 batches of y = w*x, aggregation_frequency 2, reporting after
    every batch) run as `determined.exec.harness model_def:IdentityTrial` does it -- context from the experiment
    config, Trainer.fit with periods and policy from the YAML, final validation and checkpoint -- and compared with
    the reference's closed form (test_pytorch_trial.py:1604-1642).  tests/test_launch.py runs the same thing as a
    subprocess on a GPU."""
    import json

    from oracle import gradsync_oracle as O

    from determined_clone_b200.exec import harness

    fake_dgs.install(monkeypatch)
    fix = HERE / "fixtures" / "pytorch_identity"
    report = tmp_path / "reports.json"
    monkeypatch.setenv("DET_EXPERIMENT_CONFIG", str(fix / "const.yaml"))
    monkeypatch.setenv("DET_REPORT_FILE", str(report))
    monkeypatch.setenv("DET_CHECKPOINT_DIR", str(tmp_path / "ckpts"))
    monkeypatch.chdir(fix)
    monkeypatch.syspath_prepend(str(pathlib.Path(det.__file__).parent / "compat"))
    for name in [n for n in sys.modules if n == "determined" or n.startswith("determined.") or n == "model_def"]:
        monkeypatch.delitem(sys.modules, name)
    assert harness.main(["model_def:IdentityTrial"]) == 0
    reports = json.loads(report.read_text())
    assert [r["steps_completed"] for r in reports] == list(range(1, 17))
    # avg_metrics, computed when each report was made (one batch per report): on the CPU the per-batch entries are
    # numpy views of the live parameter, exactly as `metric.cpu().detach().numpy()` gives them in the reference
    got = [r["avg_metrics"]["weight"] for r in reports]
    np.testing.assert_allclose(got, O.identity_expected_weights()[:16], rtol=1e-5, atol=1e-9)
    assert got[0] == 0.0 and got[1] == got[2] != 0.0  # no update on non-boundary batches
    # searcher.metric validation_loss, checkpoint_policy best (the default): the end-of-training validation is the
    # best (only) one, so exactly one checkpoint exists, taken at batch 16
    (ckpt,) = list((tmp_path / "ckpts").iterdir())
    assert json.loads((ckpt / "metadata.json").read_text())["steps_completed"] == 16
    saved = torch.load(ckpt / "state_dict.pth", weights_only=False)
    np.testing.assert_allclose(float(saved["models_state_dict"][0]["weight"]), got[-1], rtol=1e-6)
    sys.modules.pop("model_def", None)


# ---- randomized differential test: engine bookkeeping vs the reference's lines over plain torch --------------------
class _Branchy(torch.nn.Module):
    """A trunk plus a side branch that only some batches use: its parameters get no gradient on the others."""

    def __init__(self, width, depth):
        super().__init__()
        self.inp = torch.nn.Linear(6, width)
        self.mid = torch.nn.ModuleList([torch.nn.Linear(width, width) for _ in range(depth)])
        self.norm = torch.nn.LayerNorm(width)
        self.side = torch.nn.Linear(width, width, bias=False)
        self.out = torch.nn.Linear(width, 3)

    def forward(self, x, use_side):
        h = torch.tanh(self.inp(x))
        for layer in self.mid:
            h = torch.tanh(layer(h)) + h
        h = self.norm(h)
        if use_side:
            h = h + self.side(h)
        return self.out(h)


def _random_optimizer(rng, model):
    decay = [p for p in model.parameters() if p.ndim >= 2]
    rest = [p for p in model.parameters() if p.ndim < 2]
    groups = [{"params": decay, "weight_decay": float(rng.choice([0.0, 0.01, 0.1])), "lr": float(rng.choice([0.01, 0.03]))},
              {"params": rest, "weight_decay": 0.0}]
    kind = rng.choice(["sgd", "sgd_plain", "adam", "adamw"])
    if kind == "sgd":
        return torch.optim.SGD(groups, lr=0.05, momentum=0.9, nesterov=bool(rng.integers(2)), dampening=0.0)
    if kind == "sgd_plain":
        return torch.optim.SGD(groups, lr=0.05, momentum=float(rng.choice([0.0, 0.5])), dampening=float(rng.choice([0.0, 0.2])))
    cls = torch.optim.Adam if kind == "adam" else torch.optim.AdamW
    return cls(groups, lr=0.01, betas=(0.9, float(rng.choice([0.95, 0.999]))), eps=1e-8)


@pytest.mark.parametrize("seed", range(12))
def test_random_trials_match_the_reference_lines(monkeypatch, seed):
    """Random model shape, optimizer kind and hyper-parameters per group, aggregation frequency, clipping by norm /
    by value / none, a StepLR stepped every optimizer step, and batches that leave some parameters without a
    gradient: this harness (emulated kernels) against oracle/torch_port.RefContext, which restates the reference's
    lines over torch's own optimizers.  Catches bookkeeping slips the fixed fixtures cannot: range tables with gaps,
    per-parameter Adam step counters drifting apart, momentum buffers created late, rates changing between steps."""
    from oracle import torch_port

    fake_dgs.install(monkeypatch)
    rng = np.random.default_rng(seed)
    torch.manual_seed(seed)
    base = _Branchy(int(rng.choice([8, 16, 24])), int(rng.integers(1, 4)))
    agg = int(rng.choice([1, 2, 3]))
    clip = [None, ("norm", 0.5), ("value", 0.05)][int(rng.integers(3))]
    steps = 6
    g = torch.Generator().manual_seed(seed)
    data = [(torch.randn(5, 6, generator=g), torch.randn(5, 3, generator=g), bool(rng.integers(2))) for _ in range(steps * agg)]
    # within one aggregation window the side branch is used by all micro-batches or by none (otherwise the reference
    # itself divides a partial sum by the full window length, which is fine, but then every step has every gradient)
    data = [(x, y, data[(i // agg) * agg][2]) for i, (x, y, _) in enumerate(data)]

    def clip_fn(params):
        params = [p for p in params if p.grad is not None]
        if clip and clip[0] == "norm":
            torch.nn.utils.clip_grad_norm_(params, clip[1])
        elif clip:
            torch.nn.utils.clip_grad_value_(params, clip[1])

    results = []
    for which in ("ours", "ref"):
        ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 5}, aggregation_frequency=agg) if which == "ours" \
            else torch_port.RefContext(aggregation_frequency=agg)
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(_random_optimizer(np.random.default_rng(seed + 100), model))
        sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.5)
        trace = []
        for i, (x, y, use_side) in enumerate(data):
            ctx._current_batch_idx = i
            ctx.backward(torch.nn.functional.mse_loss(model(x, use_side), y))
            ctx.step_optimizer(opt, clip_grads=clip_fn if clip else None)
            if (i + 1) % agg == 0:
                sched.step()
                trace.append([_np(p) for p in model.parameters()])
        state = {n: {k: (_np(v) if torch.is_tensor(v) and v.ndim else float(v)) for k, v in opt.state[p].items()}
                 for n, p in model.named_parameters() if p in opt.state}
        results.append((trace, state))
        if which == "ours":
            ctx.close()
    (ours, ours_state), (ref, ref_state) = results
    for s, (pa, pb) in enumerate(zip(ours, ref)):
        for a, b in zip(pa, pb):
            assert_close(a, b, "float32", f"seed {seed}: after optimizer step {s}")  # relative to the tensor's magnitude
    assert ours_state.keys() == ref_state.keys()
    for name in ref_state:
        assert ours_state[name].keys() == ref_state[name].keys(), name
        for k, want in ref_state[name].items():
            if k == "step":
                assert ours_state[name][k] == want, f"seed {seed}: {name}.step"  # integer counter: exact
            else:
                assert_close(ours_state[name][k], want, "float32", f"seed {seed}: {name}.{k}")


# ---- usage patterns beyond "one loss, one optimizer" ----------------------------------------------------------------
@pytest.mark.parametrize("agg,stray", [(1, "clear"), (1, "freeze"), (2, "freeze"), (1, "keep")])
def test_two_optimizers_two_losses_and_two_backward_passes_per_step(monkeypatch, agg, stray):
    """A GAN-shaped trial: two models with an optimizer each; the first optimizer's step sees the sum of TWO
    backward passes per batch (backward_passes_per_step=2, reference _pytorch_context.py:313-378), the second's loss
    flows through the first model too.  What happens to the gradients that pass leaves on the first model is the
    trial's business, as in the reference: "clear" them afterwards (they must then not show up in the next window),
    "freeze" the first model during the pass (none arise), or "keep" them (they add to the next window's gradient).
    Against the reference's lines over plain torch."""
    from oracle import torch_port

    fake_dgs.install(monkeypatch)
    torch.manual_seed(3)
    gen0, disc0 = torch.nn.Linear(4, 6), torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 1))
    g = torch.Generator().manual_seed(3)
    data = [torch.randn(7, 4, generator=g) for _ in range(4 * agg)]
    results = []
    for which in ("ours", "ref"):
        ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 7}, aggregation_frequency=agg) if which == "ours" \
            else torch_port.RefContext(aggregation_frequency=agg)
        gen, disc = ctx.wrap_model(copy.deepcopy(gen0)), ctx.wrap_model(copy.deepcopy(disc0))
        opt_d = ctx.wrap_optimizer(torch.optim.Adam(disc.parameters(), lr=0.01), backward_passes_per_step=2)
        opt_g = ctx.wrap_optimizer(torch.optim.SGD(gen.parameters(), lr=0.05, momentum=0.9))
        trace = []
        for i, z in enumerate(data):
            ctx._current_batch_idx = i
            fake_sample = gen(z).detach()
            ctx.backward(disc(fake_sample).mean())          # discriminator: two losses, two backward passes
            ctx.backward((disc(z @ torch.ones(4, 6)) - 1).pow(2).mean())
            ctx.step_optimizer(opt_d)
            for p in disc.parameters():
                p.requires_grad_(stray != "freeze")
            ctx.backward(-disc(gen(z)).mean())               # generator: through the discriminator
            ctx.step_optimizer(opt_g)
            for p in disc.parameters():
                p.requires_grad_(True)
                if stray == "clear":                         # the generator pass left gradients on the discriminator
                    p.grad = None
            if (i + 1) % agg == 0:
                trace.append([_np(p) for p in list(gen.parameters()) + list(disc.parameters())])
        results.append(trace)
        if which == "ours":
            ctx.close()
    for s, (pa, pb) in enumerate(zip(*results)):
        for a, b in zip(pa, pb):
            assert_close(a, b, "float32", f"after optimizer step {s}")


def test_gradients_stay_visible_without_auto_zero_grads(monkeypatch):
    """step_optimizer(auto_zero_grads=False): after the step p.grad holds the synchronized gradients the step used
    (reference :923-924 simply does not call zero_grad), until the trial clears them itself."""
    fake_dgs.install(monkeypatch)
    torch.manual_seed(4)
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4})
    model = ctx.wrap_model(torch.nn.Linear(3, 2))
    twin = copy.deepcopy(model)
    opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1))
    ref_opt = torch.optim.SGD(twin.parameters(), lr=0.1)
    x = torch.randn(4, 3)
    for i in range(3):
        ctx._current_batch_idx = i
        ctx.backward(model(x).pow(2).mean())
        ctx.step_optimizer(opt, auto_zero_grads=False)
        twin(x).pow(2).mean().backward()
        ref_opt.step()
        for p, q in zip(model.parameters(), twin.parameters()):
            assert p.grad is not None
            assert_close(_np(p.grad), _np(q.grad), "float32", f"gradient left on the parameter, step {i}")
            assert_close(_np(p), _np(q), "float32", f"parameter, step {i}")
        for p in list(model.parameters()) + list(twin.parameters()):
            p.grad = None
    with pytest.raises(det.errors.InvalidExperimentException, match="auto_zero_grads must be set to true"):
        ctx2 = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4}, aggregation_frequency=2)
        m2 = ctx2.wrap_model(torch.nn.Linear(3, 2))
        o2 = ctx2.wrap_optimizer(torch.optim.SGD(m2.parameters(), lr=0.1))
        ctx2._current_batch_idx = 0
        ctx2.step_optimizer(o2, auto_zero_grads=False)
    ctx.close()


def test_optimizer_membership_changes_after_wrapping_fail_loudly(monkeypatch):
    """The flat buffers are laid out at wrap_optimizer(): a parameter group added later, or a parameter unfrozen
    later, would silently never be synchronized or stepped -- both are refused when the next step finds them.
    Parameters that stay frozen, gradients from a plain loss.backward() and steps without any gradient are fine."""
    fake_dgs.install(monkeypatch)
    torch.manual_seed(0)
    x = torch.randn(4, 3)

    def make(freeze_second=False, both=False):
        ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4})
        m = ctx.wrap_model(torch.nn.Sequential(torch.nn.Linear(3, 4), torch.nn.Linear(4, 2)))
        if freeze_second:
            for p in m[1].parameters():
                p.requires_grad_(False)
        params = list(m.parameters()) if both or freeze_second else list(m[0].parameters())
        return ctx, m, ctx.wrap_optimizer(torch.optim.SGD(params, lr=0.1, momentum=0.9))

    ctx, m, opt = make()
    w0 = m[0].weight.detach().clone()
    ctx._current_batch_idx = 0
    m(x).pow(2).mean().backward()            # not through context.backward(): picked up by the step
    ctx.step_optimizer(opt)
    assert not torch.equal(w0, m[0].weight.detach()) and m[0].weight.grad is None
    ctx._current_batch_idx = 1
    w1 = m[0].weight.detach().clone()
    ctx.step_optimizer(opt)                  # no gradient at all: nothing moves (torch's step skips such parameters)
    assert torch.equal(w1, m[0].weight.detach())
    opt.add_param_group({"params": list(m[1].parameters())})
    ctx._current_batch_idx = 2
    ctx.backward(m(x).pow(2).mean())
    with pytest.raises(det.errors.InvalidExperimentException, match="parameters changed after wrap_optimizer"):
        ctx.step_optimizer(opt)
    ctx.close()

    ctx, m, opt = make(freeze_second=True)   # frozen parameters in the optimizer are fine as long as they stay frozen
    for i in range(2):
        ctx._current_batch_idx = i
        v0 = m[1].weight.detach().clone()
        ctx.backward(m(x).pow(2).mean())
        ctx.step_optimizer(opt)
        assert torch.equal(v0, m[1].weight.detach())
    for p in m[1].parameters():
        p.requires_grad_(True)
    ctx._current_batch_idx = 2
    ctx.backward(m(x).pow(2).mean())
    with pytest.raises(det.errors.InvalidExperimentException, match="requires_grad=False when its optimizer was wrapped"):
        ctx.step_optimizer(opt)
    ctx.close()


_OPTIMIZER_VARIANTS = {
    # name: (constructor, does the fused step stand in for optimizer.step()?)
    "adam_maximize": (lambda ps: torch.optim.Adam(ps, lr=0.01, maximize=True), True),
    "adam_decoupled_wd": (lambda ps: torch.optim.Adam(ps, lr=0.01, weight_decay=0.1, decoupled_weight_decay=True), True),
    "adam_l2_wd": (lambda ps: torch.optim.Adam(ps, lr=0.01, weight_decay=0.1), True),
    "adam_foreach_off": (lambda ps: torch.optim.Adam(ps, lr=0.01, foreach=False), True),
    "sgd_maximize": (lambda ps: torch.optim.SGD(ps, lr=0.01, momentum=0.9, maximize=True), True),
    "sgd_dampening": (lambda ps: torch.optim.SGD(ps, lr=0.01, momentum=0.9, dampening=0.3), True),
    "sgd_plain": (lambda ps: torch.optim.SGD(ps, lr=0.01), True),
    # what the kernels do not implement keeps torch's own step, on our gradient views
    "adam_amsgrad": (lambda ps: torch.optim.Adam(ps, lr=0.01, amsgrad=True), False),
    "adamw_amsgrad": (lambda ps: torch.optim.AdamW(ps, lr=0.01, amsgrad=True), False),
    "adam_tensor_lr": (lambda ps: torch.optim.Adam(ps, lr=torch.tensor(0.01)), False),
    "rmsprop": (lambda ps: torch.optim.RMSprop(ps, lr=0.01), False),
    "nadam": (lambda ps: torch.optim.NAdam(ps, lr=0.01), False),
}


@pytest.mark.parametrize("name", sorted(_OPTIMIZER_VARIANTS))
def test_optimizer_options_are_fused_or_left_to_torch(monkeypatch, name):
    """Every optimizer option either is implemented by the step entry points (same trajectory as torch's own step)
    or makes the engine leave optimizer.step() to torch -- never a fused step that silently ignores the option."""
    make, want_fused = _OPTIMIZER_VARIANTS[name]
    fake = fake_dgs.install(monkeypatch)
    torch.manual_seed(0)
    base = torch.nn.Sequential(torch.nn.Linear(3, 4), torch.nn.Tanh(), torch.nn.Linear(4, 2))
    x, y = torch.randn(6, 3), torch.randn(6, 2)
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 6})
    model = ctx.wrap_model(copy.deepcopy(base))
    opt = ctx.wrap_optimizer(make(model.parameters()))
    twin = copy.deepcopy(base)
    twin_opt = make(twin.parameters())
    assert ctx._engines[opt].can_fuse() == want_fused
    for i in range(5):
        ctx._current_batch_idx = i
        ctx.backward(torch.nn.functional.mse_loss(model(x), y))
        ctx.step_optimizer(opt)
        torch.nn.functional.mse_loss(twin(x), y).backward()
        twin_opt.step()
        twin_opt.zero_grad()
        for p, q in zip(model.parameters(), twin.parameters()):
            assert_close(_np(p), _np(q), "float32", f"{name}: step {i}")
    kinds = {e[0] for e in fake.log}
    # fused optimizers on one slot take the single-launch step (flatten + step in one entry point); the others keep
    # flatten -> torch's own step
    assert (bool(kinds & {"direct_adam", "direct_sgd"}) == want_fused) and (("flatten" in kinds) != want_fused)
    ctx.close()


def test_parameters_moved_out_of_the_flat_buffer_are_noticed(monkeypatch):
    """Parameters are views of one flat buffer per optimizer; rebinding their .data after wrap_optimizer() (model.to,
    .half(), p.data = ...) would leave the step updating memory the model no longer reads.  The first and last
    parameter of every buffer are checked at each step: whole-model moves are caught, in-place changes are fine,
    and a channels_last model trains to the same weights (its parameters become plain contiguous views)."""
    fake_dgs.install(monkeypatch)
    torch.manual_seed(0)
    base = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(8, 4, 3, padding=1))
    base = base.to(memory_format=torch.channels_last)
    x = torch.randn(2, 3, 8, 8).to(memory_format=torch.channels_last)
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 2})
    model = ctx.wrap_model(copy.deepcopy(base))
    opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9))
    twin = copy.deepcopy(base)
    twin_opt = torch.optim.SGD(twin.parameters(), lr=0.1, momentum=0.9)
    for i in range(3):
        ctx._current_batch_idx = i
        ctx.backward(model(x).pow(2).mean())  # conv weight gradients arrive in channels_last strides
        ctx.step_optimizer(opt)
        twin(x).pow(2).mean().backward()
        twin_opt.step()
        twin_opt.zero_grad()
    for p, q in zip(model.parameters(), twin.parameters()):
        assert_close(_np(p), _np(q), "float32", "channels_last model")
    with torch.no_grad():
        model[0].weight.copy_(twin[0].weight * 0.5)  # in place: still the buffer's memory
    ctx._current_batch_idx = 3
    ctx.backward(model(x).pow(2).mean())
    ctx.step_optimizer(opt)
    model.double()  # rebinds every parameter's .data
    ctx._current_batch_idx = 4
    with pytest.raises(det.errors.InvalidExperimentException, match="no longer lives in the optimizer's flat parameter buffer"):
        ctx.step_optimizer(opt)
    ctx.close()


def test_cleared_adam_state_restarts_from_step_zero_with_zero_moments(monkeypatch):
    """optimizer.state.clear() (or load_state_dict of an empty state) between steps: torch starts over -- step 1, zero
    moments -- and so must the fused step, whose flat exp_avg / exp_avg_sq buffers still hold the old moments."""
    fake_dgs.install(monkeypatch)
    torch.manual_seed(3)
    base = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 2))
    x, y = torch.randn(9, 5), torch.randn(9, 2)
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 9})
    model = ctx.wrap_model(copy.deepcopy(base))
    opt = ctx.wrap_optimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.1))
    twin = copy.deepcopy(base)
    twin_opt = torch.optim.AdamW(twin.parameters(), lr=1e-2, weight_decay=0.1)
    for i in range(6):
        if i == 3:
            opt.state.clear()
            twin_opt.state.clear()
        ctx._current_batch_idx = i
        ctx.backward(torch.nn.functional.mse_loss(model(x), y))
        ctx.step_optimizer(opt)
        torch.nn.functional.mse_loss(twin(x), y).backward()
        twin_opt.step()
        twin_opt.zero_grad()
        for p, q in zip(model.parameters(), twin.parameters()):
            assert_close(_np(p), _np(q), "float32", f"step {i}")
            assert float(opt.state[p]["step"]) == float(twin_opt.state[q]["step"]) == (i % 3) + 1
            assert_close(_np(opt.state[p]["exp_avg_sq"]), _np(twin_opt.state[q]["exp_avg_sq"]), "float32", f"exp_avg_sq step {i}")
    ctx.close()
