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
        loss = torch.nn.functional.mse_loss(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.opt, clip_grads=self._clip)
        if (batch_idx + 1) % self.g.agg == 0:
            self.rec["params"].append([_np(p) for p in self.plist])
            self.rec["state"].append([{k: (_np(v) if torch.is_tensor(v) else v) for k, v in self.opt.state[p].items()} for p in self.plist])
        return {"loss": loss}


FP32_ONE_SLOT = [n for n, sc in SCENARIOS.items() if sc["slots"] == 1 and sc["dtype"] == "float32"]


@pytest.mark.parametrize("name", FP32_ONE_SLOT)
def test_single_slot_scenarios_through_the_whole_python_path(monkeypatch, name):
    """The reference's recorded runs (10 optimizer steps each) reproduced by this harness with emulated kernels:
    synchronized gradients, clip norm, parameters, optimizer state, step counters, reported metrics."""
    fake = fake_dgs.install(monkeypatch)
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
    for s in range(g.steps):
        want_avg = g.per_step("avg_grads", s)
        for j, i in enumerate(opt_order_to_model):
            assert_close(rec["avg"][s][j], want_avg[i], "float32", f"{name}: synced grad, step {s}, param {i}")
        for i in range(N_PARAMS):
            assert_close(rec["params"][s][i], g.per_step("params", s)[i], "float32", f"{name}: parameter {i} after step {s}")
            st = rec["state"][s][i]
            if g.hp["optimizer"] == "sgd":
                assert_close(st["momentum_buffer"], g.state(s, i, "momentum_buffer"), "float32", f"{name}: momentum {i} step {s}")
            elif g.hp["optimizer"] in ("adam", "adamw"):
                assert_close(st["exp_avg"], g.state(s, i, "exp_avg"), "float32", f"{name}: exp_avg {i} step {s}")
                assert_close(st["exp_avg_sq"], g.state(s, i, "exp_avg_sq"), "float32", f"{name}: exp_avg_sq {i} step {s}")
                assert float(st["step"]) == float(g.state(s, i, "step")) == s + 1
        if g.clip is not None and g.clip[0] == "norm":
            assert_close(float(rec["norm"][s]), g.z["total_norm"][s], "float32", f"{name}: total norm step {s}")
    assert [n for n, _ in reports] == g.z["reported_steps"].tolist()
    for (n, r), want_avg, want_batches in zip(reports, g.z["reported_avg_loss"], g.z["reported_batch_loss"]):
        assert_close(r["avg_metrics"]["loss"], want_avg, "float32", f"{name}: reported avg loss at {n}")
        assert_close([float(b["loss"]) for b in r["batch_metrics"]], want_batches, "float32", f"{name}: per-batch loss at {n}")
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
