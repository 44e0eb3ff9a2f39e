"""Checkpoint interop with the REFERENCE harness.

tests/golden/ckpt_adamw_a2/ holds a checkpoint written by the unmodified reference controller's _save after 3
optimizer steps, and continuation.npz holds what the reference itself did after restoring it (see
tests/golden/make_golden_ckpt.py).  Here the same files are restored by this repo's controller onto the GPU
and training continues through the CUDA path: batch positions exactly, losses / clip norms / parameters /
optimizer state within the fp32 tolerance of the other parity tests."""
import pathlib
import pickle

import numpy as np
import pytest

from golden_io import assert_close
from golden_model import GoldenNet, make_data, make_optimizer

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import determined_clone_b200 as det  # noqa: E402
from determined_clone_b200 import pytorch  # noqa: E402

CKPT = pathlib.Path(__file__).parent / "golden" / "ckpt_adamw_a2"
# the generator's constants (make_golden_ckpt.py)
HP = dict(optimizer="adamw", lr=3e-3, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1, global_batch_size=8)
AGG, SAVE_AT, RESUME_STEPS, CLIP = 2, 3, 3, 1.0


def _np(t):
    return t.detach().float().cpu().numpy().copy()


class CkptTrial(pytorch.PyTorchTrial):
    def __init__(self, context, rec):
        self.context, self.rec = context, rec
        self.model = context.wrap_model(GoldenNet())  # fresh random init: everything must come from the file
        self.plist = list(self.model.parameters())
        self.opt = context.wrap_optimizer(make_optimizer(self.model, HP))
        x, y = make_data((SAVE_AT + RESUME_STEPS) * AGG, HP["global_batch_size"])
        self.ds = torch.utils.data.TensorDataset(x, y)

    def build_training_data_loader(self):
        return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

    def _clip(self, params):
        self.rec["norm"].append(torch.nn.utils.clip_grad_norm_(list(params), CLIP))

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x), y)
        self.context.backward(loss)
        self.rec["loss"].append(loss.detach())
        self.rec["batch_idx"].append(batch_idx)
        self.context.step_optimizer(self.opt, clip_grads=self._clip)
        if (batch_idx + 1) % AGG == 0:
            self.rec["params"].append([_np(p) for p in self.plist])
        return {"loss": loss}


def _resume(rec, path, max_batches):
    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(hparams=HP, slots_per_trial=1, aggregation_frequency=AGG, trial_seed=17)
    trial = CkptTrial(ctx, rec)
    # the reference's local-training controller that wrote the file has trial_id 0 (_pytorch_trial.py:225-226)
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=max_batches, reporting_period=AGG, trial_id=0)
    ctl._load(path)
    return ctx, trial, ctl


def test_resume_from_a_checkpoint_the_reference_wrote(lib, tmp_path):
    z = np.load(CKPT / "continuation.npz")
    rec = {"norm": [], "loss": [], "batch_idx": [], "params": []}
    ctx, trial, ctl = _resume(rec, CKPT, (SAVE_AT + RESUME_STEPS) * AGG)
    # counters restored from the reference's trial_state.pkl: exact
    assert ctl.batches_trained == ctl.state.batches_trained == SAVE_AT * AGG
    assert (ctl.state.trial_id, ctl.state.last_ckpt, ctl.state.epochs_trained) == (0, SAVE_AT * AGG, 0)
    np.testing.assert_array_equal(np.concatenate([_np(p).ravel() for p in trial.plist]), z["saved_params_check"])
    launches0 = det._lib.calls.dgs_launch_count()
    ctl.run()
    torch.cuda.synchronize()
    assert det._lib.calls.dgs_launch_count() > launches0, "the CUDA extension did not run"

    assert rec["batch_idx"] == z["batch_idx"].tolist()  # the loader resumed at the same batch: exact
    assert ctl.batches_trained == int(z["batches_trained"])
    assert_close([float(v) for v in rec["loss"]], z["loss"], "float32", "per-batch loss after resume")
    assert_close([float(v) for v in rec["norm"]], z["total_norm"], "float32", "clip norm after resume")
    n_params = len(trial.plist)
    for s in range(RESUME_STEPS):
        for i in range(n_params):
            assert_close(rec["params"][s][i], z[f"params/{s}/{i}"], "float32", f"parameter {i}, step {s} after resume")
    for i, p in enumerate(trial.plist):
        st = trial.opt.state[p]
        assert float(st["step"]) == float(z[f"state/{i}/step"]) == SAVE_AT + RESUME_STEPS  # integer counter: exact
        assert_close(_np(st["exp_avg"]), z[f"state/{i}/exp_avg"], "float32", f"exp_avg {i} after resume")
        assert_close(_np(st["exp_avg_sq"]), z[f"state/{i}/exp_avg_sq"], "float32", f"exp_avg_sq {i} after resume")

    # ... and back: what this harness saves has the reference's files and keys, loadable by plain torch objects the
    # way the reference's _load does it (models_state_dict[i] -> load_state_dict, same for optimizers)
    ctl._save(tmp_path / "ours")
    ctx.close()
    saved = torch.load(tmp_path / "ours" / "state_dict.pth", map_location="cpu", weights_only=False)
    theirs = torch.load(CKPT / "state_dict.pth", map_location="cpu", weights_only=False)
    assert set(theirs) <= set(saved)
    assert list(saved["models_state_dict"][0]) == list(theirs["models_state_dict"][0])
    plain = GoldenNet()
    plain.load_state_dict(saved["models_state_dict"][0])
    plain_opt = make_optimizer(plain, HP)
    plain_opt.load_state_dict(saved["optimizers_state_dict"][0])
    for i, p in enumerate(plain.parameters()):
        np.testing.assert_array_equal(_np(p), rec["params"][-1][i])
        assert float(plain_opt.state[p]["step"]) == SAVE_AT + RESUME_STEPS
    with (tmp_path / "ours" / "trial_state.pkl").open("rb") as f:
        state = pickle.load(f)
    with (CKPT / "trial_state.pkl").open("rb") as f:
        assert set(state) == set(pickle.load(f))
    assert state["batches_trained"] == state["last_ckpt"] == (SAVE_AT + RESUME_STEPS) * AGG


def test_checkpoint_of_another_trial_only_warm_starts(lib):
    """trial_id differs -> weights and optimizer state load, counters start from zero (reference :1273-1279)."""
    rec = {"norm": [], "loss": [], "batch_idx": [], "params": []}
    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(hparams=HP, slots_per_trial=1, aggregation_frequency=AGG, trial_seed=17)
    trial = CkptTrial(ctx, rec)
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=AGG, reporting_period=AGG, trial_id=7)
    ctl._load(CKPT)
    assert ctl.batches_trained == 0 and ctl.state.batches_trained == 0 and ctl.state.trial_id == 7
    z = np.load(CKPT / "continuation.npz")
    np.testing.assert_array_equal(np.concatenate([_np(p).ravel() for p in trial.plist]), z["saved_params_check"])
    ctl.run()
    assert rec["batch_idx"] == [0, 1]
    assert float(trial.opt.state[trial.plist[0]]["step"]) == SAVE_AT + 1  # Adam's own counter came with the state
    ctx.close()
