#!/usr/bin/env python
"""Generate tests/golden/ckpt_adamw_a2/: a checkpoint WRITTEN BY THE REFERENCE harness, plus what the
reference does after restoring it.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_ckpt.py

Stage 1: the unmodified reference controller (see make_golden.py for how it is driven) trains
golden_model.GoldenNet for SAVE_AT optimizer steps (AdamW, aggregation_frequency 2, clip_grad_norm_ 1.0) and
checkpoints through its own _save (reference _pytorch_trial.py:1312-1385).  state_dict.pth, trial_state.pkl
and load_data.json of that checkpoint are committed as they were written; the `code/` copy is not.

Stage 2: a fresh reference trial restores that checkpoint through the reference's run()/_load (:606-612,
:1117-1271) and trains RESUME_STEPS more optimizer steps.  continuation.npz records the per-batch losses,
clip norms, parameters after every step and the final optimizer state of that second stage: the values a
drop-in harness must reproduce after loading the same files.
"""
import os
import pathlib
import shutil
import sys
import tempfile
from typing import Any, Dict, List

import numpy as np

HERE = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
import make_golden  # noqa: E402

NAME = "ckpt_adamw_a2"
HP = dict(optimizer="adamw", lr=3e-3, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1, global_batch_size=8)
AGG, SAVE_AT, RESUME_STEPS, CLIP = 2, 3, 3, 1.0


def stage(storage: str, resume_from: Any, rec: Dict[str, List[Any]]) -> str:
    import torch

    import determined as det
    from determined import core, pytorch
    from golden_model import TRIAL_SEED, GoldenNet, make_data, make_optimizer

    torch.set_num_threads(1)
    start = 0 if resume_from is None else SAVE_AT
    stop = SAVE_AT if resume_from is None else SAVE_AT + RESUME_STEPS

    class RefCkptTrial(pytorch.PyTorchTrial):
        def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
            self.context = context
            self.model = context.wrap_model(GoldenNet())
            self.plist = list(self.model.parameters())
            self.opt = context.wrap_optimizer(make_optimizer(self.model, HP))
            x, y = make_data((SAVE_AT + RESUME_STEPS) * AGG, HP["global_batch_size"])
            self.ds = torch.utils.data.TensorDataset(x, y)

        def build_training_data_loader(self) -> pytorch.DataLoader:
            return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

        def build_validation_data_loader(self) -> pytorch.DataLoader:
            return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

        def _clip(self, params: Any) -> None:
            rec["norm"].append(make_golden._to_np(torch.nn.utils.clip_grad_norm_(list(params), CLIP)))

        def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
            x, y = batch
            loss = torch.nn.functional.mse_loss(self.model(x), y)
            self.context.backward(loss)
            rec["loss"].append(float(loss))
            rec["batch_idx"].append(batch_idx)
            self.context.step_optimizer(self.opt, clip_grads=self._clip)
            if (batch_idx + 1) % AGG == 0:
                rec["params"].append([make_golden._to_np(p) for p in self.plist])
            return {"loss": loss}

        def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
            x, y = batch
            return {"val_loss": torch.nn.functional.mse_loss(self.model(x), y)}

    core_context = core._dummy_init(checkpoint_storage=storage)
    core_context.preempt.start()
    core_context.train._trial_id = "1"
    exp_config = {
        "scheduling_unit": AGG,
        "resources": {"native_parallel": False, "slots_per_trial": 1},
        "hyperparameters": HP,
        "optimizations": {"aggregation_frequency": AGG, "gradient_compression": False,
                          "average_training_metrics": True, "average_aggregated_gradients": True},
        "checkpoint_policy": "all",
        "perform_initial_validation": False,
        "checkpoint_storage": {"type": "shared_fs", "host_path": storage},
        "searcher": {"metric": "val_loss", "smaller_is_better": True},
        "min_checkpoint_period": {"batches": 0},
        "min_validation_period": {"batches": 0},
        "data": {},
    }
    pytorch._PyTorchTrialController.pre_execute_hook(TRIAL_SEED, det._DistributedBackend())
    context = pytorch.PyTorchTrialContext(
        core_context=core_context, trial_seed=TRIAL_SEED, hparams=HP, slots_per_trial=1, num_gpus=0,
        exp_conf=exp_config, aggregation_frequency=AGG, steps_completed=start * AGG, managed_training=True,
        debug_enabled=False, enable_tensorboard_logging=False,
    )
    context._set_default_gradient_compression(False)
    context._set_default_average_aggregated_gradients(True)
    trial = RefCkptTrial(context)
    controller = pytorch._PyTorchTrialController(
        trial_inst=trial, context=context, max_length=pytorch.Batch(stop * AGG),
        checkpoint_period=pytorch.Batch(sys.maxsize), validation_period=pytorch.Batch(sys.maxsize),
        searcher_metric_name="val_loss", reporting_period=pytorch.Batch(AGG), local_training=True,
        latest_checkpoint=resume_from, steps_completed=start * AGG, smaller_is_better=True, test_mode=False,
        checkpoint_policy="all", step_zero_validation=False, det_profiler=None, global_batch_size=None,
    )
    controller.run()
    rec["state"] = [
        {k: (make_golden._to_np(v) if torch.is_tensor(v) else np.asarray(v)) for k, v in trial.opt.state[p].items()}
        for p in trial.plist
    ]
    rec["batches_trained"] = controller.state.batches_trained
    ckpts = sorted((d for d in pathlib.Path(storage).iterdir() if (d / "state_dict.pth").exists()), key=lambda d: d.stat().st_mtime)
    return ckpts[-1].name


def main() -> None:
    # the reference calls torch.load(path, map_location="cpu") (:1131) and was written for torch 1.x, where that
    # unpickles everything; torch >= 2.6 defaults to weights_only=True and rejects the numpy RNG state the
    # reference itself stored.  This restores the torch 1.x default for this process; the reference is untouched.
    os.environ["TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD"] = "1"
    make_golden._install_shims()
    out = HERE / NAME
    with tempfile.TemporaryDirectory() as storage:
        first: Dict[str, List[Any]] = {"norm": [], "loss": [], "batch_idx": [], "params": []}
        uuid = stage(storage, None, first)
        assert first["batches_trained"] == SAVE_AT * AGG
        out.mkdir(exist_ok=True)
        for f in ("state_dict.pth", "trial_state.pkl", "load_data.json"):
            shutil.copyfile(os.path.join(storage, uuid, f), out / f)
        second: Dict[str, List[Any]] = {"norm": [], "loss": [], "batch_idx": [], "params": []}
        stage(storage, uuid, second)
        assert second["batch_idx"][0] == SAVE_AT * AGG, second["batch_idx"]
        z: Dict[str, np.ndarray] = {
            "loss": np.asarray(second["loss"], dtype=np.float64),
            "batch_idx": np.asarray(second["batch_idx"], dtype=np.int64),
            "total_norm": np.asarray(second["norm"], dtype=np.float32),
            "batches_trained": np.asarray(second["batches_trained"], dtype=np.int64),
            "saved_params_check": np.concatenate([p.ravel() for p in first["params"][-1]]),
        }
        for s, params in enumerate(second["params"]):
            for i, p in enumerate(params):
                z[f"params/{s}/{i}"] = p
        for i, st in enumerate(second["state"]):
            for k, v in st.items():
                z[f"state/{i}/{k}"] = v
        np.savez_compressed(out / "continuation.npz", **z)
    print("wrote", out, sorted(f.name for f in out.iterdir()))


if __name__ == "__main__":
    main()
