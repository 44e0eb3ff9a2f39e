#!/usr/bin/env python
"""Generate tests/golden/lr_schedule.json: learning rates the REFERENCE controller produces when it steps a wrapped
LR scheduler automatically (reference _pytorch_trial.py:845-870), for every step mode x aggregation_frequency x
scheduler frequency below.  Run in the build container only (needs /root/reference):

    python tests/golden/make_golden_lr.py

The trial is a one-weight model trained by the unmodified reference harness on CPU (driven as in make_golden.py);
its scheduler is StepLR(step_size=1, gamma=0.5) on lr 1.0, so every scheduler step halves the rate and the number of
steps taken so far can be read off the recorded value exactly.  Recorded: the rate in force when train_batch(i)
starts, for i = 0..BATCHES-1, then the rate after the last batch."""
import json
import pathlib
import sys
import tempfile
from typing import Any, Dict, List

HERE = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
import make_golden  # noqa: E402

BATCHES, EPOCH_LEN = 24, 6
CASES = [(mode, agg, freq) for mode in ("STEP_EVERY_BATCH", "STEP_EVERY_OPTIMIZER_STEP", "STEP_EVERY_EPOCH", "MANUAL_STEP")
         for agg in (1, 2, 4) for freq in (1, 2, 3)]


def run_case(mode: str, agg: int, freq: int) -> List[float]:
    import torch

    import determined as det
    from determined import core, pytorch

    seen: List[float] = []

    class LrTrial(pytorch.PyTorchTrial):
        def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
            self.context = context
            self.model = context.wrap_model(torch.nn.Linear(1, 1, bias=False))
            self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=1.0))
            self.sched = context.wrap_lr_scheduler(
                torch.optim.lr_scheduler.StepLR(self.opt, step_size=1, gamma=0.5),
                step_mode=getattr(pytorch.LRScheduler.StepMode, mode), frequency=freq,
            )

        def build_training_data_loader(self) -> pytorch.DataLoader:
            ds = torch.utils.data.TensorDataset(torch.zeros(EPOCH_LEN * 2, 1))
            return pytorch.DataLoader(ds, batch_size=2)

        def build_validation_data_loader(self) -> pytorch.DataLoader:
            return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.zeros(2, 1)), batch_size=2)

        def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
            seen.append(self.opt.param_groups[0]["lr"])
            loss = self.model(batch[0]).sum()
            self.context.backward(loss)
            self.context.step_optimizer(self.opt)
            return {"loss": loss}

        def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
            return {"val_loss": 0.0}

    with tempfile.TemporaryDirectory() as storage:
        hp = {"global_batch_size": 2}
        core_context = core._dummy_init(checkpoint_storage=storage)
        core_context.preempt.start()
        core_context.train._trial_id = "1"
        exp_config = {
            "scheduling_unit": BATCHES, "resources": {"native_parallel": False, "slots_per_trial": 1}, "hyperparameters": hp,
            "optimizations": {"aggregation_frequency": agg, "gradient_compression": False, "average_training_metrics": True,
                              "average_aggregated_gradients": True},
            "checkpoint_policy": "none", "perform_initial_validation": False,
            "checkpoint_storage": {"type": "shared_fs", "host_path": storage},
            "searcher": {"metric": "val_loss", "smaller_is_better": True},
            "min_checkpoint_period": {"batches": 0}, "min_validation_period": {"batches": 0}, "data": {},
        }
        pytorch._PyTorchTrialController.pre_execute_hook(17, det._DistributedBackend())
        context = pytorch.PyTorchTrialContext(
            core_context=core_context, trial_seed=17, hparams=hp, slots_per_trial=1, num_gpus=0, exp_conf=exp_config,
            aggregation_frequency=agg, steps_completed=0, managed_training=True, debug_enabled=False, enable_tensorboard_logging=False,
        )
        trial = LrTrial(context)
        controller = pytorch._PyTorchTrialController(
            trial_inst=trial, context=context, max_length=pytorch.Batch(BATCHES), checkpoint_period=pytorch.Batch(sys.maxsize),
            validation_period=pytorch.Batch(sys.maxsize), searcher_metric_name="val_loss", reporting_period=pytorch.Batch(BATCHES),
            local_training=True, latest_checkpoint=None, steps_completed=0, smaller_is_better=True, test_mode=False,
            checkpoint_policy="none", step_zero_validation=False, det_profiler=None, global_batch_size=None,
        )
        controller.run()
        seen.append(trial.opt.param_groups[0]["lr"])
    return seen


def main() -> None:
    make_golden._install_shims()
    out = {"batches": BATCHES, "epoch_len": EPOCH_LEN, "cases": []}
    for mode, agg, freq in CASES:
        out["cases"].append({"mode": mode, "aggregation_frequency": agg, "frequency": freq, "lr": run_case(mode, agg, freq)})
    (HERE / "lr_schedule.json").write_text(json.dumps(out, indent=1))
    print("wrote", HERE / "lr_schedule.json", len(out["cases"]), "cases")


if __name__ == "__main__":
    main()
