"""Legacy entrypoint `model_def:TrialClass` (reference harness/determined/exec/harness.py:26-203 and
exec/launch.py:29-63): build the context from the experiment config, construct the trial, train for the
searcher's length.

    DET_EXPERIMENT_CONFIG=exp.yaml python -m determined_clone_b200.exec.harness model_def:MyTrial
"""
import importlib
import os
import sys
from typing import Any, List

from determined_clone_b200 import _info
from determined_clone_b200.pytorch import _trainer


def load_trial_class(entrypoint: str) -> Any:
    module, _, name = entrypoint.partition(":")
    if not name:
        raise ValueError(f"entrypoint must look like module:TrialClass, got {entrypoint!r}")
    sys.path.insert(0, os.getcwd())
    return getattr(importlib.import_module(module), name)


def main(argv: List[str]) -> int:
    if len(argv) != 1:
        print("usage: python -m determined_clone_b200.exec.harness module:TrialClass   (with DET_EXPERIMENT_CONFIG set)", file=sys.stderr)
        return 2
    if _info.get_cluster_info() is None:
        print("DET_EXPERIMENT_CONFIG must point at the experiment YAML", file=sys.stderr)
        return 2
    trial_class = load_trial_class(argv[0])
    with _trainer.init() as context:
        trial = trial_class(context)
        trainer = _trainer.Trainer(trial, context)
        conf = context.get_experiment_config()
        gbs = context._global_batch_size

        def period(key: str) -> Any:  # {"batches": n} | {"epochs": n} | {"records": n}; absent or 0 = never
            spec = {k: v for k, v in (conf.get(key) or {}).items() if v}
            return _trainer.TrainUnit._from_values(**spec, global_batch_size=gbs) if spec else None

        context._set_is_pre_trainer()
        reports = trainer.fit(
            checkpoint_period=period("min_checkpoint_period"),
            validation_period=period("min_validation_period"),
            reporting_period=_trainer.Batch(int(conf.get("scheduling_unit", 100))),
            checkpoint_policy=conf.get("checkpoint_policy", "best"),
            latest_checkpoint=_info.get_cluster_info().latest_checkpoint,
            step_zero_validation=bool(conf.get("perform_initial_validation", False)),
        )
        if context.distributed.rank == 0:
            for report in reports:
                print({k: report[k] for k in ("steps_completed", "avg_metrics", "samples_per_second") if k in report}, flush=True)
            out = os.environ.get("DET_REPORT_FILE")  # machine-readable copy for tests and tools
            if out:
                import json

                def plain(v: Any) -> Any:
                    try:
                        return None if v is None else float(v)
                    except (TypeError, ValueError):
                        return str(v)

                with open(out, "w") as f:
                    json.dump([{"steps_completed": r["steps_completed"],
                                "avg_metrics": {k: plain(v) for k, v in r["avg_metrics"].items()},
                                "batch_metrics": [{k: plain(v) for k, v in b.items()} for b in r["batch_metrics"]]}
                               for r in reports], f)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
