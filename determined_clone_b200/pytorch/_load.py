"""Rebuild a trained PyTorchTrial from a checkpoint directory, outside a training run (reference
harness/determined/pytorch/_load.py: CheckpointLoadContext :14-65, load_trial_from_checkpoint_path :68-237).

A checkpoint names its trial class three ways, newest first: load_data.json written by _save (`trial_cls_spec`,
plus experiment_config / hparams for trials that were launched through the legacy `entrypoint: module:Class`);
metadata.json of releases up to 0.17.7, where the class is the experiment config's entrypoint; or the caller passes
trial_class.  The class is imported from the checkpoint's code/ directory when it is not given."""
import importlib
import json
import logging
import pathlib
import warnings
from typing import Any, Dict, Optional, Type

import torch

from determined_clone_b200 import _import, errors
from determined_clone_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_clone_b200.pytorch._pytorch_trial import PyTorchTrial

logger = logging.getLogger("determined.pytorch")


class CheckpointLoadContext(PyTorchTrialContext):
    """The context a Trial class is constructed with when it is only being loaded: wrap_model / wrap_optimizer /
    wrap_lr_scheduler register what the trial builds so that the saved state can be put back, but no gradient
    engine is created -- it needs no GPU, and it cannot train (step_optimizer refuses)."""

    def __init__(self, hparams: Optional[Dict] = None, exp_conf: Optional[Dict[str, Any]] = None) -> None:
        super().__init__(
            core_context=None, trial_seed=0, hparams=hparams, slots_per_trial=1, num_gpus=0, exp_conf=exp_conf,
            aggregation_frequency=1, steps_completed=0, managed_training=False, debug_enabled=False,
        )
        self._load_only = True


def _class_from_spec(spec: str, code_dir: pathlib.Path) -> Type[PyTorchTrial]:
    module, _, qualname = spec.partition(":")
    with _import.import_from_path(code_dir):
        obj: Any = importlib.import_module(module)
        for part in qualname.split("."):
            obj = getattr(obj, part)
    return obj


def load_trial_from_checkpoint_path(
    path: str,
    trial_class: Optional[Type[PyTorchTrial]] = None,
    trial_kwargs: Optional[Dict[str, Any]] = None,
    torch_load_kwargs: Optional[Dict[str, Any]] = None,
    **kwargs: Any,
) -> Any:
    """Returns an instance of the trial class with the checkpoint's model weights loaded (for checkpoints of the
    pre-0.12.13 `build_model()` API: the bare model).  trial_kwargs are passed to the class after the context;
    torch_load_kwargs go to torch.load (e.g. {"map_location": "cpu"})."""
    trial_kwargs = trial_kwargs or {}
    torch_load_kwargs = torch_load_kwargs or {}
    if kwargs:
        if torch_load_kwargs:
            raise ValueError("kwargs may not be set if torch_load_kwargs is also set")
        warnings.warn(
            "Passing **kwargs to load_trial_from_checkpoint_path() is deprecated, and support for "
            "doing so will be removed in the future.  Please pass a dict of kwargs as "
            "`torch_load_kwargs` instead.",
            FutureWarning,
            stacklevel=2,
        )
        torch_load_kwargs = kwargs

    ckpt_dir = pathlib.Path(path)
    maybe_build_model = False  # only checkpoints old enough to lack load_data.json can predate wrap_model()
    if (ckpt_dir / "load_data.json").exists():
        info = json.loads((ckpt_dir / "load_data.json").read_text())
        if info["trial_type"] != "PyTorchTrial":
            logger.warning("Checkpoint does not appear to be a valid PyTorchTrial checkpoint, continuing anyway...")
        experiment_config, hparams, spec = info["experiment_config"], info["hparams"], info["trial_cls_spec"]
        is_trainer = info.get("is_trainer", False)
        if info.get("trial_in_main", False):
            module, _, qualname = spec.partition(":")
            if module == "":
                raise ValueError(
                    f"Unable to load trial class {qualname}, which was defined in the entrypoint "
                    "script, but the entrypoint script did not appear to be an importable module. "
                    "Define your Trial class in an importable module instead."
                )
            logger.warning(
                f"Importing trial class {qualname}, which was defined in the entrypoint script. "
                f"Assuming that entrypoint is importable via `import {module}`.  Otherwise, define "
                "your Trial class in an importable file."
            )
    elif (ckpt_dir / "metadata.json").exists():
        meta = json.loads((ckpt_dir / "metadata.json").read_text())
        if not meta.get("framework", "").startswith("torch") and "torch_version" not in meta:
            logger.warning("Checkpoint does not appear to be a valid PyTorchTrial checkpoint, continuing anyway...")
        experiment_config, hparams = meta["experiment_config"], meta["hparams"]
        spec, is_trainer, maybe_build_model = experiment_config["entrypoint"], False, True
    else:
        raise AssertionError(
            "Checkpoint does not have either load_data.json or metadata.json.  Checkpoints written "
            "by Determined 0.17.7 and earlier did not save enough information to be loaded "
            "directly from the files in the checkpoint.  Instead, a metadata.json was written "
            "during the call to Checkpoint.download().  If you are reading an old checkpoint "
            "directly from checkpoint storage, you can either use Checkpoint.download() instead or "
            "you can use Checkpoint.write_metadata_file('metadata.json') to create a suitable "
            "metadata file for loading a legacy checkpoint."
        )

    if not is_trainer:
        # launched as `entrypoint: module:Class`: such checkpoints always carry their config and hyperparameters
        assert experiment_config is not None and hparams is not None
    if trial_class is None:
        try:
            trial_class = _class_from_spec(spec, ckpt_dir / "code")
        except Exception as e:
            raise ValueError(
                f"Automatic import logic failed to import Trial class as {spec}. "
                "You will need to provide your Trial class via the trial_class argument intead."
            ) from e
    context = CheckpointLoadContext(hparams, experiment_config)
    if experiment_config:
        opts = experiment_config.get("optimizations", {}) or {}
        context._aggregation_frequency = int(opts.get("aggregation_frequency", 1))
        context._set_default_gradient_compression(bool(opts.get("gradient_compression", False)))
        context._set_default_average_aggregated_gradients(bool(opts.get("average_aggregated_gradients", True)))
    trial = trial_class(context, **trial_kwargs)

    torch_load_kwargs.setdefault("weights_only", False)  # the file also holds numpy / python RNG state
    checkpoint = torch.load(str(ckpt_dir / "state_dict.pth"), **torch_load_kwargs)
    build_model = getattr(trial, "build_model", None) if maybe_build_model else None
    if callable(build_model):
        # the pre-0.12.13 API: the trial never called wrap_model(); hand back the bare model with its weights
        model = build_model()
        model.load_state_dict(checkpoint["model_state_dict"] if "model_state_dict" in checkpoint else checkpoint["models_state_dict"][0])
        return model
    if "model_state_dict" in checkpoint:
        raise errors.InvalidCheckpointException()
    for model, saved in zip(context.models, checkpoint["models_state_dict"]):
        saved = dict(saved)
        torch.nn.modules.utils.consume_prefix_in_state_dict_if_present(saved, "module.")  # written by a DDP-wrapped model
        model.load_state_dict(saved)
    return trial
