"""_make_local_execution_env(): what the reference's tests (and its checkpoint-export path) use to get a
(core.Context, environment) pair without a cluster (reference harness/determined/_execution.py:163-206).  The
environment here carries the few fields a PyTorchTrialContext is built from."""
import time
import types
from typing import Any, Dict, Optional, Tuple


def _make_local_execution_env(managed_training: bool, test_mode: bool, config: Optional[Dict[str, Any]], checkpoint_dir: str,
                              tensorboard_path: Any = None, hparams: Optional[Dict[str, Any]] = None,
                              limit_gpus: Optional[int] = None) -> Tuple[Any, Any]:
    from determined_clone_b200 import core

    config = dict(config or {})
    seed = (config.get("reproducibility") or {}).get("experiment_seed")
    env = types.SimpleNamespace(
        experiment_config=config,
        hparams=hparams if hparams is not None else dict(config.get("hyperparameters", {})),
        trial_seed=int(seed) if seed is not None else int(time.time()),
        latest_checkpoint=None, steps_completed=0, managed_training=managed_training, test_mode=test_mode, on_cluster=False,
    )
    return core._dummy_init(checkpoint_storage=checkpoint_dir, tensorboard_path=tensorboard_path), env
