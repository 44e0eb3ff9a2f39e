"""Minimal stand-in for det.get_cluster_info() on a single box without a master
(reference harness/determined/_info.py, ClusterInfo/TrialInfo): the single-box launcher passes the
experiment config through DET_EXPERIMENT_CONFIG (a YAML path), and the trial reads hyper-parameters,
slots, the searcher length and the optimizations block from it exactly as it would on a cluster
(reference harness/determined/exec/harness.py:147-201)."""
import os
from typing import Any, Dict, List, Optional


class TrialInfo:
    def __init__(self, config: Dict[str, Any]) -> None:
        self._config = config
        self.hparams: Dict[str, Any] = dict(config.get("hyperparameters", {}))
        self.trial_seed = int(config.get("reproducibility", {}).get("experiment_seed", 17))
        self._steps_completed = 0
        self._debug = bool(config.get("debug", False))
        self.trial_id, self.experiment_id = 1, 1


class ClusterInfo:
    task_type = "TRIAL"

    def __init__(self, config: Dict[str, Any]) -> None:
        self.trial = TrialInfo(config)
        self.container_addrs: List[str] = ["127.0.0.1"]
        self.container_rank = 0
        slots = int(config.get("resources", {}).get("slots_per_trial", 1))
        self.slot_ids = list(range(slots))
        self.gpu_uuids = [str(i) for i in self.slot_ids]
        self.allocation_id = "local"
        self.latest_checkpoint: Optional[str] = None


def load_experiment_config(path: str) -> Dict[str, Any]:
    import yaml

    with open(path) as f:
        return yaml.safe_load(f) or {}


def get_cluster_info() -> Optional[ClusterInfo]:
    """ClusterInfo of the local 'cluster' when launched by determined_clone_b200.launch, else None."""
    path = os.environ.get("DET_EXPERIMENT_CONFIG")
    if not path:
        return None
    return ClusterInfo(load_experiment_config(path))
