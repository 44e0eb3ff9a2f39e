"""`import determined as det` for unmodified trial code: put determined_clone_b200/compat on PYTHONPATH and
`determined`, `determined.pytorch`, `determined.core` resolve to this repo's implementation of the per-step
path.  Only names on that path exist; the rest of the Determined SDK (CLI, client, searchers) is out of scope."""
import sys

import determined_clone_b200 as _impl
from determined_clone_b200 import constants, core, errors, ipc, pytorch, util  # noqa: F401
from determined_clone_b200._execution import _make_local_execution_env  # noqa: F401
from determined_clone_b200._import import import_from_path  # noqa: F401
from determined_clone_b200._info import ClusterInfo, get_cluster_info  # noqa: F401
from determined_clone_b200._trial_controller import _DistributedBackend  # noqa: F401

__version__ = _impl.__version__
LOG_FORMAT, InvalidHP = _impl.LOG_FORMAT, _impl.InvalidHP
InvalidExperimentException = errors.InvalidExperimentException
for _name in ("constants", "core", "errors", "ipc", "pytorch", "util"):
    sys.modules[f"determined.{_name}"] = getattr(_impl, _name) if _name in ("core", "pytorch") else globals()[_name]
