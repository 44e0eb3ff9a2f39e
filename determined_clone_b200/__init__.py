"""B200-native replacement for the per-step data-parallel path of Determined's PyTorch harness.

    import determined_clone_b200 as det
    from determined_clone_b200 import pytorch, core

`pytorch` and `core` expose the same names the reference's determined.pytorch / determined.core expose
on this path (PyTorchTrial, PyTorchTrialContext, MetricReducer, DistributedContext, ...).  The compute
lives in csrc/libdgs.so (hand-written sm_100a kernels + NCCL); see include/dgs.h and DESIGN.md.
"""
from determined_clone_b200 import errors, util  # noqa: F401
from determined_clone_b200._execution import _make_local_execution_env  # noqa: F401
from determined_clone_b200._import import import_from_path  # noqa: F401
from determined_clone_b200._info import ClusterInfo, get_cluster_info  # noqa: F401
from determined_clone_b200._trial_controller import _DistributedBackend  # noqa: F401

__version__ = "0.1.0"

# what train.py scripts pass to logging.basicConfig(format=...) (reference determined/__init__.py:26-30)
LOG_FORMAT = "%(levelname)s: [%(process)s] %(name)s: %(message)s"


class InvalidHP(Exception):
    """Raised by trial code to say that this hyper-parameter combination cannot train (reference
    determined/_execution.py).  Without a searcher to report to, it simply ends the run."""


def __getattr__(name: str):  # lazy: importing the package must not import torch
    if name in ("pytorch", "core"):
        import importlib

        return importlib.import_module(f"determined_clone_b200.{name}")
    raise AttributeError(name)
