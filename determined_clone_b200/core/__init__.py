from determined_clone_b200.core._distributed import DistributedContext, DummyDistributedContext
from determined_clone_b200.core._context import (
    Context,
    DummyCheckpointContext,
    DummyPreemptContext,
    DummyTrainContext,
    _dummy_init,
)

__all__ = ["Context", "DistributedContext", "DummyCheckpointContext", "DummyDistributedContext", "DummyPreemptContext",
           "DummyTrainContext", "_dummy_init"]
