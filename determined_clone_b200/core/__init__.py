from determined_clone_b200.core._distributed import DistributedContext, DummyDistributedContext

__all__ = ["DistributedContext", "DummyDistributedContext"]
