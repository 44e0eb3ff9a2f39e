"""Which data-parallel backend the launcher selected (reference harness/determined/_trial_controller.py:10-27).
Both switches lead to the same engine here: Horovod and torch DDP are replaced by libdgs."""
import os


class _DistributedBackend:
    HOROVOD = "USE_HOROVOD"
    DEEPSPEED = "USE_DEEPSPEED"
    TORCH = "USE_TORCH_DISTRIBUTED"

    def use_horovod(self) -> bool:
        return bool(os.environ.get(self.HOROVOD, None))

    def use_deepspeed(self) -> bool:
        return bool(os.environ.get(self.DEEPSPEED, None))

    def use_torch(self) -> bool:
        return bool(os.environ.get(self.TORCH, None))
