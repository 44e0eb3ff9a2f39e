"""LR scheduler wrapper with Determined's step modes (reference
harness/determined/pytorch/_lr_scheduler.py:7-85).  Schedulers mutate optimizer.param_groups[*]["lr"];
the fused step kernel re-reads the groups on every optimizer step, so nothing else is needed."""
import enum
from typing import Any, Dict, List


class LRScheduler:
    class StepMode(enum.Enum):
        STEP_EVERY_EPOCH = 1
        STEP_EVERY_BATCH = 2
        MANUAL_STEP = 3
        STEP_EVERY_OPTIMIZER_STEP = 4

    def __init__(self, scheduler: Any, step_mode: "LRScheduler.StepMode", frequency: int = 1) -> None:
        if not isinstance(step_mode, LRScheduler.StepMode):
            raise TypeError(f"step_mode must be an LRScheduler.StepMode. Got {type(step_mode)}.")
        self._scheduler, self._step_mode, self._frequency = scheduler, step_mode, frequency

    def step(self, *args: Any, **kwargs: Any) -> None:
        self._scheduler.step(*args, **kwargs)

    def get_last_lr(self) -> List:
        return self._scheduler.get_last_lr()

    def load_state_dict(self, state_dict: Dict[Any, Any]) -> None:
        self._scheduler.load_state_dict(state_dict)

    def state_dict(self) -> Dict[Any, Any]:
        return self._scheduler.state_dict()
