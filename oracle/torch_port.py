"""Port of the reference's per-step training path onto plain torch, as the reference executes it on CPU.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/gradsync_oracle.py for the rules): used by bench.py's
cpu_baseline and `--impl reference` legs and pinned by tests/test_oracle_golden.py.  Nothing in
determined_clone_b200/ imports it.

The reference's hot path is ~150 lines of Python that call torch (a third-party dependency, not under
/root/reference): this module restates those lines in the same order with the same torch calls, so
that timing it on the host cores is timing what `PyTorchTrial on CPU` does per batch, minus the
control plane.  Single process: with one slot the reference uses no DDP wrapper at all
(harness/determined/pytorch/_pytorch_context.py:295-298).

  RefContext.backward                _pytorch_context.py:718-816  (loss.backward(); grads accumulate in .grad)
  RefContext._average_gradients      _pytorch_context.py:818-825  (one div_ per parameter)
  RefContext.step_optimizer          _pytorch_context.py:827-924  (average -> clip -> step -> zero_grad)
  RefContext._should_communicate...  _pytorch_context.py:711-716
  train_batch_like_reference         _pytorch_trial.py:877-939    (to_device, train_batch, metric.cpu() per batch)
"""
from typing import Any, Callable, Dict, List, Optional

import numpy as np
import torch


class RefContext:
    def __init__(self, aggregation_frequency: int = 1, device: str = "cpu", hparams: Optional[Dict[str, Any]] = None,
                 slots_per_trial: int = 1) -> None:
        self._aggregation_frequency = aggregation_frequency
        self.device = torch.device(device)
        self._current_batch_idx: Optional[int] = None
        self._optimizer_configs: Dict[Any, Dict[str, Any]] = {}
        self.models: List[torch.nn.Module] = []
        self.optimizers: List[torch.optim.Optimizer] = []
        self._hparams = hparams or {}
        self._slots = slots_per_trial

    def get_per_slot_batch_size(self) -> int:
        return int(self._hparams["global_batch_size"]) // self._slots

    def wrap_model(self, model: torch.nn.Module) -> torch.nn.Module:
        model = model.to(self.device)
        self.models.append(model)
        return model

    def wrap_optimizer(self, optimizer: torch.optim.Optimizer, backward_passes_per_step: int = 1,
                       average_aggregated_gradients: bool = True) -> torch.optim.Optimizer:
        self._optimizer_configs[optimizer] = {"average_aggregated_gradients": average_aggregated_gradients}
        self.optimizers.append(optimizer)
        return optimizer

    def _should_communicate_and_update(self) -> bool:
        if self._current_batch_idx is None:
            raise RuntimeError("Training hasn't started.")
        return (self._current_batch_idx + 1) % self._aggregation_frequency == 0

    def backward(self, loss: torch.Tensor) -> None:
        ddp = [m for m in self.models if isinstance(m, torch.nn.parallel.DistributedDataParallel)]
        if ddp and not self._should_communicate_and_update():
            # DDP all-reduces on every backward unless told otherwise (_pytorch_context.py:802-812)
            import contextlib

            with contextlib.ExitStack() as stack:
                for m in ddp:
                    stack.enter_context(m.no_sync())
                loss.backward()
        else:
            loss.backward()

    @staticmethod
    def _average_gradients(parameters: Any, divisor: int) -> None:
        if divisor == 1:
            return
        divisor_value = float(divisor)
        for p in filter(lambda param: param.grad is not None, parameters):
            p.grad.data.div_(divisor_value)

    def step_optimizer(self, optimizer: torch.optim.Optimizer, clip_grads: Optional[Callable] = None,
                       auto_zero_grads: bool = True) -> None:
        if not self._should_communicate_and_update():
            return
        parameters = [p for group in optimizer.param_groups for p in group.get("params", [])]
        if self._optimizer_configs[optimizer]["average_aggregated_gradients"]:
            self._average_gradients(parameters=parameters, divisor=self._aggregation_frequency)
        if clip_grads is not None:
            clip_grads(parameters)
        optimizer.step()
        if auto_zero_grads:
            optimizer.zero_grad()


def train_batch_like_reference(ctx: RefContext, trial: Any, batch: Any, batch_idx: int) -> Dict[str, Any]:
    """What _PyTorchTrialController._train_batch does around the user's train_batch."""
    ctx._current_batch_idx = batch_idx
    batch = [b.to(ctx.device) for b in batch] if isinstance(batch, (list, tuple)) else batch.to(ctx.device)
    metrics = trial.train_batch(batch=batch, epoch_idx=0, batch_idx=batch_idx)
    if isinstance(metrics, torch.Tensor):
        metrics = {"loss": metrics}
    for name, metric in metrics.items():
        if isinstance(metric, torch.Tensor):
            metric = metric.cpu().detach().numpy()  # the per-batch device->host read of the reference
        metrics[name] = metric
    return metrics


def make_metrics(batch_metrics: List[Dict[str, Any]]) -> Dict[str, Any]:
    """avg over batches as util.make_metrics does (harness/determined/util.py:171-196)."""
    names = list(batch_metrics[0].keys()) if batch_metrics else []
    return {"avg_metrics": {k: float(np.mean([np.asarray(b[k]) for b in batch_metrics])) for k in names},
            "batch_metrics": batch_metrics}
