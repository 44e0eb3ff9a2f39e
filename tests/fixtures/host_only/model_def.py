"""A trial without an optimizer, so the legacy entrypoint, the experiment-config plumbing and the training loop
(periods, validation, checkpoint policy) can run end to end on a box without a GPU.  Written against `determined`
names only, as user code is."""
from typing import Any, Dict

import torch

from determined import pytorch


class HostOnlyTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
        self.context = context
        self.model = context.wrap_model(torch.nn.Linear(1, 1, bias=False))
        with torch.no_grad():
            self.model.weight.fill_(float(context.get_hparam("start")))

    def build_training_data_loader(self) -> pytorch.DataLoader:
        data = torch.utils.data.TensorDataset(torch.arange(24).float())
        return pytorch.DataLoader(data, batch_size=self.context.get_per_slot_batch_size())

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        data = torch.utils.data.TensorDataset(torch.arange(8).float())
        return pytorch.DataLoader(data, batch_size=self.context.get_per_slot_batch_size())

    def train_batch(self, batch: pytorch.TorchData, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        with torch.no_grad():
            self.model.weight.sub_(0.5)  # "training": the validation metric below falls by 0.5 per batch
        return {"loss": float(self.model.weight)}

    def evaluate_batch(self, batch: pytorch.TorchData) -> Dict[str, Any]:
        return {"distance": abs(float(self.model.weight))}
