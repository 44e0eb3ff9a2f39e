"""y = w * x trained with MSE and SGD: every weight has a closed form, which is how the reference checks
2-slot x aggregation_frequency=2 training end to end (its fixture pytorch_identity/model_def.py and
test_pytorch_trial.py:980-1001,1604-1642).  Written against `determined` names only, as user code is."""
from typing import Any, Dict

import torch

from determined import pytorch


class IdentityDataset(torch.utils.data.Dataset):
    def __init__(self, n: int = 64) -> None:
        self.x = [float(i) * 0.1 + 1.0 for i in range(n)]

    def __len__(self) -> int:
        return len(self.x)

    def __getitem__(self, i: int) -> Any:
        v = torch.tensor([self.x[i]], dtype=torch.float64)
        return v, v


class IdentityTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
        self.context = context
        model = torch.nn.Linear(1, 1, bias=False).to(torch.float32)
        model.weight.data.fill_(0.0)
        self.model = context.wrap_model(model)
        self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=context.get_hparam("learning_rate")))

    def build_training_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(IdentityDataset(), batch_size=self.context.get_per_slot_batch_size())

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(IdentityDataset(20), batch_size=self.context.get_per_slot_batch_size())

    def train_batch(self, batch: pytorch.TorchData, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x.float()), y.float())
        self.context.backward(loss)
        self.context.step_optimizer(self.opt)
        return {"loss": loss, "weight": self.model.weight.detach().reshape(())}

    def evaluate_batch(self, batch: pytorch.TorchData, batch_idx: int = 0) -> Dict[str, Any]:
        x, y = batch
        return {"validation_loss": torch.nn.functional.mse_loss(self.model(x.float()), y.float())}
