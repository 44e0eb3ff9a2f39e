"""Model, data and optimizer definitions shared by the golden-vector generator (which runs them
through the REFERENCE harness) and by the parity tests (which run them through this repo's harness).

Pure torch: nothing here imports either harness.
"""
from typing import Any, Dict, List, Tuple

import torch
from torch import nn

IN_DIM, HIDDEN, OUT_DIM = 24, 40, 8
DATA_SEED = 1234
TRIAL_SEED = 17


class GoldenNet(nn.Module):
    """Small MLP with a LayerNorm so that there are 1-D parameters for a no-decay group."""

    def __init__(self) -> None:
        super().__init__()
        self.fc1 = nn.Linear(IN_DIM, HIDDEN)
        self.norm = nn.LayerNorm(HIDDEN)
        self.fc2 = nn.Linear(HIDDEN, OUT_DIM)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.fc2(torch.tanh(self.norm(self.fc1(x))))


def make_data(num_batches: int, global_batch: int) -> Tuple[torch.Tensor, torch.Tensor]:
    g = torch.Generator().manual_seed(DATA_SEED)
    n = num_batches * global_batch
    x = torch.randn(n, IN_DIM, generator=g)
    y = torch.randn(n, OUT_DIM, generator=g)
    return x, y


def param_groups(model: nn.Module, weight_decay: float) -> List[Dict[str, Any]]:
    decay = [p for p in model.parameters() if p.ndim >= 2]
    no_decay = [p for p in model.parameters() if p.ndim < 2]
    return [
        {"params": decay, "weight_decay": weight_decay},
        {"params": no_decay, "weight_decay": 0.0},
    ]


def make_optimizer(model: nn.Module, hp: Dict[str, Any]) -> torch.optim.Optimizer:
    kind = hp["optimizer"]
    groups = param_groups(model, hp.get("weight_decay", 0.0))
    if kind == "sgd":
        return torch.optim.SGD(
            groups,
            lr=hp["lr"],
            momentum=hp.get("momentum", 0.0),
            dampening=hp.get("dampening", 0.0),
            nesterov=hp.get("nesterov", False),
        )
    if kind == "adamw":
        return torch.optim.AdamW(groups, lr=hp["lr"], betas=tuple(hp["betas"]), eps=hp["eps"])
    if kind == "adam":
        return torch.optim.Adam(groups, lr=hp["lr"], betas=tuple(hp["betas"]), eps=hp["eps"])
    if kind == "adadelta":
        return torch.optim.Adadelta(groups, lr=hp["lr"])
    raise ValueError(kind)


# name -> (hparams, slots, aggregation_frequency, optimizer steps)
SCENARIOS: Dict[str, Dict[str, Any]] = {
    "sgd_n1_a1": dict(
        hp=dict(optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=5e-4, global_batch_size=8),
        slots=1, agg=1, steps=10, clip=None, dtype="float32",
    ),
    "sgd_nesterov_n1_a2_clipval": dict(
        hp=dict(optimizer="sgd", lr=0.05, momentum=0.8, dampening=0.0, nesterov=True,
                weight_decay=1e-3, global_batch_size=8),
        slots=1, agg=2, steps=10, clip=("value", 0.05), dtype="float32",
    ),
    "adamw_n1_a4_clip": dict(
        hp=dict(optimizer="adamw", lr=3e-4, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1,
                global_batch_size=8),
        slots=1, agg=4, steps=10, clip=("norm", 1.0), dtype="float32",
    ),
    "adam_n1_a1_clip": dict(
        hp=dict(optimizer="adam", lr=1e-3, betas=[0.9, 0.999], eps=1e-8, weight_decay=0.01,
                global_batch_size=8),
        slots=1, agg=1, steps=10, clip=("norm", 0.5), dtype="float32",
    ),
    "adadelta_n1_a1": dict(
        hp=dict(optimizer="adadelta", lr=1.0, global_batch_size=8),
        slots=1, agg=1, steps=10, clip=("norm", 1.0), dtype="float32",
    ),
    "sgd_n2_a2": dict(
        hp=dict(optimizer="sgd", lr=0.1, momentum=0.9, weight_decay=5e-4, global_batch_size=8),
        slots=2, agg=2, steps=10, clip=None, dtype="float32",
    ),
    "adamw_n2_a4_clip": dict(
        hp=dict(optimizer="adamw", lr=3e-4, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1,
                global_batch_size=8),
        slots=2, agg=4, steps=10, clip=("norm", 1.0), dtype="float32",
    ),
    "adamw_bf16_n1_a4_clip": dict(
        hp=dict(optimizer="adamw", lr=3e-4, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1,
                global_batch_size=8),
        slots=1, agg=4, steps=10, clip=("norm", 1.0), dtype="bfloat16",
    ),
    # optimizations.average_aggregated_gradients: false -> gradients of the 3 micro-batches are summed, not averaged
    "sgd_n1_a3_noavg": dict(
        hp=dict(optimizer="sgd", lr=0.02, momentum=0.9, weight_decay=1e-3, global_batch_size=8),
        slots=1, agg=3, steps=10, clip=("norm", 2.0), dtype="float32", average=False,
    ),
    "adamw_bf16_n2_a2_clip": dict(
        hp=dict(optimizer="adamw", lr=3e-4, betas=[0.9, 0.95], eps=1e-8, weight_decay=0.1,
                global_batch_size=8),
        slots=2, agg=2, steps=10, clip=("norm", 1.0), dtype="bfloat16",
    ),
}
