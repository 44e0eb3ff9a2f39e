"""PyTorchTrial-style workloads for BASELINE.json's configs, written against the trial-facing API only
(context.wrap_model / wrap_optimizer / backward / step_optimizer), so the same class runs under this
repo's PyTorchTrialContext on the GPU and under the torch port of the reference path on the CPU
(oracle/torch_port.RefContext).  Synthetic data and random-initialised weights: there is no network for
datasets or checkpoints.
"""
from typing import Any, Dict, Tuple

import torch
from torch import nn


class CifarResNetTrial:
    """configs[1]: CIFAR-10 ResNet (torchvision resnet18, 10 classes; the reference's
    examples/computer_vision/cifar10_pytorch directory is absent from the snapshot),
    SGD lr 0.1 momentum 0.9 weight decay 5e-4, fp32, aggregation_frequency 1."""

    NAME = "cifar10_resnet18_pytorchtrial"
    INPUT_SHAPE = (3, 32, 32)
    NUM_CLASSES = 10

    def __init__(self, context: Any) -> None:
        import torchvision

        self.context = context
        self.model = context.wrap_model(torchvision.models.resnet18(num_classes=self.NUM_CLASSES))
        self.optimizer = context.wrap_optimizer(
            torch.optim.SGD(self.model.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
        )
        self.loss_fn = nn.CrossEntropyLoss()

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        x, y = batch
        loss = self.loss_fn(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss}

    def get_batch_length(self, batch: Any) -> int:
        return int(batch[0].shape[0])

    @classmethod
    def synthetic_batch(cls, batch_size: int, generator: torch.Generator) -> Tuple[torch.Tensor, torch.Tensor]:
        x = torch.randn(batch_size, *cls.INPUT_SHAPE, generator=generator)
        y = torch.randint(0, cls.NUM_CLASSES, (batch_size,), generator=generator)
        return x, y
