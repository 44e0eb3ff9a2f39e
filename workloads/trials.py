"""PyTorchTrial-style workloads for BASELINE.json's configs, written against the trial-facing API only
(context.wrap_model / wrap_optimizer / backward / step_optimizer), so the same class runs under this
repo's PyTorchTrialContext on the GPU, under the torch port of the reference path on the CPU
(oracle/torch_port.RefContext) and under torch DDP on the GPU.  Synthetic data and random-initialised
weights: there is no network for datasets or checkpoints.
"""
from typing import Any, Dict, Tuple

import torch
from torch import nn


class _VisionTrial:
    NAME = ""
    INPUT_SHAPE: Tuple[int, ...] = ()
    NUM_CLASSES = 10
    PER_SLOT_BATCH = 128
    AGG = 1
    DTYPE = "f32"
    STEP_KERNEL = "dgs::sgd_kernel<float,float,true>"
    STEP_BYTES_PER_ELEM = 20  # read g,p,buf + write p,buf (fp32)
    OPTIMIZER = ""
    WEIGHT_DECAY = 5e-4

    def _net(self) -> nn.Module:
        raise NotImplementedError

    def __init__(self, context: Any) -> None:
        self.context = context
        self.model = context.wrap_model(self._net())
        self.optimizer = context.wrap_optimizer(
            torch.optim.SGD(self.model.parameters(), lr=0.1, momentum=0.9, weight_decay=self.WEIGHT_DECAY)
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


class MnistTrial(_VisionTrial):
    """configs[0]: the MNIST tutorial trial (reference examples/tutorials/mnist_pytorch/model.py:17-35 and
    train.py:28-72, hyper-parameters of const.yaml): two 3x3 convolutions, max-pool, two linear layers,
    NLL loss, Adadelta lr 1.0, batch 64.  Dropout is set to 0 so that runs are comparable across back ends.
    Adadelta is not one of the fused optimizers: its step() stays in torch on the engine's gradient views."""

    NAME = "mnist_pytorch_tutorial_trial"
    INPUT_SHAPE = (1, 28, 28)
    NUM_CLASSES = 10
    PER_SLOT_BATCH = 64
    STEP_KERNEL = "torch Adadelta (not fused)"
    OPTIMIZER = "adadelta lr 1.0"

    def __init__(self, context: Any) -> None:
        self.context = context
        n1, n2 = 32, 64
        net = nn.Sequential(
            nn.Conv2d(1, n1, 3, 1), nn.ReLU(), nn.Conv2d(n1, n2, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Dropout2d(0.0),
            nn.Flatten(), nn.Linear(144 * n2, 128), nn.ReLU(), nn.Dropout(0.0), nn.Linear(128, 10), nn.LogSoftmax(dim=1),
        )
        self.model = context.wrap_model(net)
        self.optimizer = context.wrap_optimizer(torch.optim.Adadelta(self.model.parameters(), lr=1.0))
        self.loss_fn = nn.NLLLoss()


class CifarResNetTrial(_VisionTrial):
    """configs[1]: CIFAR-10 ResNet (torchvision resnet18, 10 classes; the reference's
    examples/computer_vision/cifar10_pytorch directory is absent from the snapshot),
    SGD lr 0.1 momentum 0.9 weight decay 5e-4, fp32, aggregation_frequency 1."""

    NAME = "cifar10_resnet18_pytorchtrial"
    INPUT_SHAPE = (3, 32, 32)
    NUM_CLASSES = 10
    PER_SLOT_BATCH = 128
    OPTIMIZER = "sgd_momentum lr 0.1 m 0.9 wd 5e-4"

    def _net(self) -> nn.Module:
        import torchvision

        return torchvision.models.resnet18(num_classes=self.NUM_CLASSES)


class ResNet50Trial(_VisionTrial):
    """configs[2]: ResNet-50, synthetic 224x224 batches, fp32 SGD + momentum, per-slot batch 64."""

    NAME = "resnet50_pytorchtrial"
    INPUT_SHAPE = (3, 224, 224)
    NUM_CLASSES = 1000
    PER_SLOT_BATCH = 64
    WEIGHT_DECAY = 1e-4
    OPTIMIZER = "sgd_momentum lr 0.1 m 0.9 wd 1e-4"

    def _net(self) -> nn.Module:
        import torchvision

        return torchvision.models.resnet50(num_classes=self.NUM_CLASSES)


class GPT2MediumTrial:
    """configs[3]: GPT-2 medium causal LM (n_embd 1024, 24 layers, 16 heads, vocab 50257, tied head),
    parameters in bf16, AdamW lr 3e-4 betas (0.9, 0.95) eps 1e-8 weight decay 0.1 (none on 1-D
    parameters), aggregation_frequency 4, clip_grad_norm_ 1.0, per-slot batch 4 x 1024 tokens."""

    NAME = "gpt2_medium_causal_lm_pytorchtrial"
    PER_SLOT_BATCH = 4
    SEQ_LEN = 1024
    VOCAB = 50257
    AGG = 4
    DTYPE = "bf16"
    STEP_KERNEL = "dgs::adam_kernel<bf16,bf16>"
    STEP_BYTES_PER_ELEM = 14  # read g,p,m,v + write p,m,v (bf16)
    OPTIMIZER = "adamw lr 3e-4 betas (0.9,0.95) wd 0.1, clip_grad_norm_ 1.0"

    def __init__(self, context: Any) -> None:
        from transformers import GPT2Config, GPT2LMHeadModel

        self.context = context
        cfg = GPT2Config(vocab_size=self.VOCAB, n_positions=self.SEQ_LEN, n_embd=1024, n_layer=24, n_head=16,
                         resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, use_cache=False)
        model = GPT2LMHeadModel(cfg).to(torch.bfloat16)
        self.model = context.wrap_model(model)
        decay = [p for p in self.model.parameters() if p.ndim >= 2]
        no_decay = [p for p in self.model.parameters() if p.ndim < 2]
        self.optimizer = context.wrap_optimizer(torch.optim.AdamW(
            [{"params": decay, "weight_decay": 0.1}, {"params": no_decay, "weight_decay": 0.0}],
            lr=3e-4, betas=(0.9, 0.95), eps=1e-8,
        ))

    @staticmethod
    def _clip(params: Any) -> None:
        torch.nn.utils.clip_grad_norm_(params, 1.0)

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        tokens = batch[0] if isinstance(batch, (list, tuple)) else batch
        loss = self.model(input_ids=tokens, labels=tokens).loss
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self._clip)
        return {"loss": loss}

    def get_batch_length(self, batch: Any) -> int:
        tokens = batch[0] if isinstance(batch, (list, tuple)) else batch
        return int(tokens.shape[0])

    @classmethod
    def synthetic_batch(cls, batch_size: int, generator: torch.Generator) -> Tuple[torch.Tensor]:
        return (torch.randint(0, cls.VOCAB, (batch_size, cls.SEQ_LEN), generator=generator),)


WORKLOADS = {"mnist": MnistTrial, "cifar_resnet18": CifarResNetTrial, "resnet50": ResNet50Trial, "gpt2_medium": GPT2MediumTrial}
