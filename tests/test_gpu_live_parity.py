"""Both arms on the SAME GPU and the same inputs: a trial under this repo's context (libdgs kernels) beside the torch
port of the reference's context lines (torch.optim + clip_grad_norm_), compared after every optimizer step by
workloads/parity.py -- the checker bench.py runs on the BASELINE.json workloads at every N.  No CPU-recorded
fixture is involved, so the comparison is element-wise: in per-step synced mode an fp32 element may differ from
the reference's by 1e-5 of its own value plus 1e-5 of the tensor's rms, nothing more.  16-bit: the reference
accumulates micro-batches in bf16, and Adam's first steps on zero-initialised biases follow the SIGN of gradients
that are rounding noise; the bound there is the reference's distance from its own twin (micro-batches of a window in
reverse order), see workloads/parity.py."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from determined_clone_b200 import pytorch  # noqa: E402
from workloads import parity as P  # noqa: E402


class _Base:
    NAME, PER_SLOT_BATCH, AGG, DTYPE = "", 32, 1, "f32"

    def get_batch_length(self, batch):
        return int(batch[0].shape[0])


class ConvBnSgd(_Base):
    """fp32, SGD momentum + weight decay, batch-norm buffers, no clipping, A = 1 (the single-launch step on one GPU)."""
    NAME = "conv_bn_sgd"

    def __init__(self, context):
        self.context = context
        net = torch.nn.Sequential(torch.nn.Conv2d(3, 16, 3, padding=1), torch.nn.BatchNorm2d(16), torch.nn.ReLU(),
                                  torch.nn.Conv2d(16, 16, 3, padding=1), torch.nn.BatchNorm2d(16), torch.nn.ReLU(),
                                  torch.nn.AdaptiveAvgPool2d(1), torch.nn.Flatten(), torch.nn.Linear(16, 10))
        self.model = context.wrap_model(net)
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4))

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.cross_entropy(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss}

    @classmethod
    def synthetic_batch(cls, n, gen):
        return torch.randn(n, 3, 16, 16, generator=gen), torch.randint(0, 10, (n,), generator=gen)


class MlpAdamW(_Base):
    """AdamW with two parameter groups, aggregation over A micro-batches, clip_grad_norm_."""
    NAME, AGG, DTYPE = "mlp_adamw_a2_clip", 2, "f32"
    TORCH_DTYPE = torch.float32

    def __init__(self, context):
        self.context = context
        net = torch.nn.Sequential(torch.nn.Linear(24, 96), torch.nn.LayerNorm(96), torch.nn.GELU(), torch.nn.Linear(96, 24)).to(self.TORCH_DTYPE)
        self.model = context.wrap_model(net)
        decay = [p for p in self.model.parameters() if p.ndim >= 2]
        rest = [p for p in self.model.parameters() if p.ndim < 2]
        self.optimizer = context.wrap_optimizer(torch.optim.AdamW(
            [{"params": decay, "weight_decay": 0.1}, {"params": rest, "weight_decay": 0.0}], lr=3e-3, betas=(0.9, 0.95)))

    @staticmethod
    def _clip(params):
        torch.nn.utils.clip_grad_norm_(params, 1.0)

    def train_batch(self, batch, epoch_idx, batch_idx):
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x.to(self.TORCH_DTYPE)).float(), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self._clip)
        return {"loss": loss}

    @classmethod
    def synthetic_batch(cls, n, gen):
        return torch.randn(n, 24, generator=gen), torch.randn(n, 24, generator=gen)


class MlpAdamWBf16(MlpAdamW):
    NAME, AGG, DTYPE = "mlp_adamw_bf16_a4_clip", 4, "bf16"
    TORCH_DTYPE = torch.bfloat16


def _run(trial_cls, mode):
    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": trial_cls.PER_SLOT_BATCH}, aggregation_frequency=trial_cls.AGG,
                                      device=torch.device("cuda", 0))
    return P.live_parity(trial_cls.NAME, trial_cls, ctx, torch.device("cuda", 0), 0, 1, steps=10, mode=mode)


@pytest.mark.parametrize("trial_cls", [ConvBnSgd, MlpAdamW])
def test_fp32_one_step_error_is_elementwise_within_1e5(lib, trial_cls):
    case = _run(trial_cls, "synced")
    assert case["ok"] and case["within_rtol"], case
    assert case["max_rel"] <= 1e-5 and case["max_frac_outside_elementwise"] == 0.0, case
    if case["max_norm_rel"] is not None:
        assert case["max_norm_rel"] <= 1e-5 and case["max_grad_rel"] <= 1e-5, case


@pytest.mark.parametrize("trial_cls", [ConvBnSgd, MlpAdamW])
def test_fp32_ten_free_running_steps_stay_within_1e5(lib, trial_cls):
    """north_star's 'over 10 steps': these two small fp32 trials are not chaotic over ten steps, so the plain bound holds."""
    case = _run(trial_cls, "free")
    assert case["ok"] and case["max_rel"] <= 1e-5 and case["max_loss_rel"] <= 1e-5, case


@pytest.mark.parametrize("mode", ["synced", "free"])
def test_bf16_stays_within_the_reference_twin(lib, mode):
    case = _run(MlpAdamWBf16, mode)
    assert case["ok"], case
    # the loss trajectory and the norm are within 1e-2 outright
    assert case["max_loss_rel"] <= 1e-2, case
    if mode == "synced":
        assert case["max_norm_rel"] <= 1e-2, case
