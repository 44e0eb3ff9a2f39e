"""CPU oracle for the per-step data-parallel path of Determined's PyTorch harness.

TEST INFRASTRUCTURE ONLY.  Nothing under determined_clone_b200/ may import this module; it is used
by tests/, by __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs, always
as the checker or the timed CPU baseline, never as the product path.

What it restates (paths relative to the reference tree, H/ = harness/determined/):
  * H/pytorch/_pytorch_context.py:711-716   _should_communicate_and_update
  * H/pytorch/_pytorch_context.py:802-816   backward: local `grad += new` under DDP.no_sync(), then
                                            DDP's grad * (1/N) + allreduce(sum) on the boundary pass
  * H/pytorch/_pytorch_context.py:818-825   _average_gradients: grad.div_(float(A))
  * H/pytorch/_pytorch_context.py:902       clip_grads(parameters) == torch.nn.utils.clip_grad_norm_ /
                                            clip_grad_value_
  * H/pytorch/_pytorch_context.py:914-924   optimizer.step() (+ zero_grad)
  * H/pytorch/_metric_utils.py:13-60, H/util.py:171-196, H/pytorch/_reducer.py:28-46  metric reduction

The numerical kernels the reference calls live in torch, a third-party dependency that is not under
/root/reference (pinned there as torch==1.11.0 in harness/tests/requirements/requirements-harness.txt:10;
this image has torch 2.11.0).  Their published algorithms are restated here op by op, with the same
rounding points (every torch op rounds its result to the tensor dtype):
  torch/nn/utils/clip_grad.py   _get_total_norm / _clip_grads_with_norm_ / clip_grad_value_
  torch/optim/sgd.py            _single_tensor_sgd
  torch/optim/adam.py           _single_tensor_adam (Adam and AdamW, amsgrad=False)

Parity is PINNED: tests/test_oracle_golden.py checks every function below against fixtures recorded
from the unmodified reference harness running in the build container (tests/golden/make_golden.py),
and against the known answers of the reference's own tests
(harness/tests/experiment/pytorch/test_reducer.py:18-29, test_metric_utils.py:13-86,
test_pytorch_trial.py:1604-1642).
"""
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

F32 = np.float32


# ---- dtype emulation ------------------------------------------------------------------------
def bf16_round(x: np.ndarray) -> np.ndarray:
    """Round fp32 values to the nearest bfloat16 (ties to even), returned widened to fp32."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    rounded = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    out = rounded.astype(np.uint32).view(np.float32)
    nan = np.isnan(x)
    if nan.any():
        out = np.where(nan, x, out)
    return out.reshape(x.shape)


def f16_round(x: np.ndarray) -> np.ndarray:
    return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)


def rnd(x: Any, dtype: str) -> np.ndarray:
    """Result of a torch op whose output tensor has `dtype`: fp32 math, one rounding."""
    x = np.asarray(x, dtype=np.float32)
    if dtype == "float32":
        return x
    if dtype == "bfloat16":
        return bf16_round(x)
    if dtype == "float16":
        return f16_round(x)
    raise ValueError(dtype)


# ---- gating -----------------------------------------------------------------------------------
def should_communicate_and_update(batch_idx: Optional[int], aggregation_frequency: int, managed: bool = True) -> bool:
    """H/pytorch/_pytorch_context.py:711-716."""
    if not managed:
        return True
    if batch_idx is None:
        raise RuntimeError("Training hasn't started.")
    return (batch_idx + 1) % aggregation_frequency == 0


# ---- gradient path ------------------------------------------------------------------------------
def accumulate(micro_grads: Sequence[np.ndarray], dtype: str = "float32") -> np.ndarray:
    """autograd AccumulateGrad across micro-batches: grad = g0; grad += g1; ... in the param dtype
    (H/pytorch/_pytorch_context.py:802-816)."""
    acc = rnd(micro_grads[0], dtype)
    for g in micro_grads[1:]:
        acc = rnd(acc + rnd(g, dtype), dtype)
    return acc


def ddp_allreduce_mean(per_rank: Sequence[np.ndarray], dtype: str = "float32") -> np.ndarray:
    """torch DDP on the boundary backward: each rank's bucket = grad * (1/N), then allreduce(sum)
    (the DistributedDataParallel created at H/pytorch/_pytorch_context.py:296)."""
    n = len(per_rank)
    if n == 1:
        return rnd(per_rank[0], dtype)
    inv = F32(1.0 / n)
    total = rnd(rnd(per_rank[0], dtype) * inv, dtype)
    for g in per_rank[1:]:
        total = rnd(total + rnd(rnd(g, dtype) * inv, dtype), dtype)
    return total


def average_gradients(grad: np.ndarray, divisor: int, dtype: str = "float32") -> np.ndarray:
    """H/pytorch/_pytorch_context.py:818-825: no-op for divisor 1, else grad.div_(float(divisor))."""
    if divisor == 1:
        return grad
    return rnd(grad / F32(float(divisor)), dtype)


def synced_gradients(raw: Sequence[Sequence[np.ndarray]], aggregation_frequency: int, average: bool = True,
                     dtype: str = "float32") -> np.ndarray:
    """raw[r][m] = rank r's gradient of micro-batch m of one window -> what clip_grads sees."""
    per_rank = [accumulate(micro, dtype) for micro in raw]
    g = ddp_allreduce_mean(per_rank, dtype)
    return average_gradients(g, aggregation_frequency, dtype) if average else g


def _tensor_norm(g: np.ndarray, dtype: str) -> np.float32:
    # torch._foreach_norm: fp32 accumulation, result in the tensor dtype
    s = np.sqrt(np.sum(np.square(g.astype(np.float64))))
    return rnd(F32(s), dtype)[()]


def clip_grad_norm(grads: Sequence[np.ndarray], max_norm: float, dtype: str = "float32") -> Tuple[np.float32, List[np.ndarray]]:
    """torch.nn.utils.clip_grad_norm_(params, max_norm) with norm_type 2 (torch/nn/utils/clip_grad.py):
    total = ||(||g_i||)_i||; coef = clamp(max_norm / (total + 1e-6), max=1); g_i *= coef."""
    norms = np.asarray([_tensor_norm(g, dtype) for g in grads], dtype=np.float32)
    total = rnd(F32(np.sqrt(np.sum(np.square(norms.astype(np.float64))))), dtype)[()]
    coef = rnd(F32(max_norm) / rnd(total + F32(1e-6), dtype), dtype)[()]
    coef = coef if not (coef > 1.0) else F32(1.0)
    return total, [rnd(g * coef, dtype) for g in grads]


def clip_grad_value(grads: Sequence[np.ndarray], clip_value: float) -> List[np.ndarray]:
    """torch.nn.utils.clip_grad_value_: clamp every element to [-clip_value, clip_value]."""
    c = F32(clip_value)
    return [np.clip(g, -c, c) for g in grads]


# ---- optimizers -----------------------------------------------------------------------------
def sgd_step(p: np.ndarray, g: np.ndarray, buf: Optional[np.ndarray], *, lr: float, momentum: float = 0.0,
             dampening: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False, maximize: bool = False,
             dtype: str = "float32") -> Tuple[np.ndarray, Optional[np.ndarray]]:
    """torch/optim/sgd.py _single_tensor_sgd.  `buf` None == no momentum buffer yet."""
    if maximize:
        g = -g
    if weight_decay != 0:
        g = rnd(g + F32(weight_decay) * p, dtype)
    if momentum != 0:
        if buf is None:
            buf = g.copy()
        else:
            buf = rnd(rnd(buf * F32(momentum), dtype) + F32(1 - dampening) * g, dtype)
        g = rnd(g + F32(momentum) * buf, dtype) if nesterov else buf
    p = rnd(p + F32(-lr) * g, dtype)
    return p, buf


def _lerp(a: np.ndarray, b: np.ndarray, w: float) -> np.ndarray:
    w32 = F32(w)
    if abs(w) < 0.5:
        return a + w32 * (b - a)
    return b - (b - a) * (F32(1) - w32)


def adam_step(p: np.ndarray, g: np.ndarray, m: np.ndarray, v: np.ndarray, step: int, *, lr: float,
              betas: Tuple[float, float], eps: float, weight_decay: float = 0.0, decoupled: bool = True,
              maximize: bool = False, dtype: str = "float32") -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """torch/optim/adam.py _single_tensor_adam, amsgrad=False.  `step` is the 1-based count of this
    update.  decoupled=True is AdamW."""
    b1, b2 = betas
    if maximize:
        g = -g
    if weight_decay != 0:
        if decoupled:
            p = rnd(p * F32(1 - lr * weight_decay), dtype)
        else:
            g = rnd(g + F32(weight_decay) * p, dtype)
    m = rnd(_lerp(m, g, 1 - b1), dtype)
    v = rnd(v * F32(b2), dtype)
    v = rnd(v + F32(1 - b2) * g * g, dtype)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    step_size = lr / bc1
    bc2_sqrt = bc2 ** 0.5
    denom = rnd(rnd(rnd(np.sqrt(v), dtype) / F32(bc2_sqrt), dtype) + F32(eps), dtype)
    p = rnd(p + F32(-step_size) * (m / denom), dtype)
    return p, m, v


# ---- one optimizer over several parameter tensors, many steps -----------------------------------
class OptimizerOracle:
    """State of one torch.optim.{SGD,Adam,AdamW} over a list of parameter tensors."""

    def __init__(self, kind: str, params: Sequence[np.ndarray], group_of: Sequence[int],
                 groups: Sequence[Dict[str, Any]], dtype: str = "float32") -> None:
        assert kind in ("sgd", "adam", "adamw")
        self.kind, self.dtype = kind, dtype
        self.params = [rnd(p, dtype).copy() for p in params]
        self.group_of, self.groups = list(group_of), [dict(g) for g in groups]
        self.steps = [0] * len(self.params)
        self.m = [np.zeros_like(p) for p in self.params]
        self.v = [np.zeros_like(p) for p in self.params]
        self.buf: List[Optional[np.ndarray]] = [None] * len(self.params)

    def step(self, grads: Sequence[Optional[np.ndarray]]) -> None:
        for i, g in enumerate(grads):
            if g is None:  # torch optimizers skip parameters without a gradient
                continue
            hp = self.groups[self.group_of[i]]
            if self.kind == "sgd":
                self.params[i], self.buf[i] = sgd_step(
                    self.params[i], g, self.buf[i], lr=hp["lr"], momentum=hp.get("momentum", 0.0),
                    dampening=hp.get("dampening", 0.0), weight_decay=hp.get("weight_decay", 0.0),
                    nesterov=hp.get("nesterov", False), maximize=hp.get("maximize", False), dtype=self.dtype)
            else:
                self.steps[i] += 1
                self.params[i], self.m[i], self.v[i] = adam_step(
                    self.params[i], g, self.m[i], self.v[i], self.steps[i], lr=hp["lr"], betas=tuple(hp["betas"]),
                    eps=hp["eps"], weight_decay=hp.get("weight_decay", 0.0), decoupled=self.kind == "adamw",
                    maximize=hp.get("maximize", False), dtype=self.dtype)


def train_window(opt: OptimizerOracle, raw: Sequence[Sequence[Sequence[np.ndarray]]], aggregation_frequency: int,
                 clip: Optional[Tuple[str, float]] = None, average: bool = True) -> Dict[str, Any]:
    """One optimizer step of step_optimizer (H/pytorch/_pytorch_context.py:827-924).
    raw[r][m][i]: rank r, micro-batch m of this window, parameter tensor i."""
    n_params = len(opt.params)
    avg = [synced_gradients([[raw[r][m][i] for m in range(len(raw[r]))] for r in range(len(raw))],
                            aggregation_frequency, average, opt.dtype) for i in range(n_params)]
    total = None
    clipped = avg
    if clip is not None:
        kind, val = clip
        if kind == "norm":
            total, clipped = clip_grad_norm(avg, val, opt.dtype)
        else:
            clipped = clip_grad_value(avg, val)
    opt.step(clipped)
    return {"avg_grads": avg, "total_norm": total, "clipped_grads": clipped}


# ---- metrics ------------------------------------------------------------------------------------
def process_combined_metrics_and_batches(combined: Sequence[Tuple[Dict[str, Any], int]]) -> Tuple[Dict[str, List[Any]], List[int]]:
    """H/pytorch/_metric_utils.py:13-26: drop ranks that report 0 batches, transpose to name -> per-rank list."""
    kept = [c for c in combined if c[1]]
    names = list(kept[0][0].keys())
    return {k: [c[0][k] for c in kept] for k in names}, [c[1] for c in kept]


def average_training_metrics(combined_timeseries: Dict[str, Sequence[Sequence[Any]]], combined_num_batches: Sequence[int]) -> List[Dict[str, Any]]:
    """H/pytorch/_metric_utils.py:29-60: per metric and batch, the mean over ranks of the non-None
    values; 0-d arrays stay 0-d arrays."""
    num_batches = combined_num_batches[0]
    n_proc = len(combined_num_batches)
    out: List[Dict[str, Any]] = [{} for _ in range(num_batches)]
    for name, per_proc in combined_timeseries.items():
        is_array = isinstance(per_proc[0][0], np.ndarray)
        for b in range(num_batches):
            vals = np.array([per_proc[r][b] for r in range(n_proc)])
            mean = np.mean(vals[vals != None])  # noqa: E711
            out[b][name] = np.array(mean) if is_array else mean
    return out


def make_metrics(batch_metrics: Sequence[Dict[str, Any]]) -> Dict[str, Any]:
    """H/util.py:171-196: avg_metrics[name] = mean of the non-None per-batch values, None if the
    values are not scalars."""
    names: List[str] = []
    for d in batch_metrics:
        for k in d:
            if k not in names:
                names.append(k)
    avg: Dict[str, Optional[float]] = {}
    for k in names:
        try:
            vals = np.array([d[k] for d in batch_metrics if k in d])
            avg[k] = np.mean(vals[vals != None]).item()  # noqa: E711
        except (TypeError, ValueError):
            avg[k] = None
    return {"batch_metrics": list(batch_metrics), "avg_metrics": avg}


AVG, SUM, MAX, MIN = 1, 2, 3, 4  # Reducer enum, H/pytorch/_reducer.py:9-25


def simple_reduce_metrics(reducer: int, metrics: np.ndarray, num_batches: Optional[Sequence[int]] = None) -> np.float64:
    """H/pytorch/_reducer.py:28-46."""
    if reducer == AVG:
        if num_batches and len(metrics) != len(num_batches):
            raise RuntimeError(
                f"Lengths of metrics and num_batches are not equal: {len(metrics)} != {len(num_batches)}."
            )
        return np.average(metrics, weights=num_batches)
    if reducer == SUM:
        return np.sum(metrics)
    if reducer == MAX:
        return np.max(metrics)
    if reducer == MIN:
        return np.min(metrics)
    raise NotImplementedError


def identity_expected_weights(batch_size: int = 4, epoch_size: int = 64, num_epochs: int = 3, lr: float = 0.001) -> List[float]:
    """Closed form of the reference's 2-slot x aggregation_frequency=2 identity experiment: y = w*x
    trained with MSE and SGD, weights reported after every batch
    (harness/tests/experiment/pytorch/test_pytorch_trial.py:1604-1642)."""
    weights: List[float] = []
    w = 0.0
    held: List[float] = []
    for i, start in enumerate(range(0, epoch_size * num_epochs, batch_size)):
        xs = [((start % epoch_size) + k) * 0.1 + 1.0 for k in range(batch_size)]
        if i % 2 == 0:
            held = xs
        else:
            both = held + xs
            w = w + 2.0 * lr * sum(x * (x - x * w) for x in both) / len(both)
        weights.append(w)
    return weights
