"""In-bench parity checker: the same PyTorchTrial-style workload driven for a few optimizer steps through this
repo's PyTorchTrialContext (libdgs kernels + communicator) and, beside it on the SAME devices and inputs, through the
torch port of the reference's context lines (oracle/torch_port.RefContext: torch DDP + torch.optim +
clip_grad_norm_, reference harness/determined/pytorch/_pytorch_context.py:295-296,718-924).

CHECKER ONLY (bench.py's `parity` object and tests): nothing under determined_clone_b200/ imports this module.  It
exists because the driver's `pytest -m gpu` box has one GPU: the N >= 2 evidence has to come from the run that has
N GPUs, which is the scaling bench.

What is compared after every optimizer step, on every rank (max over ranks is reported):
  * every entry of model.state_dict() -- parameters AND module buffers (batch-norm running statistics are what DDP
    re-broadcasts before each forward; integer buffers such as num_batches_tracked must be equal exactly);
  * the step's loss on this rank.
Two modes:
  free   both arms evolve independently for `steps` optimizer steps from identical initial weights (north_star's
         "over 10 steps");
  synced before every optimizer step the reference arm's parameters, buffers and optimizer state are copied into
         ours, so each comparison is the error of ONE step of the path on identical inputs (no chaotic growth through
         the network).
Error measure per floating tensor: max |ours - ref| / max |ref|  (tensor-relative), and the fraction of elements
outside |ours - ref| <= rtol * |ref| + rtol * rms(ref)  (element-wise with an rms floor).  rtol is north_star's:
1e-5 for fp32, 1e-2 for bf16.
"""
import copy
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

from oracle import torch_port


class DdpRefContext(torch_port.RefContext):
    """RefContext whose wrap_model does what the reference's does with several slots: DistributedDataParallel
    around the model on this rank's device (reference _pytorch_context.py:295-296, default DDP arguments)."""

    def __init__(self, world: int, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._world = world

    def wrap_model(self, model: torch.nn.Module) -> torch.nn.Module:
        model = model.to(self.device)
        if self._world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model)
        self.models.append(model)
        return model


def _plain(model: torch.nn.Module) -> torch.nn.Module:
    return model.module if isinstance(model, torch.nn.parallel.DistributedDataParallel) else model


def _tensor_err(a: torch.Tensor, b: torch.Tensor, rtol: float) -> Tuple[float, float]:
    """(max|a-b| / max|b|, fraction of elements outside rtol*|b| + rtol*rms(b)); integer tensors: (0 or inf, 0 or 1)."""
    if not b.dtype.is_floating_point:
        same = bool(torch.equal(a, b))
        return (0.0 if same else float("inf")), (0.0 if same else 1.0)
    a, b = a.detach().double(), b.detach().double()
    if b.numel() == 0:
        return 0.0, 0.0
    diff = (a - b).abs()
    scale = float(b.abs().max())
    rel = float(diff.max()) / scale if scale > 0 else float(diff.max())
    rms = float(b.pow(2).mean().sqrt())
    bad = float((diff > rtol * b.abs() + rtol * rms + 1e-300).double().mean())
    if not (rel == rel):  # NaN on either side
        rel, bad = float("inf"), 1.0
    return rel, bad


def _copy_arm(src_model: torch.nn.Module, src_opt: torch.optim.Optimizer, dst_model: torch.nn.Module,
              dst_opt: torch.optim.Optimizer) -> None:
    """Reference arm -> ours, in place (parameters and optimizer state of ours live in the engine's flat buffers)."""
    with torch.no_grad():
        for (_, d), (_, s) in zip(_plain(dst_model).state_dict().items(), _plain(src_model).state_dict().items()):
            d.copy_(s)
        sp = [p for g in src_opt.param_groups for p in g["params"]]
        dp = [p for g in dst_opt.param_groups for p in g["params"]]
        for s, d in zip(sp, dp):
            for k, v in src_opt.state.get(s, {}).items():
                cur = dst_opt.state.get(d, {}).get(k)
                if torch.is_tensor(v) and torch.is_tensor(cur):
                    cur.copy_(v)


def live_parity(name: str, trial_cls: Any, ours_ctx: Any, device: torch.device, rank: int, world: int, steps: int = 10,
                per_slot_batch: Optional[int] = None, mode: str = "free", reduce_max: Optional[Callable[[List[float]], List[float]]] = None,
                seed: int = 17) -> Dict[str, Any]:
    """Run `steps` optimizer steps of trial_cls under both arms and compare.  `ours_ctx` is a fresh
    PyTorchTrialContext of this repo (aggregation_frequency = trial_cls.AGG); reduce_max maps a list of floats to
    their max over ranks."""
    agg = int(getattr(trial_cls, "AGG", 1))
    batch = int(per_slot_batch or trial_cls.PER_SLOT_BATCH)
    rtol = 1e-2 if trial_cls.DTYPE == "bf16" else 1e-5
    det_before = torch.backends.cudnn.deterministic
    torch.backends.cudnn.deterministic = True  # both arms must run the same conv algorithms, bit-reproducibly
    ours_trial = ref_trial = None
    try:
        torch.manual_seed(seed)
        ours_trial = trial_cls(ours_ctx)
        torch.manual_seed(seed)
        ref_ctx = DdpRefContext(world, aggregation_frequency=agg, device=str(device))
        ref_trial = trial_cls(ref_ctx)
        # identical start: rank 0's weights everywhere (ours broadcast them at wrap_optimizer, DDP in its constructor)
        with torch.no_grad():
            for (k, d), (_, s) in zip(_plain(ref_trial.model).state_dict().items(), _plain(ours_trial.model).state_dict().items()):
                d.copy_(s)
        gen = torch.Generator().manual_seed(1234 + rank)
        n_micro = steps * agg
        data = [tuple(t.to(device) for t in trial_cls.synthetic_batch(batch, gen)) for _ in range(min(n_micro, 8))]
        per_step: List[Dict[str, float]] = []
        worst_rel, worst_bad, worst_loss, worst_what = 0.0, 0.0, 0.0, ""
        for i in range(n_micro):
            b = data[i % len(data)]
            boundary = (i + 1) % agg == 0
            if mode == "synced" and i % agg == 0:
                _copy_arm(ref_trial.model, ref_trial.optimizer, ours_trial.model, ours_trial.optimizer)
            ref_ctx._current_batch_idx = i
            ours_ctx._current_batch_idx = i
            lr = ref_trial.train_batch(b, 0, i)["loss"].detach().float()
            lo = ours_trial.train_batch(b, 0, i)["loss"].detach().float()
            loss_rel = float((lo - lr).abs() / lr.abs().clamp_min(1e-30))
            worst_loss = max(worst_loss, loss_rel if loss_rel == loss_rel else float("inf"))
            if not boundary:
                continue
            step_rel, step_bad = 0.0, 0.0
            so, sr = _plain(ours_trial.model).state_dict(), _plain(ref_trial.model).state_dict()
            for k in sr:
                rel, bad = _tensor_err(so[k], sr[k], rtol)
                if rel > step_rel:
                    step_rel = rel
                    if rel > worst_rel:
                        worst_what = k
                step_bad = max(step_bad, bad)
            worst_rel, worst_bad = max(worst_rel, step_rel), max(worst_bad, step_bad)
            per_step.append({"max_rel": step_rel, "frac_outside": step_bad, "loss_rel": loss_rel})
        # every rank must end with identical parameters (integer check: exact)
        flat = torch.cat([p.detach().reshape(-1).float() for p in _plain(ours_trial.model).parameters()])
        replicas_equal = 1.0
        if world > 1:
            import torch.distributed as dist

            ref0 = flat.clone()
            dist.broadcast(ref0, 0)
            replicas_equal = 1.0 if bool(torch.equal(ref0, flat)) else 0.0
        vals = [worst_rel, worst_bad, worst_loss, -replicas_equal] + [s["max_rel"] for s in per_step]
        if reduce_max is not None:
            vals = reduce_max(vals)
        worst_rel, worst_bad, worst_loss, neg_eq = vals[:4]
        ok = bool(worst_rel <= rtol and worst_loss <= rtol and neg_eq == -1.0)
        return {"case": name, "mode": mode, "optimizer_steps": steps, "aggregation_frequency": agg, "dtype": trial_cls.DTYPE,
                "rtol": rtol, "max_rel": worst_rel, "max_frac_outside_elementwise": worst_bad, "max_loss_rel": worst_loss,
                "replicas_identical": neg_eq == -1.0, "worst_tensor": worst_what, "max_rel_per_step": vals[4:], "ok": ok}
    finally:
        torch.backends.cudnn.deterministic = det_before
        ours_ctx.close()
        del ours_trial, ref_trial
