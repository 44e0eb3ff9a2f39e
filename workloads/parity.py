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
Reference self-noise ("twin"): a third arm runs the reference's lines again on an input that differs from the
reference arm's by the least a rounding can express -- the micro-batches of every accumulation window in reverse
order (A > 1: the same mathematical sum), or initial parameters moved by one ulp (A = 1) -- and is compared with the
reference arm the same way.  Where that disagreement already exceeds rtol (a chaotic configuration; Adam's first
steps on zero-initialised 16-bit biases, where the update is lr * sign(g) and the sign of a near-zero 16-bit
gradient is rounding noise) no implementation other than a bit-identical one can be within rtol of the reference,
and the case passes when this path is no further from the reference than 1.5 x (per-step synced mode; 4 x when
free-running, where both distances are draws from a chaotic process) the reference is from its own twin.
Both numbers are always reported.

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


def _compare_models(ours: torch.nn.Module, ref: torch.nn.Module, rtol: float) -> Tuple[float, float, str]:
    worst, bad, what = 0.0, 0.0, ""
    so, sr = _plain(ours).state_dict(), _plain(ref).state_dict()
    for k in sr:
        rel, b = _tensor_err(so[k], sr[k], rtol)
        if rel > worst:
            worst, what = rel, k
        bad = max(bad, b)
    return worst, bad, what


def live_parity(name: str, trial_cls: Any, ours_ctx: Any, device: torch.device, rank: int, world: int, steps: int = 10,
                per_slot_batch: Optional[int] = None, mode: str = "free", reduce_max: Optional[Callable[[List[float]], List[float]]] = None,
                seed: int = 17, twin: bool = True) -> Dict[str, Any]:
    """Run `steps` optimizer steps of trial_cls under both arms (and the reference's twin) and compare.  `ours_ctx` is
    a fresh PyTorchTrialContext of this repo (aggregation_frequency = trial_cls.AGG); reduce_max maps a list of
    floats to their max over ranks."""
    agg = int(getattr(trial_cls, "AGG", 1))
    batch = int(per_slot_batch or trial_cls.PER_SLOT_BATCH)
    rtol = 1e-2 if trial_cls.DTYPE == "bf16" else 1e-5
    det_before = torch.backends.cudnn.deterministic
    torch.backends.cudnn.deterministic = True  # both arms must run the same conv algorithms, bit-reproducibly
    ours_trial = ref_trial = twin_trial = None
    try:
        torch.manual_seed(seed)
        ours_trial = trial_cls(ours_ctx)
        torch.manual_seed(seed)
        ref_ctx = DdpRefContext(world, aggregation_frequency=agg, device=str(device))
        ref_trial = trial_cls(ref_ctx)
        arms = [(ref_ctx, ref_trial)]
        if twin:
            torch.manual_seed(seed)
            twin_ctx = DdpRefContext(world, aggregation_frequency=agg, device=str(device))
            twin_trial = trial_cls(twin_ctx)
            arms.append((twin_ctx, twin_trial))
        # identical start: rank 0's weights everywhere (ours broadcast them at wrap_optimizer, DDP in its constructor)
        with torch.no_grad():
            for _, t in arms:
                for (k, d), (_, s) in zip(_plain(t.model).state_dict().items(), _plain(ours_trial.model).state_dict().items()):
                    d.copy_(s)
            if twin and agg == 1:
                for p in _plain(twin_trial.model).parameters():
                    p.copy_(torch.nextafter(p, torch.full_like(p, float("inf"))))
        gen = torch.Generator().manual_seed(1234 + rank)
        n_micro = steps * agg
        data = [tuple(t.to(device) for t in trial_cls.synthetic_batch(batch, gen)) for _ in range(min(n_micro, 8))]
        # gradients as the clip function sees them (synchronized, averaged) and the total norm, where the trial clips
        seen: Dict[int, List[Any]] = {0: [], 1: [], 2: []}
        if hasattr(trial_cls, "_clip"):
            def recorder(slot: int, inner: Callable) -> Callable:
                def clip(params: Any) -> None:
                    params = list(params)
                    # sliced routes: a rank holds only its 1/N of the averaged gradients, there is nothing whole to read
                    sliced = slot == 0 and any(e.sharded for e in ours_ctx._engines.values())
                    grads = None if sliced else [p.grad.detach().float().clone() for p in params]
                    norm = torch.nn.utils.clip_grad_norm_(params, 1.0)
                    seen[slot].append((grads, float(norm)))
                return clip
            for slot, t in enumerate([ours_trial] + [t for _, t in arms]):
                t._clip = recorder(slot, t._clip)
        worst = {"rel": 0.0, "bad": 0.0, "loss": 0.0, "what": "", "grad": 0.0, "norm": 0.0}
        noise = {"rel": 0.0, "bad": 0.0, "loss": 0.0, "grad": 0.0, "norm": 0.0}
        per_step: List[float] = []
        noise_per_step: List[float] = []
        for w in range(steps):
            if mode == "synced":
                _copy_arm(ref_trial.model, ref_trial.optimizer, ours_trial.model, ours_trial.optimizer)
                if twin:
                    _copy_arm(ref_trial.model, ref_trial.optimizer, twin_trial.model, twin_trial.optimizer)
                    if agg == 1:
                        with torch.no_grad():
                            for p in _plain(twin_trial.model).parameters():
                                p.copy_(torch.nextafter(p, torch.full_like(p, float("inf"))))
            for m in range(agg):
                i = w * agg + m
                b = data[i % len(data)]
                ref_ctx._current_batch_idx = ours_ctx._current_batch_idx = i
                lr = ref_trial.train_batch(b, 0, i)["loss"].detach().float()
                lo = ours_trial.train_batch(b, 0, i)["loss"].detach().float()
                loss_rel = float((lo - lr).abs() / lr.abs().clamp_min(1e-30))
                worst["loss"] = max(worst["loss"], loss_rel if loss_rel == loss_rel else float("inf"))
                if twin:
                    j = w * agg + (agg - 1 - m)  # the window's micro-batches in reverse order
                    twin_ctx._current_batch_idx = i
                    lt = twin_trial.train_batch(data[j % len(data)], 0, i)["loss"].detach().float()
                    if agg == 1:
                        noise["loss"] = max(noise["loss"], float((lt - lr).abs() / lr.abs().clamp_min(1e-30)))
            rel, bad, what = _compare_models(ours_trial.model, ref_trial.model, rtol)
            if rel > worst["rel"]:
                worst["what"] = what
            worst["rel"], worst["bad"] = max(worst["rel"], rel), max(worst["bad"], bad)
            per_step.append(rel)
            if twin:
                rel, bad, _ = _compare_models(twin_trial.model, ref_trial.model, rtol)
                noise["rel"], noise["bad"] = max(noise["rel"], rel), max(noise["bad"], bad)
                noise_per_step.append(rel)
            if seen[0] and seen[1]:
                (go, no), (gr, nr) = seen[0][-1], seen[1][-1]
                if go is not None:
                    worst["grad"] = max(worst["grad"], max(_tensor_err(a, b_, rtol)[0] for a, b_ in zip(go, gr)))
                worst["norm"] = max(worst["norm"], abs(no - nr) / max(abs(nr), 1e-30))
                if twin and seen[2]:
                    gt, nt = seen[2][-1]
                    noise["grad"] = max(noise["grad"], max(_tensor_err(a, b_, rtol)[0] for a, b_ in zip(gt, gr)))
                    noise["norm"] = max(noise["norm"], abs(nt - nr) / max(abs(nr), 1e-30))
                for v in seen.values():
                    v.clear()
        # every rank must end with identical parameters (integer check: exact)
        flat = torch.cat([p.detach().reshape(-1).float() for p in _plain(ours_trial.model).parameters()])
        replicas_equal = 1.0
        if world > 1:
            import torch.distributed as dist

            ref0 = flat.clone()
            dist.broadcast(ref0, 0)
            replicas_equal = 1.0 if bool(torch.equal(ref0, flat)) else 0.0
        head = [worst["rel"], worst["bad"], worst["loss"], worst["grad"], worst["norm"], -replicas_equal,
                noise["rel"], noise["bad"], noise["loss"], noise["grad"], noise["norm"]]
        vals = head + per_step + noise_per_step
        if reduce_max is not None:
            vals = reduce_max(vals)
        w_rel, w_bad, w_loss, w_grad, w_norm, neg_eq, n_rel, n_bad, n_loss, n_grad, n_norm = vals[:len(head)]
        per_step = vals[len(head):len(head) + steps]
        noise_per_step = vals[len(head) + steps:]

        # free-running arms are two draws from the same chaotic process: the margin over the twin is wider there
        margin = 1.5 if mode == "synced" else 4.0

        def within(x: float, floor: float) -> bool:
            return bool(x <= rtol or (twin and x <= margin * floor))

        # 16-bit gradients: the reference accumulates micro-batches in the parameter dtype, so its own result moves by
        # about 2e-2 of the tensor's largest element when the micro-batches arrive in another order (the twin); this
        # path accumulates in fp32.  Gradients and norm are therefore judged like the parameters.
        strict = False
        checks = {
            "parameters_and_buffers": within(w_rel, n_rel),
            "loss": within(w_loss, n_loss) if agg == 1 else bool(w_loss <= rtol or within(w_rel, n_rel)),
            "gradients_seen_by_clip": bool(w_grad <= rtol) if strict else within(w_grad, n_grad),
            "total_norm": bool(w_norm <= rtol) if strict else within(w_norm, n_norm),
            "replicas_identical": neg_eq == -1.0,
        }
        routes = sorted({e.route_name() for e in ours_ctx._engines.values()})
        return {"case": name, "mode": mode, "route_at_end": routes[0] if len(routes) == 1 else routes, "optimizer_steps": steps, "aggregation_frequency": agg, "dtype": trial_cls.DTYPE,
                "rtol": rtol, "max_rel": w_rel, "max_frac_outside_elementwise": w_bad, "max_loss_rel": w_loss,
                "max_grad_rel": w_grad if hasattr(trial_cls, "_clip") else None, "max_norm_rel": w_norm if hasattr(trial_cls, "_clip") else None,
                "worst_tensor": worst["what"], "max_rel_per_step": per_step,
                "reference_self_noise": {"twin": ("micro-batches of each window reversed" if agg > 1 else "initial parameters + 1 ulp") if twin else None,
                                         "max_rel": n_rel, "max_frac_outside_elementwise": n_bad, "max_loss_rel": n_loss,
                                         "max_grad_rel": n_grad, "max_norm_rel": n_norm, "max_rel_per_step": noise_per_step},
                "within_rtol": bool(w_rel <= rtol and w_loss <= rtol), "checks": checks, "ok": all(checks.values())}
    finally:
        torch.backends.cudnn.deterministic = det_before
        ours_ctx.close()
        del ours_trial, ref_trial, twin_trial
