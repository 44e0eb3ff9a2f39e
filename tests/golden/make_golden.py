#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REAL reference harness on CPU.

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py [scenario ...]

What runs: the unmodified reference package imported from /root/reference/harness
(determined.pytorch.PyTorchTrialContext, _PyTorchTrialController, torch DDP on gloo for slots > 1),
driven the way the reference's own tests drive it
(harness/tests/experiment/pytorch_utils.py:16-120 and :223-241).  Two tiny shims stand in for
pure-Python dependencies that are not installed (ruamel.yaml -> PyYAML, appdirs); they are written
to a temp dir, never into the reference tree.

What is recorded, per scenario of golden_model.SCENARIOS (all ranks):
  init_params            parameters after wrap_model (rank 0)
  raw_grads[r][m]        every rank's raw gradient of every micro-batch (tensor hook)
  avg_grads[s]           gradients handed to clip_grads, i.e. after DDP averaging and
                         _average_gradients (reference _pytorch_context.py:818-825, :902)
  total_norm[s]          value returned by torch.nn.utils.clip_grad_norm_ (norm scenarios)
  clipped_grads[s]       gradients after the user's clip function
  params[s], state[s]    parameters / optimizer state after step_optimizer
  loss[r][m]             per-rank per-micro-batch loss
  reported               chief's (steps_completed, avg_metrics, batch_metrics) as passed to
                         core.train.report_training_metrics (integers are exact)
"""
import os
import pathlib
import sys
import tempfile
from typing import Any, Dict, List

import numpy as np

HERE = pathlib.Path(__file__).resolve().parent
REFERENCE = pathlib.Path("/root/reference/harness")


def _install_shims() -> None:
    d = pathlib.Path(tempfile.mkdtemp(prefix="det_shims"))
    (d / "ruamel").mkdir()
    (d / "ruamel" / "__init__.py").write_text("")
    (d / "ruamel" / "yaml.py").write_text(
        "import yaml as _y\nsafe_load = _y.safe_load\n"
        "class YAML:\n"
        "    def __init__(self, typ=None, pure=False): pass\n"
        "    def load(self, s): return _y.safe_load(s)\n"
        "    def dump(self, d, f=None): return _y.safe_dump(d, f)\n"
    )
    (d / "appdirs.py").write_text(
        "import os\ndef user_data_dir(name=None, *a, **k): return os.path.join('/tmp', 'appdirs', name or 'x')\n"
    )
    sys.dont_write_bytecode = True
    sys.path[:0] = [str(d), str(REFERENCE), str(HERE)]


def _to_np(t: Any) -> np.ndarray:
    import torch

    t = t.detach().cpu()
    if t.dtype == torch.bfloat16:
        # store the exact bf16 values widened to fp32
        t = t.float()
    return t.numpy().copy()


def run_rank(name: str, rank: int, size: int, out_dir: str) -> None:
    _install_shims()
    import torch

    import determined as det
    from determined import core, pytorch
    from golden_model import SCENARIOS, TRIAL_SEED, GoldenNet, make_data, make_optimizer

    torch.set_num_threads(1)
    sc = SCENARIOS[name]
    hp, agg, steps = dict(sc["hp"]), sc["agg"], sc["steps"]
    dtype = getattr(torch, sc["dtype"])
    rec: Dict[str, Any] = {"raw": [], "avg": [], "norm": [], "clipped": [], "params": [], "state": [], "loss": []}

    class RefGoldenTrial(pytorch.PyTorchTrial):
        _searcher_metric = "val_loss"

        def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
            self.context = context
            self.model = context.wrap_model(GoldenNet().to(dtype))
            self.plist = list(self.model.parameters())
            rec["init"] = [_to_np(p) for p in self.plist]
            self.opt = context.wrap_optimizer(make_optimizer(self.model, hp))
            self._mb: List[Any] = [None] * len(self.plist)
            for i, p in enumerate(self.plist):
                p.register_hook(lambda g, i=i: self._mb.__setitem__(i, _to_np(g)))
            x, y = make_data(steps * agg, hp["global_batch_size"])
            self.ds = torch.utils.data.TensorDataset(x.to(dtype), y.to(dtype))

        def build_training_data_loader(self) -> pytorch.DataLoader:
            return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

        def build_validation_data_loader(self) -> pytorch.DataLoader:
            return pytorch.DataLoader(self.ds, batch_size=self.context.get_per_slot_batch_size(), shuffle=False)

        def _clip(self, params: Any) -> None:
            params = list(params)
            rec["avg"].append([_to_np(p.grad) for p in params])
            if sc["clip"] is not None:
                kind, val = sc["clip"]
                if kind == "norm":
                    rec["norm"].append(_to_np(torch.nn.utils.clip_grad_norm_(params, val)))
                else:
                    torch.nn.utils.clip_grad_value_(params, val)
            rec["clipped"].append([_to_np(p.grad) for p in params])

        def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
            x, y = batch
            loss = torch.nn.functional.mse_loss(self.model(x).float(), y.float())
            self.context.backward(loss)
            rec["raw"].append(list(self._mb))
            rec["loss"].append(float(loss))
            self.context.step_optimizer(self.opt, clip_grads=self._clip)
            if (batch_idx + 1) % agg == 0:
                rec["params"].append([_to_np(p) for p in self.plist])
                st = []
                for p in self.plist:
                    s = self.opt.state[p]
                    st.append({k: (_to_np(v) if torch.is_tensor(v) else np.asarray(v)) for k, v in s.items()})
                rec["state"].append(st)
            return {"loss": loss}

        def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
            x, y = batch
            return {"val_loss": torch.nn.functional.mse_loss(self.model(x).float(), y.float())}

    if size > 1:
        os.environ.update(
            RANK=str(rank), WORLD_SIZE=str(size), LOCAL_RANK=str(rank), LOCAL_WORLD_SIZE=str(size),
            GROUP_RANK="0", GROUP_WORLD_SIZE="1", MASTER_ADDR="127.0.0.1", MASTER_PORT="29411",
            USE_TORCH_DISTRIBUTED="1",
        )
    backend = det._DistributedBackend()
    distributed = core.DistributedContext.from_torch_distributed() if backend.use_torch() else None
    core_context = core._dummy_init(distributed=distributed, checkpoint_storage=out_dir)
    core_context.preempt.start()
    core_context.train._trial_id = "1"
    reported: List[Any] = []
    real_report = core_context.train.report_training_metrics

    def report(steps_completed: int, metrics: Dict[str, Any], batch_metrics: Any = None) -> None:
        reported.append((steps_completed, dict(metrics), [dict(b) for b in (batch_metrics or [])]))
        real_report(steps_completed=steps_completed, metrics=metrics, batch_metrics=batch_metrics)

    core_context.train.report_training_metrics = report  # type: ignore

    exp_config = {
        "scheduling_unit": agg,
        "resources": {"native_parallel": False, "slots_per_trial": size},
        "hyperparameters": hp,
        "optimizations": {"aggregation_frequency": agg, "gradient_compression": False,
                          "average_training_metrics": True, "average_aggregated_gradients": bool(sc.get("average", True))},
        "checkpoint_policy": "none",
        "perform_initial_validation": False,
        "checkpoint_storage": {"type": "shared_fs", "host_path": out_dir},
        "searcher": {"metric": "val_loss", "smaller_is_better": True},
        "min_checkpoint_period": {"batches": 0},
        "min_validation_period": {"batches": 0},
        "data": {},
    }
    pytorch._PyTorchTrialController.pre_execute_hook(TRIAL_SEED, backend)
    context = pytorch.PyTorchTrialContext(
        core_context=core_context, trial_seed=TRIAL_SEED, hparams=hp, slots_per_trial=size, num_gpus=0,
        exp_conf=exp_config, aggregation_frequency=agg, steps_completed=0, managed_training=True,
        debug_enabled=False, enable_tensorboard_logging=False,
    )
    context._set_default_gradient_compression(False)
    context._set_default_average_aggregated_gradients(bool(sc.get("average", True)))
    trial = RefGoldenTrial(context)
    controller = pytorch._PyTorchTrialController(
        trial_inst=trial, context=context, max_length=pytorch.Batch(steps * agg),
        checkpoint_period=pytorch.Batch(sys.maxsize), validation_period=pytorch.Batch(sys.maxsize),
        searcher_metric_name="val_loss", reporting_period=pytorch.Batch(agg), local_training=True,
        latest_checkpoint=None, steps_completed=0, smaller_is_better=True, test_mode=False,
        checkpoint_policy="none", step_zero_validation=False, det_profiler=None, global_batch_size=None,
    )
    controller.run()

    out: Dict[str, np.ndarray] = {}
    n_p = len(rec["init"])
    for i in range(n_p):
        out[f"init_params/{i}"] = rec["init"][i]
    for m, grads in enumerate(rec["raw"]):
        for i in range(n_p):
            out[f"raw_grads/{m}/{i}"] = grads[i]
    out["loss"] = np.asarray(rec["loss"], dtype=np.float64)
    for s in range(len(rec["avg"])):
        for i in range(n_p):
            out[f"avg_grads/{s}/{i}"] = rec["avg"][s][i]
            out[f"clipped_grads/{s}/{i}"] = rec["clipped"][s][i]
            out[f"params/{s}/{i}"] = rec["params"][s][i]
            for k, v in rec["state"][s][i].items():
                out[f"state/{s}/{i}/{k}"] = v
    if rec["norm"]:
        out["total_norm"] = np.asarray(rec["norm"], dtype=np.float32)
    if rank == 0:
        out["reported_steps"] = np.asarray([r[0] for r in reported], dtype=np.int64)
        out["reported_avg_loss"] = np.asarray([r[1]["loss"] for r in reported], dtype=np.float64)
        out["reported_batch_loss"] = np.asarray([[b["loss"] for b in r[2]] for r in reported], dtype=np.float64)
    np.savez(os.path.join(out_dir, f"{name}.rank{rank}.npz"), **out)


def generate(name: str) -> None:
    import torch.multiprocessing as mp

    sys.path.insert(0, str(HERE))
    from golden_model import SCENARIOS

    size = SCENARIOS[name]["slots"]
    with tempfile.TemporaryDirectory() as tmp:
        ctx = mp.get_context("spawn")
        procs = [ctx.Process(target=run_rank, args=(name, r, size, tmp)) for r in range(size)]
        for p in procs:
            p.start()
        for p in procs:
            p.join()
            if p.exitcode != 0:
                raise SystemExit(f"scenario {name}: a rank exited with {p.exitcode}")
        merged: Dict[str, np.ndarray] = {}
        for r in range(size):
            with np.load(os.path.join(tmp, f"{name}.rank{r}.npz")) as z:
                for k in z.files:
                    if k.startswith(("raw_grads/", "loss")):
                        merged[f"rank{r}/{k}"] = z[k]
                    elif r == 0:
                        merged[k] = z[k]
                    else:
                        # replicas must agree exactly after DDP
                        if k.startswith(("params/", "avg_grads/")) and not np.array_equal(merged[k], z[k]):
                            raise SystemExit(f"{name}: rank {r} disagrees with rank 0 on {k}")
        np.savez_compressed(HERE / f"{name}.npz", **merged)
    print(f"wrote {HERE / (name + '.npz')}  ({(HERE / (name + '.npz')).stat().st_size} bytes)")


if __name__ == "__main__":
    sys.path.insert(0, str(HERE))
    from golden_model import SCENARIOS

    for nm in sys.argv[1:] or list(SCENARIOS):
        generate(nm)
