#!/usr/bin/env python
"""Benchmark of the per-step data-parallel path (BASELINE.json metric: samples/sec at 1/2/4/8 B200;
grad-sync us/step and % of the HBM/NVLink roofline).

    python bench.py --gpus N --steps K --warmup W            # this repo's path, one process per GPU
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (torch port), host cores

Workload (config.workload): BASELINE.json configs[1], the CIFAR-10 ResNet PyTorchTrial, per-slot batch
128 (weak scaling: global batch 128*N), fp32 SGD+momentum, synthetic data.  One step = one train_batch:
forward, context.backward (flatten/scale/cast kernels + bucket allreduce on the NCCL side stream),
context.step_optimizer (fused step kernel).

  value  device-timed: inputs already resident in HBM, CUDA events around exactly K steps, max over ranks
  e2e    the same trial driven through the public API (_PyTorchTrialController.run): every step copies
         its batch from pinned host memory and reads the step's metrics back to the host
  roofline  the fused optimizer kernel (largest HBM mover of the path), CUDA events recorded inside libdgs
            immediately around each launch (dgs_timer_*) during the timed region, on the launching stream;
            algorithmic bytes per element as stated in DESIGN.md
  gradsync_microbench  BASELINE.json configs[4]: one flat 1 GiB fp32 bucket, accumulate + allreduce +
            clip + AdamW, time per step against max(HBM, NVLink) roofline (N > 1: also the opt-in sharded step)
  cpu_baseline  (N=1, rank 0) the torch port of the reference's path on the host cores, bounded sample
"""
import argparse
import ctypes
import json
import os
import pathlib
import socket
import sys
import threading
import time
from typing import Any, Dict, List, Optional, Tuple

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

PER_SLOT_BATCH = 128
# dram__bytes_read.sum + dram__bytes_write.sum of one sgd_kernel launch in this workload, from the ncu --set full capture
# profiles/r1_prof_bench_r1_summary.csv (134.19 MB read + 33.32 MB written; the rest of the 89 MB of writes is
# still in L2 when the kernel ends)
NCU_TRAFFIC_BYTES = 167.5e6
METRIC, UNIT = "samples_per_sec", "samples/s"


def parse() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference", "torch"], default="ours",
                    help="ours: this repo; reference: the reference's CPU path; torch: the reference's path on the GPUs "
                         "(torch DDP on NCCL + torch.optim), the number to beat, for comparison only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-microbench", action="store_true")
    ap.add_argument("--micro-mb", type=int, default=1024)
    ap.add_argument("--micro-only", action="store_true", help="run only the grad-sync microbench (both step variants) and print it")
    ap.add_argument("--micro-bucket-mb", type=int, nargs="*", default=[0], help="bucket caps to try with --micro-only (0 = adaptive)")
    ap.add_argument("--workload", default="cifar_resnet18", choices=["mnist", "cifar_resnet18", "resnet50", "gpt2_medium"],
                    help="cifar_resnet18 = BASELINE.json configs[1] (the default the driver runs); the others are configs[2]/[3]")
    return ap.parse_args()


def measured_peaks() -> Tuple[float, str]:
    try:
        return float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- clocks ---------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU every 20 ms while the timed region runs."""

    def __init__(self, index: int) -> None:
        self.index, self.samples, self.reasons = index, [], set()
        self.max_mhz: Optional[int] = None
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None

    def _loop(self) -> None:
        nv = self._nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self) -> "ClockSampler":
        if self._nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *_: Any) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)

    def summary(self) -> Dict[str, Any]:
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---- reference arm / cpu baseline -----------------------------------------------------------------
def run_cpu_port(steps: int, warmup: int, budget_s: float, workload: str = "cifar_resnet18") -> Dict[str, Any]:
    """The reference's per-batch path as it runs on CPU (oracle/torch_port.py), all host threads."""
    import torch

    from oracle import torch_port
    from workloads.trials import WORKLOADS

    CifarResNetTrial = WORKLOADS[workload]
    PER_SLOT_BATCH = CifarResNetTrial.PER_SLOT_BATCH

    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    torch.manual_seed(17)
    ctx = torch_port.RefContext(aggregation_frequency=CifarResNetTrial.AGG, device="cpu")
    trial = CifarResNetTrial(ctx)
    gen = torch.Generator().manual_seed(1234)
    batches = [CifarResNetTrial.synthetic_batch(PER_SLOT_BATCH, gen) for _ in range(4)]
    idx = 0
    for _ in range(max(warmup, 1)):
        t0 = time.perf_counter()
        torch_port.train_batch_like_reference(ctx, trial, batches[idx % 4], idx)
        idx += 1
        one = time.perf_counter() - t0
    # bounded sample: as many of the requested steps as fit the budget
    n = max(1, min(steps, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        torch_port.train_batch_like_reference(ctx, trial, batches[idx % 4], idx)
        idx += 1
    dt = time.perf_counter() - t0
    return {"value": n * PER_SLOT_BATCH / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} train_batch steps of {CifarResNetTrial.NAME}, batch {PER_SLOT_BATCH}, torch {torch.__version__} CPU, "
                      f"{cores} threads", "ms_per_step": dt / n * 1e3, "steps": n}


def run_reference(args: argparse.Namespace) -> None:
    if int(os.environ.get("RANK", "0")) != 0:
        return  # the CPU path is timed once, on rank 0
    from workloads.trials import WORKLOADS

    W = WORKLOADS[args.workload]
    PER_SLOT_BATCH = W.PER_SLOT_BATCH
    r = run_cpu_port(args.steps, args.warmup, budget_s=150.0, workload=args.workload)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
        "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": W.DTYPE, "data": "synthetic",
        "config": {"workload": W.NAME, "per_slot_batch": PER_SLOT_BATCH, "global_batch": PER_SLOT_BATCH,
                   "optimizer": W.OPTIMIZER, "aggregation_frequency": W.AGG, "parallelism": "cpu, 1 process"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- our arm ------------------------------------------------------------------------------------------
def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def gradsync_microbench(ctx: Any, mb: int, steps: int, peak_gbs: float, sharded: bool = False, bucket_mb: int = 0) -> Dict[str, Any]:
    """configs[4]: one flat fp32 bucket through the path: flatten*scale -> allreduce -> L2 norm -> clip+AdamW."""
    import torch

    from determined_clone_b200 import pytorch

    pieces = 64  # the flat bucket is made of 64 equal tensors, as a model's would be of many
    n_piece = mb * (1 << 20) // 4 // pieces
    n = n_piece * pieces
    world = ctx.distributed.size
    params = [torch.nn.Parameter(torch.randn(n_piece, device=ctx.device) * 0.02) for _ in range(pieces)]
    opt = torch.optim.AdamW(params, lr=3e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    eng = pytorch.GradSyncEngine(opt, ctx.device, ctx._get_comm(), aggregation_frequency=1, average_aggregated_gradients=True,
                                 sharded_step=sharded, bucket_bytes=bucket_mb << 20)
    grads = [[torch.randn(n_piece, device=ctx.device) * 1e-2 for _ in range(pieces)] for _ in range(2)]
    clip = pytorch._gradsync.ClipRequest()
    clip.max_norm = 1.0
    st = torch.cuda.current_stream()

    def step(i: int) -> None:
        eng.begin_pass(final=True)
        for p, g in zip(params, grads[i % 2]):
            p.grad = g  # what autograd leaves behind when backward produces this gradient
        eng.end_pass()
        present = eng.finish_window()
        eng.total_norm()
        eng.fused_step(present, clip)
        eng.reset_window()

    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(steps):
        step(i)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        import torch.distributed as dist

        t = torch.tensor([ms], device=ctx.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    hbm_bytes = (8 + 4 + 28) * n  # flatten (read g, write wire) + norm (read wire) + AdamW (read g,p,m,v; write p,m,v)
    t_hbm = hbm_bytes / (peak_gbs * 1e9)
    t_nvl = (2 * (world - 1) / world * 4 * n) / 770e9 if world > 1 else 0.0  # measured peer bandwidth per direction
    eng.close()
    return {"bucket_mib": mb, "dtype": "f32", "optimizer": "adamw", "aggregation_frequency": 1, "clip": 1.0, "sharded_step": eng.sharded,
            "us_per_step": ms * 1e3, "hbm_bytes_per_step": hbm_bytes, "roofline_us": max(t_hbm, t_nvl) * 1e6,
            "frac_of_roofline": max(t_hbm, t_nvl) / (ms * 1e-3), "bound": "nvlink" if t_nvl > t_hbm else "hbm"}


def run_ours(args: argparse.Namespace) -> None:
    import torch

    from determined_clone_b200 import _build, _lib, core, pytorch
    from workloads.trials import WORKLOADS

    CifarResNetTrial = WORKLOADS[args.workload]  # the trial class of the selected workload
    PER_SLOT_BATCH = CifarResNetTrial.PER_SLOT_BATCH
    AGG = CifarResNetTrial.AGG
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the grad-sync path has no CPU implementation (use --impl reference)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist  # type: ignore

        dist.init_process_group("nccl", device_id=device)
    if rank == 0 and not _lib.LIB_PATH.exists():
        _build.build()
    if dist is not None:
        dist.barrier()
    _lib.load()

    # control plane for the trial: the reference's DistributedContext (ZeroMQ through the chief)
    if world > 1:
        ports = [_free_port(), _free_port()] if rank == 0 else [0, 0]
        dist.broadcast_object_list(ports, src=0)
        dctx = core.DistributedContext(rank=rank, size=world, local_rank=local_rank, local_size=world, cross_rank=0, cross_size=1,
                                       pub_port=ports[0], pull_port=ports[1])
    else:
        dctx = core.DummyDistributedContext()

    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": PER_SLOT_BATCH * world}, slots_per_trial=world,
                                      aggregation_frequency=AGG, trial_seed=17, device=device)

    class BenchTrial(CifarResNetTrial):
        loader: Any = None

        def build_training_data_loader(self) -> Any:
            return self.loader

        def build_validation_data_loader(self) -> Any:
            return None

    if args.micro_only:
        peak, _ = measured_peaks()
        out = []
        for cap in args.micro_bucket_mb:
            for sharded in ((False, True) if world > 1 else (False,)):
                m = gradsync_microbench(ctx, args.micro_mb, 12, peak, sharded=sharded, bucket_mb=cap)
                m["bucket_cap_mib"] = cap or "adaptive"
                out.append(m)
                torch.cuda.empty_cache()
        if rank == 0:
            print(json.dumps({"n_gpus": world, "gradsync_microbench": out}), flush=True)
        ctx.close()
        dctx.close()
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    trial = BenchTrial(ctx)
    eng = ctx._engines[trial.optimizer]
    p_total = sum(fs.total for fs in eng.sets)

    gen = torch.Generator().manual_seed(1234 + rank)
    host_batches = [tuple(t.pin_memory() for t in CifarResNetTrial.synthetic_batch(PER_SLOT_BATCH, gen)) for _ in range(8)]
    dev_batches = [tuple(t.to(device) for t in b) for b in host_batches]
    stream = torch.cuda.current_stream()

    def one_step(i: int) -> Any:
        ctx._current_batch_idx = i
        return trial.train_batch(dev_batches[i % len(dev_batches)], 0, i)

    def barrier() -> None:
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed region: inputs resident in HBM ----------------------------------------------
    for i in range(args.warmup):
        one_step(i)
    barrier()
    eng.enable_step_timing(True)
    launches0 = _lib.calls.dgs_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        e0.record(stream)
        for i in range(args.warmup, args.warmup + args.steps):
            one_step(i)
        e1.record(stream)
        barrier()
    launches = _lib.calls.dgs_launch_count() - launches0
    ms_total = e0.elapsed_time(e1)
    step_kernel_ms = eng.step_timings_ms()
    eng.enable_step_timing(False)
    if dist is not None:
        t = torch.tensor([ms_total], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t)
        lt = torch.tensor([launches], device=device, dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt)
    value = args.steps * PER_SLOT_BATCH * world / (ms_total * 1e-3)

    # ---- exposed grad-sync tail: from the end of backward's last kernel to the end of the step kernel ----
    tail_us = None
    try:
        ev = []
        orig = ctx.step_optimizer

        def timed_step_optimizer(*a: Any, **k: Any) -> None:
            if not ctx._should_communicate_and_update():
                return orig(*a, **k)  # not an optimizer step: nothing to time
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(stream)
            orig(*a, **k)
            a1.record(stream)
            ev.append((a0, a1))

        ctx.step_optimizer = timed_step_optimizer  # type: ignore
        for i in range(2 * AGG + 8 - (8 % AGG if AGG > 1 else 0)):
            one_step(args.warmup + args.steps + i)
        torch.cuda.synchronize()
        ctx.step_optimizer = orig  # type: ignore
        tails = sorted(a.elapsed_time(b) for a, b in ev)
        tail_us = tails[len(tails) // 2] * 1e3
    except Exception:
        tail_us = None

    # ---- end to end through the public API: host batches in, metrics out, every step ----------------------
    e2e_steps = args.steps

    class HostLoader:
        def __iter__(self) -> Any:
            i = 0
            while True:
                yield host_batches[i % len(host_batches)]
                i += 1

    trial.loader = HostLoader()
    reports: List[Any] = []
    base = (args.warmup + args.steps + 4 * AGG + 16) // AGG * AGG  # resume on an aggregation boundary
    warm = pytorch._PyTorchTrialController(trial, ctx, max_length=base + AGG, reporting_period=1, steps_completed=base)
    warm.run()
    ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=base + AGG + e2e_steps, reporting_period=1, steps_completed=base + AGG,
                                          on_report=lambda n, r: reports.append(r))
    barrier()
    t0 = time.perf_counter()
    ctl.run()
    barrier()
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_s], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t)
    h2d = sum(t.numel() * t.element_size() for t in host_batches[0])
    d2h = 1 * 1 * 8 + 8 + 1 + 8  # per-batch mean (1x1 f64) + period mean (f64) + validity byte + int64 count
    e2e = {"value": e2e_steps * PER_SLOT_BATCH * world / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": e2e_s / e2e_steps * 1e3, "api": "_PyTorchTrialController.run(), reporting_period=1, pinned host batches"}

    # ---- roofline of the dominant kernel of the path (fused SGD step) --------------------------------
    peak, peak_src = measured_peaks()
    k_ms = sorted(step_kernel_ms)[len(step_kernel_ms) // 2] if step_kernel_ms else None
    bytes_per_elem = CifarResNetTrial.STEP_BYTES_PER_ELEM  # DESIGN.md, kernel table
    roofline = None
    if k_ms:
        achieved = bytes_per_elem * p_total / (k_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": CifarResNetTrial.STEP_KERNEL, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": NCU_TRAFFIC_BYTES if args.workload == "cifar_resnet18" else None, "peak_source": peak_src, "elements": p_total,
                    "algorithmic_bytes_per_element": bytes_per_elem, "us_per_launch": k_ms * 1e3,
                    "note": "median over the timed region of CUDA events recorded inside libdgs around each launch (dgs_timer_read); for "
                            "models smaller than L2 the wire bucket just written by the flatten kernel is partly still cached"}

    micro = None
    if not args.no_microbench:
        # one bucket = the whole buffer: the microbench has no backward pass for smaller buckets to overlap with
        micro = gradsync_microbench(ctx, args.micro_mb, 12, peak, bucket_mb=args.micro_mb)
        if world > 1:  # the opt-in reduce-scatter / sharded-step / all-gather variant of the same step
            torch.cuda.empty_cache()
            micro["sharded_variant"] = gradsync_microbench(ctx, args.micro_mb, 12, peak, sharded=True, bucket_mb=args.micro_mb)

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        r = run_cpu_port(steps=20, warmup=1, budget_s=20.0, workload=args.workload)
        cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": CifarResNetTrial.DTYPE,
            "data": "synthetic",
            "config": {"workload": CifarResNetTrial.NAME, "per_slot_batch": PER_SLOT_BATCH, "global_batch": PER_SLOT_BATCH * world,
                       "optimizer": CifarResNetTrial.OPTIMIZER, "aggregation_frequency": AGG, "parallelism": f"dp{world}",
                       "l2": "8 distinct input batches cycled; forward/backward activations between optimizer steps exceed L2",
                       "param_elements": p_total},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks.summary(),
            "grad_sync_us_per_step": tail_us, "gradsync_microbench": micro,
        }
        print(json.dumps(line), flush=True)
    ctx.close()
    dctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_torch_gpu(args: argparse.Namespace) -> None:
    """The reference's per-step path on the same GPUs: torch DDP (NCCL) + torch.optim, through the torch port of
    the reference context.  Not part of the driver contract; prints a line with impl 'torch' for comparison."""
    import torch

    from oracle import torch_port
    from workloads.trials import WORKLOADS

    CifarResNetTrial = WORKLOADS[args.workload]
    PER_SLOT_BATCH = CifarResNetTrial.PER_SLOT_BATCH
    rank, world, local_rank = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist  # type: ignore

        dist.init_process_group("nccl", device_id=device)
    torch.manual_seed(17)

    class Ctx(torch_port.RefContext):
        def wrap_model(self, model: Any) -> Any:
            model = model.to(self.device)
            if world > 1:
                model = torch.nn.parallel.DistributedDataParallel(model)  # reference _pytorch_context.py:295-296
            self.models.append(model)
            return model

    ctx = Ctx(aggregation_frequency=CifarResNetTrial.AGG, device=str(device))
    trial = CifarResNetTrial(ctx)
    gen = torch.Generator().manual_seed(1234 + rank)
    batches = [tuple(t.to(device) for t in CifarResNetTrial.synthetic_batch(PER_SLOT_BATCH, gen)) for _ in range(8)]

    def step(i: int) -> None:
        ctx._current_batch_idx = i
        m = trial.train_batch(batches[i % 8], 0, i)
        m["loss"].cpu()  # the reference reads every metric back every batch (_pytorch_trial.py:920-927)

    for i in range(args.warmup):
        step(i)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.warmup, args.warmup + args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    if rank == 0:
        print(json.dumps({"impl": "torch", "metric": METRIC, "value": args.steps * PER_SLOT_BATCH * world / (ms * 1e-3), "unit": UNIT,
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                          "config": {"workload": CifarResNetTrial.NAME, "parallelism": f"torch DDP dp{world}"}}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    # stdout carries exactly one JSON line: keep NCCL's "NCCL version ..." banner (printed to stdout when the image
    # sets NCCL_DEBUG=VERSION) out of it; an explicit NCCL_DEBUG=INFO from the caller is respected
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "torch":
        run_torch_gpu(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
