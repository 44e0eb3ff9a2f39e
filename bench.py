#!/usr/bin/env python
"""Benchmark of the per-step data-parallel path (BASELINE.json metric: samples/sec at 1/2/4/8 B200;
grad-sync us/step and % of the HBM/NVLink roofline).

    python bench.py --gpus N --steps K --warmup W                    # this repo's path, one process per GPU
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path: N gloo ranks, host cores

Workload (config.workload): BASELINE.json configs[2], the ResNet-50 PyTorchTrial -- the config the metric is written
on ("slots_per_trial=1/2/4/8") -- per-slot batch 64 (weak scaling: global batch 64*N), synthetic 224x224 batches,
fp32 SGD + momentum.  One step = one train_batch: forward, context.backward (flatten/scale/cast kernels + bucket
collectives on the communicator's side stream), context.step_optimizer (fused step kernel).

  value     device-timed: inputs already resident in HBM, CUDA events around exactly K steps, max over ranks
  e2e       the same trial driven through the public API (_PyTorchTrialController.run): every step copies its batch
            from pinned host memory and reads the step's metrics back to the host
  roofline  the fused optimizer kernel (largest HBM mover of the path; 511 MB per launch, beyond the 126 MB L2),
            CUDA events recorded inside libdgs immediately around each launch (dgs_timer_*) during the timed
            region, on the launching stream; algorithmic bytes per element as stated in DESIGN.md; `traffic` is
            the ncu --set full capture of that kernel in this workload (profiles/, file named in the object)
  parity    (after the headline measurement) N >= 2 (and N = 1): the workloads of configs[1..3] for 10 optimizer steps under this path and, beside
            it on the same GPUs and inputs, under torch DDP + torch.optim (the reference's lines,
            workloads/parity.py); at N >= 2 also the 2-slot fixtures recorded from the unmodified reference on
            ranks 0-1.  A mismatch makes the run exit non-zero.
  gradsync_microbench  BASELINE.json configs[4]: flat buckets of 1 MB .. 2 GB, fp32 and bf16, A in {1, 4}:
            accumulate + collective + clip + AdamW, time per optimizer step against max(HBM, NVLink @ 900 GB/s)
  configs   configs[3] (GPT-2 medium bf16, AdamW, A=4, clip 1.0) measured the same way as the headline
  cpu_baseline  (N=1, rank 0) the torch port of the reference's path on the host cores, bounded sample

Order and time limits: headline (value, roofline, e2e) and cpu_baseline first; parity, configs and the microbench are
additional evidence -- a failure there is recorded under `errors`, none of them is started later than 560 s into the
run, and at 780 s the line is printed with whatever is complete (_Deadline; the driver's scaling sweep allows each N
about 870 s).  A part that is cut or fails reports the cases / rows it had finished (`incomplete`).  `wall_s` says when
each part was complete.
"""
import argparse
import json
import os
import pathlib
import socket
import sys
import threading
import time
from typing import Any, Callable, Dict, List, Optional, Tuple

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC, UNIT = "samples_per_sec", "samples/s"
NVLINK_GBS = 900.0  # per direction per GPU
# ncu --set full captures of the step kernels inside these workloads (dram__bytes_read.sum + dram__bytes_write.sum
# per launch); None until a capture of THIS round's binary is committed under profiles/
NCU_TRAFFIC: Dict[str, Dict[str, Any]] = {}
try:
    NCU_TRAFFIC = json.loads((ROOT / "profiles" / "ncu_traffic.json").read_text())
except Exception:
    pass


def parse(argv: Optional[List[str]] = None) -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference", "torch"], default="ours",
                    help="ours: this repo; reference: the reference's CPU path (N gloo ranks); torch: the reference's path on "
                         "the GPUs (torch DDP on NCCL + torch.optim), the number to beat, for comparison only")
    ap.add_argument("--workload", default="resnet50", choices=["mnist", "cifar_resnet18", "resnet50", "gpt2_medium"],
                    help="resnet50 = BASELINE.json configs[2] (the default the driver runs)")
    ap.add_argument("--per-slot-batch", type=int, default=0, help="override the workload's per-slot batch (tests)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-microbench", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs[3] (GPT-2 medium) sub-measurement")
    ap.add_argument("--parity-cases", nargs="*", default=["cifar_resnet18", "resnet50", "gpt2_medium"])
    ap.add_argument("--parity-steps", type=int, default=10)
    ap.add_argument("--micro-mb", type=int, nargs="*", default=[1, 4, 16, 64, 256, 1024, 2048])
    ap.add_argument("--micro-dtypes", nargs="*", default=["f32", "bf16"], choices=["f32", "bf16"])
    ap.add_argument("--micro-aggs", type=int, nargs="*", default=[1, 4])
    ap.add_argument("--micro-only", action="store_true", help="run only the grad-sync microbench and print it")
    ap.add_argument("--micro-bucket-mb", type=int, default=0, help="bucket cap for the microbench (0 = one bucket per flat set)")
    ap.add_argument("--micro-steps", type=int, default=10)
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    return ap.parse_args(argv)


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


# ---- device plumbing (CUDA; the CPU test suite swaps in host stand-ins to exercise this file's control flow) -------
class _Cuda:
    """The few torch.cuda calls bench.py makes.  tests/test_bench_host.py replaces this object with a host-memory
    stand-in (together with tests/fake_dgs.py for the C ABI) so that the driver's command cannot rot unnoticed;
    the product and every reported number use this class."""

    emulated = False
    backend = "nccl"

    def available(self) -> bool:
        import torch

        return torch.cuda.is_available()

    def device(self, local_rank: int) -> Any:
        import torch

        torch.cuda.set_device(local_rank)
        return torch.device("cuda", local_rank)

    def sync(self) -> None:
        import torch

        torch.cuda.synchronize()

    def event(self) -> Any:
        import torch

        return torch.cuda.Event(enable_timing=True)

    def stream(self) -> Any:
        import torch

        return torch.cuda.current_stream()

    def pin(self, t: Any) -> Any:
        return t.pin_memory()

    def empty_cache(self) -> None:
        import torch

        torch.cuda.empty_cache()


DEV: Any = _Cuda()


def _workload(args: argparse.Namespace, name: Optional[str] = None) -> Any:
    from workloads.trials import WORKLOADS

    cls = WORKLOADS[name or args.workload]
    if args.per_slot_batch and (name is None or name == args.workload):
        cls = type(cls.__name__, (cls,), {"PER_SLOT_BATCH": args.per_slot_batch})
    return cls


# ---- reference arm / cpu baseline -----------------------------------------------------------------
def run_cpu_port(trial_cls: Any, steps: int, warmup: int, budget_s: float, world: int = 1, rank: int = 0,
                 threads: Optional[int] = None) -> Dict[str, Any]:
    """The reference's per-batch path as it runs on CPU (oracle/torch_port.py; with world > 1: torch DDP over gloo,
    the process group the reference creates without GPUs, harness/determined/pytorch/_pytorch_trial.py:296-300)."""
    import torch

    from oracle import torch_port
    from workloads.parity import DdpRefContext

    batch = trial_cls.PER_SLOT_BATCH
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    threads = threads or max(cores // world, 1)
    torch.set_num_threads(threads)
    torch.manual_seed(17)
    ctx = DdpRefContext(world, aggregation_frequency=trial_cls.AGG, device="cpu")
    trial = trial_cls(ctx)
    gen = torch.Generator().manual_seed(1234 + rank)
    batches = [trial_cls.synthetic_batch(batch, gen) for _ in range(4)]
    agg = trial_cls.AGG
    idx = 0

    def agree(x: float, op: str = "max") -> float:
        if world == 1:
            return x
        import torch.distributed as dist

        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.MIN)
        return float(t)

    one = 0.0
    warm_t0 = time.perf_counter()
    for w in range(max(warmup, 1)):
        for _ in range(agg):
            t0 = time.perf_counter()
            torch_port.train_batch_like_reference(ctx, trial, batches[idx % 4], idx)
            idx += 1
            one = time.perf_counter() - t0
        # the warm-up is bounded like the sample (N CPU ranks of ResNet-50 take tens of seconds per step); every rank
        # must stop at the same window, so they agree on the clock
        if agree(time.perf_counter() - warm_t0) > budget_s / 4:
            break
    # bounded sample: as many of the requested steps as fit the budget (whole aggregation windows; all ranks agree)
    n = max(agg, min(steps, int(budget_s / max(agree(one), 1e-3))) // agg * agg)
    n = int(agree(float(n), "min"))
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(n):
        torch_port.train_batch_like_reference(ctx, trial, batches[idx % 4], idx)
        idx += 1
    if world > 1:
        dist.barrier()
    dt = agree(time.perf_counter() - t0)
    return {"value": n * batch * world / dt, "unit": UNIT, "cores": cores, "threads_per_rank": threads, "ranks": world, "kind": "port",
            "sample": f"{n} train_batch steps of {trial_cls.NAME}, per-slot batch {batch}, {world} rank(s) x {threads} threads"
                      f"{' over gloo DDP' if world > 1 else ''}, torch {torch.__version__} CPU",
            "ms_per_step": dt / n * 1e3, "steps": n}


def _reference_rank(rank: int, world: int, port: int, argv: List[str]) -> None:
    """One gloo rank of the reference arm (spawned when bench.py itself was not started under torchrun)."""
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    run_reference(parse(argv))


def run_reference(args: argparse.Namespace) -> None:
    """`--impl reference`: the reference's CPU path for the same config -- N processes over gloo (torch DDP), the
    GPU arm's per-slot batch, OMP threads = floor(cores / N) -- a bounded sample on the host cores."""
    import torch

    W = _workload(args)
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world_env == 1:
        import torch.multiprocessing as mp

        mp.spawn(_reference_rank, args=(args.gpus, _free_port(), sys.argv[1:]), nprocs=args.gpus, join=True)
        return
    rank, world = int(os.environ.get("RANK", "0")), world_env
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("gloo")
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    r = run_cpu_port(W, args.steps, args.warmup, budget_s=90.0, world=world, rank=rank, threads=max(cores // world, 1))
    if rank == 0:
        line = {
            "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": W.DTYPE, "data": "synthetic",
            "config": {"workload": W.NAME, "per_slot_batch": W.PER_SLOT_BATCH, "global_batch": W.PER_SLOT_BATCH * world,
                       "optimizer": W.OPTIMIZER, "aggregation_frequency": W.AGG,
                       "parallelism": f"cpu, {world} gloo rank(s) x {r['threads_per_rank']} threads, torch DDP" if world > 1
                       else f"cpu, 1 process x {r['threads_per_rank']} threads"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "ranks", "threads_per_rank")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()



# ---- time limits ----------------------------------------------------------------------------------------
T0 = time.monotonic()  # the process's own clock; the driver allows each `--gpus N` run about 870 s in its scaling sweep


class _Deadline:
    """Keeps a measurement that has been taken from being lost to whatever comes after it.

    soft (DGS_BENCH_SOFT_S, 560 s after start): no further optional part (parity, configs[3], microbench) is started;
    rank 0 decides and the ranks follow its decision, so they stay in step.
    hard (DGS_BENCH_HARD_S, 780 s after start): a timer thread on every rank.  If the headline has been measured by
    then, rank 0 prints the JSON line with the parts that are complete (the cut is named under `errors`) and every
    rank leaves with os._exit -- a collective that never returns (a rank that died in an optional part) cannot be
    waited out.  Before the headline exists there is nothing to print: the run exits non-zero."""

    def __init__(self, rank: int, emit: Callable[[Optional[str]], None]) -> None:
        self.rank, self.emit = rank, emit
        self.soft = float(os.environ.get("DGS_BENCH_SOFT_S", "560"))
        self.hard = float(os.environ.get("DGS_BENCH_HARD_S", "780"))
        self.phase, self.rc, self.have_headline, self.printed = "start", 0, False, False
        self.lock = threading.Lock()
        self._exit: Callable[[int], None] = os._exit
        self._timer: Optional[threading.Timer] = None

    def arm(self) -> None:
        self._timer = threading.Timer(max(self.hard - (time.monotonic() - T0), 0.0), self.expire)
        self._timer.daemon = True
        self._timer.start()

    def disarm(self) -> None:
        if self._timer is not None:
            self._timer.cancel()

    def elapsed(self) -> float:
        return time.monotonic() - T0

    def print_once(self, cut: Optional[str] = None) -> None:
        with self.lock:
            if self.printed:
                return
            self.printed = True
            if self.rank == 0:
                self.emit(cut)

    def expire(self) -> None:
        if not self.have_headline:
            sys.stderr.write(f"bench.py: no headline after {self.elapsed():.0f} s (in: {self.phase}); giving up\n")
            sys.stderr.flush()
            self._exit(4)
            return
        self.print_once(f"time limit: stopped after {self.elapsed():.0f} s in `{self.phase}`; the parts that were complete are reported")
        if self.rank != 0:
            time.sleep(1.0)  # let rank 0 write its line before the launcher sees a worker leave
        sys.stdout.flush()
        self._exit(self.rc)


# ---- our arm ------------------------------------------------------------------------------------------
def _free_port() -> int:
    """A port for the control plane's listeners, taken BELOW the kernel's ephemeral range (32768+): a port the kernel
    hands out for bind(0) can be given to any outgoing connection of the other ranks (NCCL / gloo bootstrap) between
    this probe and the moment the chief binds it -- seen once on an 8-GPU run as 'Address already in use'."""
    import random

    rng = random.Random(os.getpid() ^ int(time.time() * 1e3))
    for _ in range(200):
        port = rng.randrange(20000, 29000)
        try:
            with socket.socket() as s:
                s.bind(("0.0.0.0", port))
            return port
        except OSError:
            continue
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _free_ports(n: int) -> List[int]:
    out: List[int] = []
    while len(out) < n:
        port = _free_port()
        if port not in out:
            out.append(port)
    return out


def _max_over_ranks(dist: Any, device: Any, vals: List[float]) -> List[float]:
    if dist is None:
        return vals
    import torch

    t = torch.tensor(vals, device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


# algorithmic HBM bytes per element per optimizer step, DESIGN.md section 3 (AdamW + clip_grad_norm_)
MICRO_BYTES = {  # several GPUs: the averaged gradient has to exist in memory (the wire buffer) to be exchanged
    ("f32", 1): 8 + 4 + 28, ("f32", 4): 8 + 12 + 12 + 12 + 4 + 28, ("bf16", 1): 4 + 2 + 14, ("bf16", 4): 6 + 10 + 10 + 8 + 2 + 14,
}
MICRO_BYTES_1GPU = {  # one GPU, A = 1: gradients read once for the norm and once by the step (dgs_direct_step_*), no
    # wire buffer; with an accumulator (A > 1) the three-kernel path runs, as on several GPUs
    ("f32", 1): 4 + 28, ("f32", 4): 8 + 12 + 12 + 12 + 4 + 28, ("bf16", 1): 2 + 14, ("bf16", 4): 6 + 10 + 10 + 8 + 2 + 14,
}


def gradsync_microbench(ctx: Any, dist: Any, mb: int, dtype: str, agg: int, steps: int, peak_gbs: float, bucket_mb: int = 0,
                        route: Optional[str] = None) -> Dict[str, Any]:
    """configs[4]: one flat parameter set of `mb` MiB through the whole path per optimizer step: `agg` micro-batches of
    flatten/accumulate, scale 1/(N*A) + cast, bucket collective, global L2 norm, clip + AdamW."""
    import torch

    from determined_clone_b200 import pytorch

    tdt = torch.float32 if dtype == "f32" else torch.bfloat16
    isz = 4 if dtype == "f32" else 2
    pieces = 64 if mb >= 4 else 8  # the flat set is made of equal tensors, as a model's would be of many
    n_piece = mb * (1 << 20) // isz // pieces
    n = n_piece * pieces
    world = ctx.distributed.size
    dev = ctx.device
    params = [torch.nn.Parameter((torch.randn(n_piece, device=dev) * 0.02).to(tdt)) for _ in range(pieces)]
    opt = torch.optim.AdamW(params, lr=3e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    kwargs: Dict[str, Any] = {}
    if route is not None:
        kwargs["route"] = route
    eng = pytorch.GradSyncEngine(opt, dev, ctx._get_comm(), aggregation_frequency=agg, average_aggregated_gradients=True,
                                 bucket_bytes=(bucket_mb << 20) if bucket_mb else n * isz + 1, **kwargs)
    grads = [[(torch.randn(n_piece, device=dev) * 1e-2).to(tdt) for _ in range(pieces)] for _ in range(2)]
    clip = pytorch._gradsync.ClipRequest()
    clip.max_norm = 1.0
    st = DEV.stream()

    marks: List[List[Any]] = []

    def step(i: int, timed: bool = False) -> None:
        ev = [DEV.event() for _ in range(5)] if timed else None
        if ev:
            ev[0].record(st)
        for m in range(agg):
            eng.begin_pass(final=(m == agg - 1))
            for p, g in zip(params, grads[(i + m) % 2]):
                p.grad = g  # what autograd leaves behind when backward produces this gradient
            eng.end_pass()
        if ev:
            ev[1].record(st)
        if eng.direct and eng.can_direct(True):
            present = eng.finish_window(allow_direct=True)
            if ev:
                ev[2].record(st)
                ev[3].record(st)
            eng.direct_step(present, 1.0)  # one GPU: flatten + norm + clip + step in one launch
        else:
            present = eng.finish_window()
            if ev:
                ev[2].record(st)
            eng.total_norm()
            if ev:
                ev[3].record(st)
            eng.fused_step(present, clip)
        if ev:
            ev[4].record(st)
            marks.append(ev)
        eng.reset_window()

    for i in range(3):
        step(i)
    DEV.sync()
    if dist is not None:
        dist.barrier()
    e0, e1 = DEV.event(), DEV.event()
    e0.record(st)
    for i in range(steps):
        step(i)
    e1.record(st)
    DEV.sync()
    ms = _max_over_ranks(dist, dev, [e0.elapsed_time(e1) / steps])[0]
    # where the time goes (two more steps with events between the phases; not part of the timed region above)
    for i in range(2):
        step(i, timed=True)
    DEV.sync()
    names = ["flatten_all_passes", "finish_window(exchange)", "norm", "step(+param exchange; one GPU, A=1: the single launch)"]
    phases = {nm: round(min(ev[k].elapsed_time(ev[k + 1]) for ev in marks) * 1e3, 1) for k, nm in enumerate(names)}
    hbm_bytes = (MICRO_BYTES if world > 1 else MICRO_BYTES_1GPU)[(dtype, agg)] * n
    t_hbm = hbm_bytes / (peak_gbs * 1e9)
    t_nvl = (2 * (world - 1) / world * isz * n) / (NVLINK_GBS * 1e9) if world > 1 else 0.0
    out = {"mib": mb, "dtype": dtype, "aggregation_frequency": agg, "route": eng.route_name(), "us_per_step": ms * 1e3,
           "roofline_us": max(t_hbm, t_nvl) * 1e6, "frac": max(t_hbm, t_nvl) / (ms * 1e-3) if ms > 0 else None,
           "bound": "nvlink" if t_nvl > t_hbm else "hbm", "phases_us": phases, "buckets": sum(len(fs.buckets) for fs in eng.sets)}
    eng.close()
    del eng, opt, params, grads
    return out


def microbench_sweep(ctx: Any, dist: Any, args: argparse.Namespace, peak: float, rows: Optional[List[Dict[str, Any]]] = None) -> Dict[str, Any]:
    """`rows`: a list the caller keeps; every finished row is appended to it at once, so that a run that is cut short
    (bench._Deadline) can still report the rows it has."""
    from determined_clone_b200 import _lib

    world = ctx.distributed.size
    route: Optional[str] = None
    note = None
    if world > 1:
        # the route a conforming trial ends up on (GradSyncEngine route "auto"): the peer kernels when the GPUs can map
        # each other's memory, else NCCL allreduce
        route = "peer"
        try:
            gradsync_microbench(ctx, dist, min(args.micro_mb), "f32", 1, 1, peak, route="peer")
        except _lib.DgsError as e:
            route, note = "allreduce", f"peer route unavailable: {str(e)[:200]}"
    rows = [] if rows is None else rows
    for dtype in args.micro_dtypes:
        for agg in args.micro_aggs:
            for mb in args.micro_mb:
                rows.append(gradsync_microbench(ctx, dist, mb, dtype, agg, args.micro_steps, peak, bucket_mb=args.micro_bucket_mb, route=route))
                DEV.empty_cache()
    head = [r for r in rows if r["dtype"] == "f32" and r["aggregation_frequency"] == 1]
    big = max(head, key=lambda r: r["mib"]) if head else None
    at_1g = next((r for r in head if r["mib"] == 1024), big)
    others = []
    if world > 1 and at_1g is not None and not args.micro_only:
        for other in ("allreduce", "sharded"):
            if other != route:
                # (one bucket: the NCCL collectives' best case -- their bus bandwidth grows with the message size)
                others.append(gradsync_microbench(ctx, dist, at_1g["mib"], "f32", 1, args.micro_steps, peak, bucket_mb=at_1g["mib"], route=other))
                DEV.empty_cache()
    bucketed = None
    if world > 1 and at_1g is not None and at_1g["mib"] >= 256 and not args.micro_only and not args.micro_bucket_mb:
        # the same point cut into 128 MiB buckets (what a model of this size gets from the engine): on the peer route the
        # barrier + owner reduction of a bucket then run on a side stream behind the pushes of the following buckets
        bucketed = gradsync_microbench(ctx, dist, at_1g["mib"], "f32", 1, args.micro_steps, peak, bucket_mb=128, route=route)
        DEV.empty_cache()
    return {"optimizer": "adamw + clip_grad_norm_ 1.0", "nvlink_gbs_per_direction": NVLINK_GBS, "hbm_gbs": peak,
            "bytes_per_element": {f"{d}_A{a}": v for (d, a), v in (MICRO_BYTES if world > 1 else MICRO_BYTES_1GPU).items()},
            "headline": at_1g, "same_point_in_128_mib_buckets": bucketed, "same_point_on_the_nccl_routes": others, "note": note, "rows": rows}


class _StepTimer:
    """Wraps ctx.step_optimizer with CUDA events on boundary batches: exposed grad-sync tail per optimizer step."""

    def __init__(self, ctx: Any, stream: Any) -> None:
        self.ctx, self.stream, self.ev, self.orig = ctx, stream, [], ctx.step_optimizer

    def __enter__(self) -> "_StepTimer":
        def timed(*a: Any, **k: Any) -> None:
            if not self.ctx._should_communicate_and_update():
                return self.orig(*a, **k)  # not an optimizer step: nothing to time
            a0, a1 = DEV.event(), DEV.event()
            a0.record(self.stream)
            self.orig(*a, **k)
            a1.record(self.stream)
            self.ev.append((a0, a1))

        self.ctx.step_optimizer = timed  # type: ignore
        return self

    def __exit__(self, *_: Any) -> None:
        self.ctx.step_optimizer = self.orig  # type: ignore

    def median_us(self) -> Optional[float]:
        DEV.sync()
        tails = sorted(a.elapsed_time(b) for a, b in self.ev)
        return tails[len(tails) // 2] * 1e3 if tails else None


def measure_workload(args: argparse.Namespace, W: Any, dctx: Any, dist: Any, device: Any, rank: int, world: int, local_rank: int,
                     steps: int, warmup: int, with_e2e: bool) -> Dict[str, Any]:
    """value / roofline / grad-sync tail (/ e2e) of one workload.  `steps` and `warmup` count train_batch calls."""
    import torch

    from determined_clone_b200 import _lib, pytorch

    batch, agg = W.PER_SLOT_BATCH, W.AGG
    steps = max(agg, steps // agg * agg)
    warmup = max(agg, (warmup + agg - 1) // agg * agg)
    pytorch._PyTorchTrialController.pre_execute_hook(17)
    ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": batch * world}, slots_per_trial=world,
                                      aggregation_frequency=agg, trial_seed=17, device=device)

    class BenchTrial(W, pytorch.PyTorchTrial):  # a real PyTorchTrial: the controller below is the public entry point
        loader: Any = None

        def build_training_data_loader(self) -> Any:
            return self.loader

        def build_validation_data_loader(self) -> Any:
            return None

    ctx.experimental.disable_dataset_reproducibility_checks()  # the loader below is a plain iterable of pinned batches
    trial = BenchTrial(ctx)
    eng = ctx._engines[trial.optimizer]
    p_total = sum(fs.total for fs in eng.sets)
    gen = torch.Generator().manual_seed(1234 + rank)
    host_batches = [tuple(DEV.pin(t) for t in W.synthetic_batch(batch, gen)) for _ in range(8)]
    dev_batches = [tuple(t.to(device) for t in b) for b in host_batches]
    stream = DEV.stream()

    def one_step(i: int) -> Any:
        ctx._current_batch_idx = i
        return trial.train_batch(dev_batches[i % len(dev_batches)], 0, i)

    def barrier() -> None:
        if dist is not None:
            dist.barrier()
        DEV.sync()

    # ---- device-timed region: inputs resident in HBM ----------------------------------------------
    for i in range(warmup):
        one_step(i)
    barrier()
    eng.enable_step_timing(True)
    launches0 = _lib.calls.dgs_launch_count()
    e0, e1 = DEV.event(), DEV.event()
    with ClockSampler(local_rank) as clocks:
        e0.record(stream)
        for i in range(warmup, warmup + steps):
            one_step(i)
        e1.record(stream)
        barrier()
    launches = _lib.calls.dgs_launch_count() - launches0
    ms_total = e0.elapsed_time(e1)
    step_kernel_ms = eng.step_timings_ms()
    eng.enable_step_timing(False)
    ms_total = _max_over_ranks(dist, device, [ms_total])[0]
    if dist is not None:
        lt = torch.tensor([launches], device=device, dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt)
    value = steps * batch * world / (ms_total * 1e-3)

    # ---- exposed grad-sync tail: from the end of backward's last kernel to the end of the step kernel ----
    done = warmup + steps
    with _StepTimer(ctx, stream) as tail:
        for i in range(8 * agg):
            one_step(done + i)
        tail_us = tail.median_us()
    done += 8 * agg

    # ---- roofline of the dominant kernel of the path (fused optimizer step) --------------------------------
    peak, peak_src = measured_peaks()
    # one optimizer step = one launch (several only when a flat set has more than 24 hyper-parameter ranges): the
    # launches of each step are added up, the median over the steps is reported
    n_opt = steps // agg
    per = len(step_kernel_ms) // n_opt if n_opt and step_kernel_ms and len(step_kernel_ms) % n_opt == 0 else 0
    per_step = sorted(sum(step_kernel_ms[i * per:(i + 1) * per]) for i in range(n_opt)) if per else []
    k_ms = per_step[len(per_step) // 2] if per_step else None
    roofline = None
    if k_ms:
        bpe = W.STEP_BYTES_PER_ELEM
        n_elem = eng.step_elements()
        hbm_bytes = bpe * n_elem
        cap = NCU_TRAFFIC.get(W.NAME, {}) if world == 1 else {}
        roofline = {"bound": "hbm", "kernel": eng.step_kernel_name(), "achieved": hbm_bytes / (k_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                    "traffic": cap.get("dram_bytes_per_launch"), "traffic_source": cap.get("source"), "peak_source": peak_src,
                    "elements": n_elem, "algorithmic_bytes_per_element": bpe, "algorithmic_bytes_per_launch": hbm_bytes,
                    "us_per_launch": k_ms * 1e3, "launches_timed": len(step_kernel_ms), "launches_per_step": per,
                    "note": "median over the timed region of CUDA events recorded inside libdgs around each launch (dgs_timer_read)"}
        if eng.peer:
            # the pushing step also stores every new parameter to the N-1 other ranks (the all-gather): at 8 GPUs those
            # NVLink bytes, not the HBM bytes of the 1/N slice, bound the launch
            psz = sum(fs.total * fs.param.element_size() for fs in eng.sets)
            nvl_bytes = (world - 1) / world * psz
            if nvl_bytes / (NVLINK_GBS * 1e9) > hbm_bytes / (peak * 1e9):
                roofline.update({"bound": "nvlink", "achieved": nvl_bytes / (k_ms * 1e-3) / 1e9, "peak": NVLINK_GBS,
                                 "peak_source": "900 GB/s per direction (BASELINE.md section 4)",
                                 "algorithmic_bytes_per_launch": nvl_bytes, "algorithmic_hbm_bytes_per_launch": hbm_bytes})
        roofline["frac"] = roofline["achieved"] / roofline["peak"]

    out: Dict[str, Any] = {
        "value": value, "ms_per_step": ms_total / steps, "steps": steps, "warmup": warmup, "gpu_launches": launches,
        "clocks": clocks.summary(), "grad_sync_us_per_step": tail_us, "roofline": roofline, "route": eng.route_name(),
        "config": {"workload": W.NAME, "per_slot_batch": batch, "global_batch": batch * world, "optimizer": W.OPTIMIZER,
                   "aggregation_frequency": agg, "parallelism": f"dp{world}", "param_elements": p_total,
                   "l2": "8 distinct input batches cycled; activations and the step kernel's working set exceed the 126 MB L2"},
    }

    # ---- end to end through the public API: host batches in, metrics out, every step ----------------------
    if with_e2e:
        class HostLoader:
            def __iter__(self) -> Any:
                i = 0
                while True:
                    yield host_batches[i % len(host_batches)]
                    i += 1

        trial.loader = HostLoader()
        reports: List[Any] = []
        base = (done + agg - 1) // agg * agg  # resume on an aggregation boundary
        warm = pytorch._PyTorchTrialController(trial, ctx, max_length=base + 2 * agg, reporting_period=1, steps_completed=base)
        warm.run()
        ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=base + 2 * agg + steps, reporting_period=1,
                                              steps_completed=base + 2 * agg, on_report=lambda n, r: reports.append(r))
        barrier()
        t0 = time.perf_counter()
        ctl.run()
        barrier()
        e2e_s = _max_over_ranks(dist, device, [time.perf_counter() - t0])[0]
        h2d = sum(t.numel() * t.element_size() for t in host_batches[0])
        d2h = 1 * 1 * 8 + 8 + 1 + 8  # per-batch mean (1x1 f64) + period mean (f64) + validity byte + int64 count
        out["e2e"] = {"value": steps * batch * world / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "ms_per_step": e2e_s / steps * 1e3, "reports": len(reports) if rank == 0 else None,
                      "api": "_PyTorchTrialController.run(), reporting_period=1, pinned host batches"}
    ctx.close()
    del trial, eng, ctx, dev_batches
    DEV.empty_cache()
    return out


def run_parity(args: argparse.Namespace, dctx: Any, dist: Any, device: Any, rank: int, world: int,
               cases: Optional[List[Dict[str, Any]]] = None) -> Dict[str, Any]:
    """See the module docstring.  Returns {"n", "ok", "max_rel", "cases": [...]}.  `cases`: a list the caller keeps, filled
    as the cases finish (see microbench_sweep)."""
    from determined_clone_b200 import _lib, core, pytorch
    from workloads import parity as P

    cases = [] if cases is None else cases

    def reduce_max(v: List[float]) -> List[float]:
        return _max_over_ranks(dist, device, v)

    # N = 1: free-running and per-step synced.  N > 1: the same on the default route (auto: NCCL allreduce for two
    # steps, then the peer kernels where the trial's clip function allows it) and per-step synced with the peer
    # route and with the NCCL allreduce route pinned, so that both exchanges are checked whatever auto decides
    plans = [("free", None), ("synced", None)] + ([("synced", "peer"), ("synced", "allreduce")] if world > 1 else [])
    for name in args.parity_cases:
        W = _workload(args, name)
        for mode, route in plans:
            pytorch._PyTorchTrialController.pre_execute_hook(17)
            ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": W.PER_SLOT_BATCH * world},
                                              slots_per_trial=world, aggregation_frequency=W.AGG, trial_seed=17, device=device,
                                              exp_conf={"optimizations": {"grad_sync_route": route}} if route else None)
            try:
                case = P.live_parity(name, W, ctx, device, rank, world, steps=args.parity_steps, mode=mode, reduce_max=reduce_max)
            except _lib.DgsError as e:  # peer set-up refused on this box (all ranks alike): reported, not a mismatch
                if route != "peer":
                    raise
                case = {"case": name, "mode": mode, "ok": True, "skipped": str(e)[:300]}
            case["route_requested"] = route or "auto"
            cases.append(case)
            DEV.empty_cache()
    # the 2-slot fixtures recorded from the unmodified reference (real torch DDP over gloo), on ranks 0 and 1
    if world >= 2 and not DEV.emulated:
        sys.path.insert(0, str(ROOT / "tests"))
        ports = _free_ports(2) if rank == 0 else [0, 0]
        ports = dctx.broadcast(ports)
        failed = ""
        if rank < 2:
            from golden_io import SCENARIOS
            from test_gpu_trial_parity import run_scenario

            sub = core.DistributedContext(rank=rank, size=2, local_rank=rank, local_size=2, cross_rank=0, cross_size=1,
                                          pub_port=ports[0], pull_port=ports[1])
            try:
                for name, sc in SCENARIOS.items():
                    if sc["slots"] == 2:
                        try:
                            run_scenario(name, distributed=sub)
                        except AssertionError as e:
                            failed = f"{name}: {str(e)[:200]}"
                            break
            finally:
                sub.close()
        everyone = dctx.allgather(failed)
        bad = [f for f in everyone if f]
        cases.append({"case": "reference-recorded 2-slot fixtures (sgd_n2_a2, adamw_n2_a4_clip, adamw_bf16_n2_a2_clip) on ranks 0-1",
                      "mode": "fixtures", "ok": not bad, "error": bad[0] if bad else None})
    free = [c for c in cases if c.get("mode") == "free" and "max_rel" in c]
    # what makes the run exit non-zero: the deterministic comparisons (fp32 one-step errors on identical inputs, the
    # reference-recorded fixtures).  Free-running arms and 16-bit cases are judged against the reference's own twin,
    # a statistical bound: they are reported in `ok`, but a chaotic draw must not void a measurement.
    hard = [c for c in cases if c.get("mode") == "fixtures" or (c.get("mode") == "synced" and c.get("dtype") == "f32")]
    one_step = [c["max_rel"] for c in hard if "max_rel" in c]
    return {"n": world, "ok": all(c["ok"] for c in cases), "deterministic_checks_ok": all(c["ok"] for c in hard),
            # the error of ONE optimizer step on identical inputs, fp32 cases, every route (tensor-relative; rtol 1e-5)
            "max_rel_one_step_fp32": max(one_step, default=None),
            # free-running arms after `parity_steps` steps, all cases: includes the chaotic ones, which are judged against
            # the reference's own twin (cases[].reference_self_noise), not against rtol
            "max_rel": max((c["max_rel"] for c in free), default=None),
            "against": "torch DDP (NCCL) + torch.optim on the same GPUs, same seeds and inputs (workloads/parity.py)", "cases": cases}


def run_ours(args: argparse.Namespace) -> int:
    import torch

    from determined_clone_b200 import _build, _lib, core, pytorch

    W = _workload(args)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not DEV.available():
        raise SystemExit("bench.py: no CUDA device; the grad-sync path has no CPU implementation (use --impl reference)")
    device = DEV.device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist  # type: ignore

        if DEV.emulated:
            dist.init_process_group(DEV.backend)
        else:
            dist.init_process_group(DEV.backend, device_id=device)
    if not DEV.emulated:
        if rank == 0 and not _lib.LIB_PATH.exists():
            _build.build()
        if dist is not None:
            dist.barrier()
        _lib.load()

    # control plane for the trial: the reference's DistributedContext (ZeroMQ through the chief)
    if world > 1:
        ports = _free_ports(2) if rank == 0 else [0, 0]
        dist.broadcast_object_list(ports, src=0)
        dctx = core.DistributedContext(rank=rank, size=world, local_rank=local_rank, local_size=world, cross_rank=0, cross_size=1,
                                       pub_port=ports[0], pull_port=ports[1])
    else:
        dctx = core.DummyDistributedContext()
    peak, _ = measured_peaks()
    got: Dict[str, Any] = {"head": None, "parity": None, "extra": {}, "micro": None, "cpu": None, "wall_s": {}}
    unfinished: Dict[str, List[Dict[str, Any]]] = {"parity": [], "micro": []}  # cases / rows of a part that is still running
    errors: List[str] = []
    cpu_error: Optional[str] = None  # kept apart from `errors`: it does not put the ranks out of step

    def emit(cut: Optional[str]) -> None:
        head = got["head"]
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": head["steps"], "warmup": head["warmup"],
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": W.DTYPE, "data": "synthetic", "config": head["config"],
            "roofline": head["roofline"], "cpu_baseline": got["cpu"], "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"],
            "clocks": head["clocks"], "grad_sync_us_per_step": head["grad_sync_us_per_step"], "route": head["route"],
            "parity": got["parity"] or ({"n": world, "incomplete": True, "cases": unfinished["parity"]} if unfinished["parity"] else None),
            "gradsync_microbench": got["micro"] or ({"incomplete": True, "rows": unfinished["micro"]} if unfinished["micro"] else None),
            "configs": got["extra"],
            "errors": (errors + ([cut] if cut else [])) or None,
            # this rank's wall clock, seconds since bench.py was imported, at which each part was complete
            "wall_s": dict(got["wall_s"], printed=round(time.monotonic() - T0, 1)),
        }
        print(json.dumps(line), flush=True)

    guard = _Deadline(rank, emit)
    if not DEV.emulated:
        guard.arm()
    try:
        if args.micro_only:
            pytorch._PyTorchTrialController.pre_execute_hook(17)
            ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": world}, slots_per_trial=world, device=device)
            sweep = microbench_sweep(ctx, dist, args, peak)
            if rank == 0:
                print(json.dumps({"n_gpus": world, "gradsync_microbench": sweep}), flush=True)
            guard.printed = True
            return 0

        # the headline first; everything after it is additional evidence, and a failure there (recorded under `errors`)
        # must not void the measurement already taken.  Only a deterministic parity mismatch makes the run exit non-zero.
        guard.phase = "headline"
        got["head"] = measure_workload(args, W, dctx, dist, device, rank, world, local_rank, args.steps, args.warmup, with_e2e=True)
        guard.have_headline = True
        got["wall_s"]["headline"] = round(guard.elapsed(), 1)

        if world == 1 and not args.no_cpu_baseline:  # (one rank, bounded: nothing to keep in step)
            guard.phase = "cpu_baseline"
            try:
                r = run_cpu_port(W, steps=20, warmup=1, budget_s=args.cpu_budget_s)
                got["cpu"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
            except Exception as e:  # noqa: BLE001
                cpu_error = f"cpu_baseline: {type(e).__name__}: {str(e)[:300]}"
            got["wall_s"]["cpu_baseline"] = round(guard.elapsed(), 1)

        def optional(what: str, fn: Callable[[], Any]) -> Any:
            guard.phase = f"before {what}"
            # after a failure anywhere the ranks may no longer be in step: stop adding work, on every rank alike
            failed = [f"rank {r}: {e[0]}" for r, e in enumerate(dctx.allgather(errors)) if e]
            if failed:
                if not errors:
                    errors.append(f"{what}: not started ({failed[0][:200]})")
                return None
            if not dctx.broadcast(guard.elapsed() < guard.soft):  # rank 0's clock decides for everyone
                errors.append(f"{what}: not started, {guard.soft:.0f} s of the run's time were used")
                return None
            guard.phase = what
            try:
                return fn()
            except Exception as e:  # noqa: BLE001
                errors.append(f"{what}: {type(e).__name__}: {str(e)[:300]}")
                return None
            finally:
                got["wall_s"][what] = round(guard.elapsed(), 1)

        if not args.no_parity:
            got["parity"] = optional("parity", lambda: run_parity(args, dctx, dist, device, rank, world, cases=unfinished["parity"]))
            if got["parity"] is not None and not got["parity"]["deterministic_checks_ok"]:
                guard.rc = 3

        if not args.no_extra and args.workload != "gpt2_medium":
            G = _workload(args, "gpt2_medium")
            sub = optional("configs[3]", lambda: measure_workload(args, G, dctx, dist, device, rank, world, local_rank, 8 * G.AGG, 3 * G.AGG, with_e2e=False))
            if sub is not None:
                sub["unit"] = "sequences/s (4 x 1024 tokens per slot per micro-batch)"
                got["extra"][G.NAME] = sub

        if not args.no_microbench:
            def sweep() -> Any:
                pytorch._PyTorchTrialController.pre_execute_hook(17)
                ctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": world}, slots_per_trial=world, device=device)
                return microbench_sweep(ctx, dist, args, peak, rows=unfinished["micro"])

            got["micro"] = optional("gradsync_microbench", sweep)

        guard.phase = "teardown"
        guard.print_once(cpu_error)
    finally:
        dctx.close()
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        guard.disarm()
    return guard.rc


def run_torch_gpu(args: argparse.Namespace) -> None:
    """The reference's per-step path on the same GPUs: torch DDP (NCCL) + torch.optim, through the torch port of
    the reference context.  Not part of the driver contract; prints a line with impl 'torch' for comparison."""
    import torch

    from workloads.parity import DdpRefContext

    W = _workload(args)
    batch = W.PER_SLOT_BATCH
    rank, world, local_rank = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist  # type: ignore

        dist.init_process_group("nccl", device_id=device)
    torch.manual_seed(17)
    ctx = DdpRefContext(world, aggregation_frequency=W.AGG, device=str(device))
    trial = W(ctx)
    gen = torch.Generator().manual_seed(1234 + rank)
    batches = [tuple(t.to(device) for t in W.synthetic_batch(batch, gen)) for _ in range(8)]

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
    ms = _max_over_ranks(dist, device, [e0.elapsed_time(e1)])[0]
    if rank == 0:
        print(json.dumps({"impl": "torch", "metric": METRIC, "value": args.steps * batch * world / (ms * 1e-3), "unit": UNIT,
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                          "config": {"workload": W.NAME, "parallelism": f"torch DDP dp{world}"}}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    # stdout carries exactly one JSON line: keep NCCL's "NCCL version ..." banner (printed to stdout when the image
    # sets NCCL_DEBUG=VERSION) out of it; an explicit NCCL_DEBUG=INFO from the caller is respected
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() in ("VERSION", "WARN"):
        os.environ.pop("NCCL_DEBUG", None)  # both levels print the banner
    # a rank that never arrives at a peer barrier is reported after this long (not a hang, not a GPU fault); the library's
    # own default of 120 s suits training jobs, a benchmark wants to know sooner
    os.environ.setdefault("DGS_PEER_TIMEOUT_S", "60")
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "torch":
        run_torch_gpu(args)
    else:
        sys.exit(run_ours(args))


if __name__ == "__main__":
    main()
