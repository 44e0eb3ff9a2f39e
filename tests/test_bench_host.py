"""The driver's command must not rot: bench.py's own control flow (parity pass, timed region, grad-sync tail, e2e
through _PyTorchTrialController, microbench sweep, cpu baseline, the JSON line) runs here on the CPU with the C ABI
replaced by tests/fake_dgs.py and bench.DEV replaced by a host stand-in -- at N=1 in process and at N=2 as two gloo
ranks.  Nothing here says anything about speed or about the kernels; it only keeps `python bench.py --gpus N` from
breaking unnoticed when the engine or the controller change (round 1 lost its whole measurement that way)."""
import json
import os
import pathlib
import socket
import sys
import time

import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
ARGV = ["--steps", "2", "--warmup", "1", "--workload", "cifar_resnet18", "--per-slot-batch", "8", "--parity-cases", "cifar_resnet18",
        "--parity-steps", "2", "--micro-mb", "1", "--micro-steps", "2", "--no-extra", "--cpu-budget-s", "1"]


class _HostEvent:
    def record(self, stream=None):
        self.t = time.perf_counter()

    def elapsed_time(self, other):
        return max((other.t - self.t) * 1e3, 1e-6)


class HostDev:
    emulated = True
    backend = "gloo"

    def available(self):
        return True

    def device(self, local_rank):
        return torch.device("cpu")

    def sync(self):
        pass

    def event(self):
        return _HostEvent()

    def stream(self):
        return None

    def pin(self, t):
        return t

    def empty_cache(self):
        pass


def _check_line(line):
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks", "parity", "gradsync_microbench"):
        assert key in line, key
    assert line["value"] > 0 and line["e2e"]["value"] > 0 and line["e2e"]["h2d_bytes_per_step"] > 0
    assert line["gpu_launches"] > 0
    assert line["config"]["workload"] == "cifar10_resnet18_pytorchtrial" and line["config"]["global_batch"] == 8 * line["n_gpus"]
    assert line["parity"]["ok"] and line["parity"]["n"] == line["n_gpus"]
    want = [("free", "auto"), ("synced", "auto")] + ([("synced", "peer"), ("synced", "allreduce")] if line["n_gpus"] > 1 else [])
    assert [(c["mode"], c["route_requested"]) for c in line["parity"]["cases"]] == want
    rows = line["gradsync_microbench"]["rows"]
    assert {(r["dtype"], r["aggregation_frequency"]) for r in rows} == {("f32", 1), ("f32", 4), ("bf16", 1), ("bf16", 4)}


def test_bench_ours_single_process(monkeypatch, capsys):
    sys.path.insert(0, str(ROOT / "tests"))
    import fake_dgs

    import bench

    fake_dgs.install(monkeypatch)
    monkeypatch.setattr(bench, "DEV", HostDev())
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    torch.set_num_threads(2)
    assert bench.run_ours(bench.parse(ARGV)) == 0
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    _check_line(line)
    assert line["n_gpus"] == 1 and line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["value"] > 0
    walls = line["wall_s"]  # when each part was complete: headline first, the line last
    assert list(walls) == ["headline", "cpu_baseline", "parity", "gradsync_microbench", "printed"] and sorted(walls.values()) == list(walls.values())


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank(rank, world, port, out_dir):
    for p in (str(ROOT), str(ROOT / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import fake_dgs

    import bench

    torch.set_num_threads(1)
    fake_dgs.install()
    bench.DEV = HostDev()
    if rank == 0:
        sys.stdout = open(os.path.join(out_dir, "line.json"), "w")
    rc = bench.run_ours(bench.parse(ARGV))
    sys.stdout.flush()
    assert rc == 0


def test_bench_ours_two_gloo_ranks(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_rank, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    line = json.loads((tmp_path / "line.json").read_text().strip().splitlines()[-1])
    _check_line(line)
    assert line["n_gpus"] == 2 and line["cpu_baseline"] is None and line["config"]["parallelism"] == "dp2"


def _ref_rank(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT))
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import bench

    if rank == 0:
        sys.stdout = open(os.path.join(out_dir, "ref.json"), "w")
    bench.run_reference(bench.parse(["--impl", "reference", "--gpus", str(world), "--steps", "2", "--warmup", "1",
                                     "--workload", "cifar_resnet18", "--per-slot-batch", "2"]))
    sys.stdout.flush()


def test_bench_reference_arm_runs_n_gloo_ranks(tmp_path):
    """--impl reference at N=2 is two gloo ranks with DDP, at the GPU arm's per-slot batch (global batch = 2 x per slot)."""
    import torch.multiprocessing as mp

    mp.spawn(_ref_rank, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    line = json.loads((tmp_path / "ref.json").read_text().strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["config"]["global_batch"] == 4
    assert line["cpu_baseline"]["ranks"] == 2 and line["value"] > 0 and line["e2e"]["h2d_bytes_per_step"] == 0


def test_time_limits_keep_the_headline(monkeypatch, capsys):
    """Past the soft limit no optional part is started (the line still carries headline, e2e and cpu baseline,
    and names what was left out); at the hard limit the timer prints what is complete exactly once and leaves."""
    sys.path.insert(0, str(ROOT / "tests"))
    import fake_dgs

    import bench

    fake_dgs.install(monkeypatch)
    monkeypatch.setattr(bench, "DEV", HostDev())
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("DGS_BENCH_SOFT_S", "0")
    torch.set_num_threads(2)
    assert bench.run_ours(bench.parse(ARGV)) == 0
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert line["value"] > 0 and line["e2e"]["value"] > 0 and line["cpu_baseline"]["value"] > 0
    assert line["parity"] is None and line["gradsync_microbench"] is None
    assert len(line["errors"]) == 1 and line["errors"][0].startswith("parity: not started")

    # a part that fails half-way: what it had finished is reported as incomplete, the failure is named, nothing else starts
    def broken_parity(args, dctx, dist, device, rank, world, cases=None):
        cases.append({"case": "first", "ok": True})
        raise RuntimeError("second case fell over")

    monkeypatch.setenv("DGS_BENCH_SOFT_S", "100000")
    monkeypatch.setattr(bench, "run_parity", broken_parity)
    assert bench.run_ours(bench.parse(ARGV + ["--no-cpu-baseline"])) == 0
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert line["value"] > 0 and line["parity"] == {"n": 1, "incomplete": True, "cases": [{"case": "first", "ok": True}]}
    assert line["gradsync_microbench"] is None and line["errors"][0] == "parity: RuntimeError: second case fell over"

    printed, left = [], []
    guard = bench._Deadline(0, printed.append)
    guard._exit = left.append
    guard.phase = "headline"
    guard.expire()  # nothing measured yet: nothing to print, non-zero exit
    assert printed == [] and left == [4]
    guard.have_headline, guard.phase, guard.rc = True, "gradsync_microbench", 3
    guard.expire()
    guard.print_once()  # the normal end of the run after the timer fired: the line is not printed twice
    assert len(printed) == 1 and "gradsync_microbench" in printed[0] and left == [4, 3]
    other = bench._Deadline(1, printed.append)  # ranks other than 0 print nothing
    other._exit = left.append
    other.have_headline = True
    monkeypatch.setattr(bench.time, "sleep", lambda s: None)
    other.expire()
    assert len(printed) == 1 and left == [4, 3, 0]
