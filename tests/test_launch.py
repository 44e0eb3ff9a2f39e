"""Single-box launch layer and Trainer API: argv construction, experiment-config plumbing, and
the reference's end-to-end closed-form check run through the same YAML + entrypoint a cluster would use."""
import json
import os
import pathlib
import shlex
import subprocess
import sys

import numpy as np
import pytest
import yaml

from determined_clone_b200 import _info, pytorch
from determined_clone_b200.launch import torch_distributed as launch
from oracle import gradsync_oracle as O

ROOT = pathlib.Path(__file__).resolve().parent.parent
FIX = ROOT / "tests" / "fixtures" / "pytorch_identity"


def test_launch_command_line():
    # reference harness/tests/launch/test_torch_distributed.py asserts the same torchrun shape
    cmd = launch.create_launch_cmd(1, 8, 0, "127.0.0.1", ["--rdzv_conf", "x=1"])
    assert cmd[1:4] == ["-m", "torch.distributed.run", "--nnodes"]
    assert cmd[cmd.index("--nproc_per_node") + 1] == "8" and cmd[cmd.index("--max_restarts") + 1] == "0"
    assert cmd[cmd.index("--master_addr") + 1] == "127.0.0.1" and cmd[-2:] == ["--rdzv_conf", "x=1"]
    assert launch.parse_args(["--a", "1", "--", "python3", "train.py"]) == (["--a", "1"], ["python3", "train.py"])
    assert launch.parse_args(["python3", "train.py"]) == ([], ["python3", "train.py"])
    # the cases of the reference's test_parse_args (harness/tests/launch/test_torch_distributed.py:9-29)
    harness = ["python3", "-m", "determined.exec.harness", "my_module:MyTrial"]
    assert launch.parse_args("--trial my_module:MyTrial".split()) == ([], harness)
    assert launch.parse_args("script arg".split()) == ([], ["script", "arg"])
    assert launch.parse_args("-- script -- arg".split()) == ([], ["script", "--", "arg"])
    assert launch.parse_args("override1 override2 -- script arg".split()) == (["override1", "override2"], ["script", "arg"])


@pytest.mark.parametrize("line,complaint", [("--trial my_module:MyTrial script", "extra arguments"), ("", "empty script"),
                                            ("--asdf 1 script", "unrecognized arguments")])
def test_launch_command_line_errors(capsys, line, complaint):
    with pytest.raises(SystemExit) as exit_info:
        launch.parse_args(line.split())
    assert exit_info.value.code != 0 and complaint in capsys.readouterr().err


def test_cluster_info_from_experiment_yaml(monkeypatch):
    monkeypatch.delenv("DET_EXPERIMENT_CONFIG", raising=False)
    assert _info.get_cluster_info() is None
    monkeypatch.setenv("DET_EXPERIMENT_CONFIG", str(FIX / "distributed.yaml"))
    info = _info.get_cluster_info()
    assert info.trial.hparams == {"global_batch_size": 4, "learning_rate": 0.001}
    assert len(info.slot_ids) == 2 and info.task_type == "TRIAL"
    assert info.trial._config["optimizations"]["aggregation_frequency"] == 2


def test_train_units():
    assert pytorch.TrainUnit._from_values(batches=3).value == 3
    assert isinstance(pytorch.TrainUnit._from_searcher_unit(5, "epochs"), pytorch.Epoch)
    r = pytorch.TrainUnit._from_searcher_unit(64, "records", 4)
    assert isinstance(r, pytorch.Record) and r.global_batch_size == 4
    with pytest.raises(ValueError):
        pytorch.TrainUnit._from_values(batches=1, epochs=1)
    ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4})
    t = pytorch.Trainer(None, ctx)  # type: ignore
    assert t._to_batches(pytorch.Batch(7), None) == 7 and t._to_batches(pytorch.Record(64), None) == 16
    assert t._to_batches(pytorch.Epoch(2), 16) == 32
    with pytest.raises(ValueError):
        t._to_batches(pytorch.Epoch(1), None)


def _run_experiment(config: str, tmp_path: pathlib.Path, FIX: pathlib.Path = FIX, extra_env=None) -> list:
    """Run the experiment's own `entrypoint:` the way a container would, on this box."""
    conf = yaml.safe_load((FIX / config).read_text())
    entry = conf["entrypoint"]
    argv = shlex.split(entry) if " " in entry else ["python3", "-m", "determined.exec.harness", entry]
    argv = [sys.executable if a == "python3" else a for a in argv]
    report = tmp_path / "reports.json"
    env = dict(os.environ, DET_EXPERIMENT_CONFIG=str(FIX / config), DET_REPORT_FILE=str(report),
               PYTHONPATH=os.pathsep.join([str(ROOT / "determined_clone_b200" / "compat"), str(ROOT), os.environ.get("PYTHONPATH", "")]),
               C10D_PORT="29433", **(extra_env or {}))
    subprocess.run(argv, cwd=FIX, env=env, check=True, timeout=600)
    return json.loads(report.read_text())


@pytest.mark.gpu
@pytest.mark.parametrize("config", ["const.yaml", "distributed.yaml"])
def test_identity_experiment_matches_closed_form(lib, config, tmp_path):
    """16 batches of y = w*x with aggregation_frequency 2 on 1 slot (global batch 4) and on 2 slots through
    `determined.launch.torch_distributed`: the weight after every batch equals the reference's closed form
    (test_pytorch_trial.py:1604-1642), and the chief reports after every batch (scheduling_unit 1)."""
    import torch

    slots = yaml.safe_load((FIX / config).read_text())["resources"]["slots_per_trial"]
    if torch.cuda.device_count() < slots:
        pytest.skip(f"needs {slots} GPUs")
    reports = _run_experiment(config, tmp_path)
    assert [r["steps_completed"] for r in reports] == list(range(1, 17))  # integer counters: exact
    got = [r["batch_metrics"][0]["weight"] for r in reports]
    want = O.identity_expected_weights()[:16]
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-9)
    assert got[0] == 0.0 and got[1] == got[2] != 0.0  # no update on non-boundary batches


def test_legacy_entrypoint_drives_validation_and_checkpoint_policy_from_the_config(tmp_path):
    """`entrypoint: model_def:Trial` with min_validation_period, checkpoint_policy: best, perform_initial_validation and
    scheduling_unit read from the experiment YAML as the reference's harness does (exec/harness.py:186-200).  The
    trial has no optimizer, so this runs on the CPU.  The validation metric |w| goes 2.0 (initial, at batch 0, where
    the reference too considers the checkpoint current), 1.0, 0.0, 1.0: 'best' checkpoints after the validations at
    batches 2 and 4 but not after the one at 6; the checkpoint at 6 is the end-of-training one."""
    import pickle

    host = FIX.parent / "host_only"
    ckpts = tmp_path / "ckpts"
    reports = _run_experiment("const.yaml", tmp_path, FIX=host, extra_env={"DET_CHECKPOINT_DIR": str(ckpts)})
    # reports at scheduling_unit (3, 6) and wherever a validation closes the period (2, 4)
    assert [r["steps_completed"] for r in reports] == [2, 3, 4, 6]
    assert [b["loss"] for r in reports for b in r["batch_metrics"]] == [1.5, 1.0, 0.5, 0.0, -0.5, -1.0]
    saved = sorted(ckpts.iterdir(), key=lambda d: d.stat().st_mtime)
    steps = []
    for d in saved:
        with (d / "trial_state.pkl").open("rb") as f:
            state = pickle.load(f)
        steps.append((state["batches_trained"], state["last_val"]))
    assert steps == [(2, 2), (4, 4), (6, 6)]  # (6, 6): not "best" (|w| = 1.0 again), written because training ended
    meta = json.loads((saved[-1] / "metadata.json").read_text())
    assert meta["steps_completed"] == 6 and meta["format"] == "pickle"
    load_data = json.loads((saved[-1] / "load_data.json").read_text())
    assert load_data["trial_cls_spec"] == "model_def:HostOnlyTrial" and load_data["hparams"]["start"] == 2.0
    assert load_data["experiment_config"]["searcher"]["metric"] == "distance"


def test_entrypoint_dispatch_and_launch(tmp_path):
    """exec/launch.py: a legacy `module:Class` entrypoint becomes the harness (1 slot) or the torch.distributed
    launcher (several); strings are shell command lines, lists are argvs (reference exec/launch.py:29-46).  Then the
    host-only experiment runs end to end through `python -m determined.exec.launch`."""
    from determined_clone_b200.exec import launch as exec_launch

    assert exec_launch.entrypoint_argv("model_def:MyTrial", 1) == ["python3", "-m", "determined.exec.harness", "model_def:MyTrial"]
    assert exec_launch.entrypoint_argv("pkg.model_def:MyTrial", 8) == ["python3", "-m", "determined.launch.torch_distributed", "--trial", "pkg.model_def:MyTrial"]
    assert exec_launch.entrypoint_argv("python3 train.py --lr 0.1", 8) == ["sh", "-c", "python3 train.py --lr 0.1"]
    assert exec_launch.entrypoint_argv(["python3", "train.py"], 2) == ["python3", "train.py"]

    host = FIX.parent / "host_only"
    report = tmp_path / "reports.json"
    env = dict(os.environ, DET_EXPERIMENT_CONFIG=str(host / "const.yaml"), DET_REPORT_FILE=str(report), DET_CHECKPOINT_DIR=str(tmp_path / "ckpts"),
               PYTHONPATH=os.pathsep.join([str(ROOT / "determined_clone_b200" / "compat"), str(ROOT), os.environ.get("PYTHONPATH", "")]))
    subprocess.run([sys.executable, "-m", "determined.exec.launch"], cwd=host, env=env, check=True, timeout=600)
    assert [r["steps_completed"] for r in json.loads(report.read_text())] == [2, 3, 4, 6]
    missing = subprocess.run([sys.executable, "-m", "determined.exec.launch"], cwd=host, capture_output=True, text=True,
                             env={k: v for k, v in env.items() if k != "DET_EXPERIMENT_CONFIG"})
    assert missing.returncode == 2 and "DET_EXPERIMENT_CONFIG" in missing.stderr


def test_horovod_launcher_command_line():
    """determined.launch.horovod, the reference's default launch layer (harness/determined/launch/horovod.py:205-281):
    same command line; overrides for horovodrun are accepted and dropped (there is no horovodrun here)."""
    from determined_clone_b200.launch import horovod

    assert horovod.parse_args(["--autohorovod", "--trial", "model_def:MyTrial"]) == (
        [], ["python3", "-m", "determined.exec.harness", "model_def:MyTrial"], True)
    assert horovod.parse_args(["--mpi-args=-v", "--", "python3", "train.py", "--lr", "1"]) == (
        ["--mpi-args=-v"], ["python3", "train.py", "--lr", "1"], False)
    for bad in (["--trial", "m:T", "extra"], []):
        with pytest.raises(SystemExit) as e:
            horovod.parse_args(bad)
        assert e.value.code == 1


def test_experiment_yaml_that_names_the_horovod_launcher_runs_unchanged(tmp_path):
    """entrypoint: python3 -m determined.launch.horovod --autohorovod --trial model_def:HostOnlyTrial (one slot): the
    script runs directly, as under the reference's --autohorovod, and produces the reports of the legacy entrypoint."""
    host = FIX.parent / "host_only"
    reports = _run_experiment("horovod.yaml", tmp_path, FIX=host, extra_env={"DET_CHECKPOINT_DIR": str(tmp_path / "ckpts")})
    assert [r["steps_completed"] for r in reports] == [2, 3, 4, 6]
    assert [b["loss"] for r in reports for b in r["batch_metrics"]] == [1.5, 1.0, 0.5, 0.0, -0.5, -1.0]
