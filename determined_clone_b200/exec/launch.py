"""Run an experiment's `entrypoint:` on this box (reference harness/determined/exec/launch.py:29-63).

    DET_EXPERIMENT_CONFIG=exp.yaml python -m determined_clone_b200.exec.launch

The reference turns a legacy `module:TrialClass` entrypoint into its Horovod launcher; here the same entrypoint goes
to the torch.distributed launcher (one worker per slot over NCCL) when the experiment asks for several slots, and
straight to the harness for one.  A string entrypoint is a shell command line, a list is an argv -- both run as
they are, with the signals a scheduler may send forwarded to them."""
import re
import signal
import subprocess
import sys
from typing import Any, Dict, List, Union

from determined_clone_b200 import _info

_LEGACY_TRIAL = re.compile(r"^[a-zA-Z0-9_.]+:[a-zA-Z0-9_]+$")  # module.path:ClassName and nothing else


def entrypoint_argv(entrypoint: Union[str, List[str]], slots: int) -> List[str]:
    if isinstance(entrypoint, str) and _LEGACY_TRIAL.match(entrypoint):
        if slots > 1:
            return ["python3", "-m", "determined.launch.torch_distributed", "--trial", entrypoint]
        return ["python3", "-m", "determined.exec.harness", entrypoint]
    if isinstance(entrypoint, str):
        return ["sh", "-c", entrypoint]
    return list(entrypoint)


def launch(experiment_config: Dict[str, Any]) -> int:
    slots = int((experiment_config.get("resources") or {}).get("slots_per_trial", 1))
    argv = entrypoint_argv(experiment_config["entrypoint"], slots)
    argv = [sys.executable if a == "python3" else a for a in argv]
    child = subprocess.Popen(argv)
    forwarded = [getattr(signal, n) for n in ("SIGINT", "SIGTERM", "SIGHUP", "SIGUSR1", "SIGUSR2") if hasattr(signal, n)]
    previous = {s: signal.signal(s, lambda signum, frame: child.send_signal(signum)) for s in forwarded}
    try:
        return child.wait()
    finally:
        for s, handler in previous.items():
            signal.signal(s, handler)


if __name__ == "__main__":
    info = _info.get_cluster_info()
    if info is None:
        print("DET_EXPERIMENT_CONFIG must point at the experiment YAML", file=sys.stderr)
        sys.exit(2)
    sys.exit(launch(info.trial._config))
