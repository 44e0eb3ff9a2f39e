"""Single-box stand-in for `python3 -m determined.launch.torch_distributed <script...>`
(reference harness/determined/launch/torch_distributed.py:15-106): turns one container into one worker
process per slot with torch.distributed.run and sets USE_TORCH_DISTRIBUTED, so that an experiment YAML whose
entrypoint uses that launcher runs unchanged on one 8 x B200 box.  There is no master here: the experiment
config reaches the workers through DET_EXPERIMENT_CONFIG (see _info.py); the pid-server / log-redirect
wrappers of the reference are cluster plumbing and are left out."""
import argparse
import os
import subprocess
import sys
from typing import List, Tuple

from determined_clone_b200 import _info

C10D_PORT = int(str(os.getenv("C10D_PORT", "29400")))


def create_launch_cmd(num_nodes: int, proc_per_node: int, node_rank: int, master_addr: str, override_args: List[str]) -> List[str]:
    cmd = [
        sys.executable, "-m", "torch.distributed.run",
        "--nnodes", str(num_nodes),
        "--nproc_per_node", str(proc_per_node),
        "--node_rank", str(node_rank),
        "--max_restarts", "0",
        "--master_addr", master_addr,
        "--master_port", str(C10D_PORT),
        "--no-python",
    ]
    cmd.extend(override_args)
    return cmd


def parse_args(args: List[str]) -> Tuple[List[str], List[str]]:
    """[TORCH_OVERRIDES... --] (--trial module:TrialClass | SCRIPT...)  ->  (overrides, script), with the cases of
    the reference (launch/torch_distributed.py:108-160, its test_parse_args): `--trial X` stands for the legacy
    harness entrypoint and takes no further arguments; an empty script and unknown options are usage errors
    (exit status 1 / 2)."""
    overrides: List[str] = []
    if "--" in args:
        cut = args.index("--")
        overrides, args = args[:cut], args[cut + 1:]
    parser = argparse.ArgumentParser(
        usage="%(prog)s [[TORCH_OVERRIDES...] --] (--trial TRIAL)|(SCRIPT...)",
        description="Launch a script under pytorch distributed, one worker per slot of this box",
    )
    parser.add_argument("--trial", help="a Trial class (module:Class) to train through the legacy harness entrypoint; excludes SCRIPT")
    parser.add_argument("script", metavar="SCRIPT...", nargs=argparse.REMAINDER, help="script to launch for training")
    parsed = parser.parse_args(args)
    script = list(parsed.script or [])
    if parsed.trial is not None:
        if script:
            parser.print_usage()
            print("error: extra arguments to --trial:", script, file=sys.stderr)
            sys.exit(1)
        script = ["python3", "-m", "determined.exec.harness", parsed.trial]
    elif not script:
        parser.print_usage()
        print("error: empty script is not allowed", file=sys.stderr)
        sys.exit(1)
    return overrides, script


def main(override_args: List[str], script: List[str]) -> int:
    info = _info.get_cluster_info()
    slots = len(info.slot_ids) if info is not None else int(os.environ.get("DET_SLOTS", "0"))
    if slots < 1:
        import torch

        slots = max(torch.cuda.device_count(), 1)
    os.environ["USE_TORCH_DISTRIBUTED"] = "True"
    os.environ["DET_CHIEF_IP"] = "127.0.0.1"
    cmd = create_launch_cmd(1, slots, 0, "127.0.0.1", override_args) + script
    return subprocess.Popen(cmd).wait()


if __name__ == "__main__":
    ov, sc = parse_args(sys.argv[1:])
    sys.exit(main(ov, sc))
