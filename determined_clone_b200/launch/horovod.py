"""Single-box stand-in for `python3 -m determined.launch.horovod [HVD_OVERRIDES... --] [--autohorovod] (--trial X | SCRIPT...)`,
the reference's default launch layer (reference harness/determined/launch/horovod.py: parse_args :205-281, main :94-204).

The reference wraps the script in horovodrun (sshd on the worker machines, a pid server, one process per slot) and
the harness then all-reduces gradients through Horovod.  On this path the per-step exchange is libdgs' own kernels
and communicator, so there is no Horovod to start: an experiment YAML whose entrypoint names this launcher gets
  * one slot (or --autohorovod with one slot): the script, run directly (the reference does the same, :102-105);
  * several slots: one worker per slot under torch.distributed.run -- the torch_distributed launcher of this
    package -- with the same DET_CHIEF_IP contract.
HVD_OVERRIDES are horovodrun options; they have nothing to act on here and are reported and dropped."""
import argparse
import logging
import os
import subprocess
import sys
from typing import List, Tuple

from determined_clone_b200 import _info
from determined_clone_b200.launch import torch_distributed

logger = logging.getLogger("determined.launch.horovod")


def parse_args(args: List[str]) -> Tuple[List[str], List[str], bool]:
    """-> (hvd_args, script, autohorovod), with the reference's cases: overrides end at the first `--`; `--trial X`
    stands for the legacy harness entrypoint and excludes SCRIPT; an empty script is a usage error (exit 1)."""
    hvd_args: List[str] = []
    if "--" in args:
        cut = args.index("--")
        hvd_args, args = args[:cut], args[cut + 1:]
    parser = argparse.ArgumentParser(
        usage="%(prog)s [[HVD_OVERRIDES...] --] (--trial TRIAL)|(SCRIPT...)",
        description="Launch a script with one worker per slot of this box (the reference launches it under horovodrun)",
    )
    parser.add_argument("--trial", help="a Trial class (module:Class) to train through the legacy harness entrypoint; excludes SCRIPT")
    parser.add_argument("script", metavar="SCRIPT...", nargs=argparse.REMAINDER, help="script to launch for training")
    parser.add_argument("--autohorovod", action="store_true", help=argparse.SUPPRESS)
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
    return hvd_args, script, bool(parsed.autohorovod)


def main(hvd_args: List[str], script: List[str], autohorovod: bool) -> int:
    if hvd_args:
        logger.warning(f"horovodrun overrides {hvd_args} ignored: this launch layer starts no horovodrun")
    info = _info.get_cluster_info()
    slots = len(info.slot_ids) if info is not None else int(os.environ.get("DET_SLOTS", "0") or 0)
    if slots <= 1:
        # one slot: no launcher at all.  (Without --autohorovod the reference would still wrap one process in
        # horovodrun; the trial code cannot tell the difference here, there being no Horovod either way.)
        os.environ["DET_CHIEF_IP"] = "127.0.0.1"
        argv = [sys.executable if a == "python3" else a for a in script]
        return subprocess.Popen(argv).wait()
    return torch_distributed.main([], script)


if __name__ == "__main__":
    _hvd, _script, _auto = parse_args(sys.argv[1:])
    sys.exit(main(_hvd, _script, _auto))
