"""`python3 -m determined.exec.launch` -> determined_clone_b200.exec.launch."""
import sys

from determined_clone_b200 import _info
from determined_clone_b200.exec.launch import entrypoint_argv, launch  # noqa: F401

if __name__ == "__main__":
    _i = _info.get_cluster_info()
    if _i is None:
        print("DET_EXPERIMENT_CONFIG must point at the experiment YAML", file=sys.stderr)
        sys.exit(2)
    sys.exit(launch(_i.trial._config))
