"""`python3 -m determined.launch.torch_distributed ...` -> determined_clone_b200.launch.torch_distributed."""
import sys

from determined_clone_b200.launch.torch_distributed import main, parse_args

if __name__ == "__main__":
    _ov, _sc = parse_args(sys.argv[1:])
    sys.exit(main(_ov, _sc))
