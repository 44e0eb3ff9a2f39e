"""`python3 -m determined.launch.horovod ...` -> determined_clone_b200.launch.horovod."""
import sys

from determined_clone_b200.launch.horovod import main, parse_args

if __name__ == "__main__":
    _hvd, _script, _auto = parse_args(sys.argv[1:])
    sys.exit(main(_hvd, _script, _auto))
