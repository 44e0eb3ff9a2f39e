"""`python3 -m determined.exec.harness module:Trial` -> determined_clone_b200.exec.harness."""
import sys

from determined_clone_b200.exec.harness import main

if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
