"""The reference's OWN unit tests for this path, run unmodified against this package.

Where /root/reference exists (the build container), its test files are executed in place -- nothing is copied --
with `determined` resolving to determined_clone_b200/compat and `tests.parallel` (the reference's threads-as-ranks
helper) to the reference's tests package.  What runs:

  experiment/pytorch/test_pytorch_data.py     samplers, DataLoader adaptation, data_length, to_device   (24 tests)
  experiment/pytorch/test_reducer.py          _simple_reduce_metrics, custom reducer slot order          (5 tests)
  experiment/pytorch/test_metric_utils.py     metric combination / averaging / reducers                  (4 tests)
  experiment/pytorch/test_pytorch_context.py  context edge cases                                         (3 tests)
  test_ipc.py (ZMQ + DistributedContext)      broadcast server/client, collectives for 1x1 .. 4x4 ranks  (9 tests)

test_ipc.py's PIDServer tests belong to the launcher's process supervision and are deselected."""
import os
import pathlib
import re
import subprocess
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
REF_TESTS = pathlib.Path("/root/reference/harness/tests")

CASES = [
    ("experiment/pytorch/test_pytorch_data.py", None, 24),
    ("experiment/pytorch/test_reducer.py", None, 5),
    ("experiment/pytorch/test_metric_utils.py", None, 4),
    ("experiment/pytorch/test_pytorch_context.py", None, 3),
    ("test_ipc.py", "not pid and not PID", 9),
]


@pytest.mark.skipif(not REF_TESTS.exists(), reason="the reference tree is only present in the build container")
@pytest.mark.parametrize("rel,select,want", CASES)
def test_reference_unit_tests_pass_against_this_package(tmp_path, rel, select, want):
    env = dict(os.environ, PYTHONDONTWRITEBYTECODE="1",
               PYTHONPATH=os.pathsep.join([str(ROOT / "determined_clone_b200" / "compat"), str(ROOT), str(REF_TESTS.parent)]))
    cmd = [sys.executable, "-m", "pytest", "-p", "no:cacheprovider", "-q", "--import-mode=importlib", "--noconftest",
           "--rootdir", str(tmp_path), "-c", os.devnull, str(REF_TESTS / rel)]
    if select:
        cmd += ["-k", select]
    r = subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=900)
    tail = r.stdout[-3000:] + r.stderr[-2000:]
    assert r.returncode == 0, tail
    m = re.search(r"(\d+) passed", r.stdout)
    assert m and int(m.group(1)) == want and "failed" not in r.stdout, tail
