"""import_from_path: import modules that live in a checkpoint's code/ directory without leaving them behind
(reference harness/determined/_import.py:70-130).  Inside the block the directory is first on sys.path; on exit
every module that was imported from under it is dropped from sys.modules again, so a second checkpoint with a file
of the same name imports its own copy."""
import contextlib
import os
import pathlib
import sys
from typing import Any, Iterator


@contextlib.contextmanager
def import_from_path(path: Any) -> Iterator[None]:
    root = str(pathlib.Path(path).resolve())
    before = set(sys.modules)
    sys.path.insert(0, root)
    try:
        yield
    finally:
        try:
            sys.path.remove(root)
        except ValueError:
            pass
        for name in set(sys.modules) - before:
            origin = getattr(sys.modules[name], "__file__", None)
            if origin and os.path.realpath(origin).startswith(root + os.sep):
                del sys.modules[name]
