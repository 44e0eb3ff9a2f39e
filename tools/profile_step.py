#!/usr/bin/env python
"""Host-side cost of one train_batch through this repo's context (cProfile), to find Python overhead in the
hooks / engine when the workload is launch-bound."""
import cProfile
import pathlib
import pstats
import sys

import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from determined_clone_b200 import pytorch  # noqa: E402
from workloads.trials import CifarResNetTrial  # noqa: E402

ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 128}, device=torch.device("cuda", 0))
trial = CifarResNetTrial(ctx)
gen = torch.Generator().manual_seed(0)
batch = tuple(t.cuda() for t in CifarResNetTrial.synthetic_batch(128, gen))


def steps(n, start):
    for i in range(start, start + n):
        ctx._current_batch_idx = i
        trial.train_batch(batch, 0, i)
    torch.cuda.synchronize()


steps(10, 0)
pr = cProfile.Profile()
pr.enable()
steps(50, 10)
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
