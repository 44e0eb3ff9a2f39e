"""world_size-2 CPU tests over gloo of what the N>1 path relies on:

  * the oracle's model of the reference's multi-rank gradient path (local accumulation under no_sync,
    DDP's grad/N + allreduce, the division by aggregation_frequency) against REAL torch DDP on gloo, live,
    on random data -- the same check the golden fixtures make offline;
  * the host-side collectives of this repo (DistributedContext over ZeroMQ between real processes, the
    training-metric averaging and custom reducers built on them) against torch.distributed's own
    all_gather_object.
No GPU and no libdgs compute call is involved.
"""
import os
import socket
import sys
import pathlib

import numpy as np
import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, size: int, port: int, zmq_ports, out_dir: str) -> None:
    for p in (str(ROOT), str(ROOT / "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    from determined_clone_b200 import core, pytorch
    from determined_clone_b200.pytorch import _metric_utils
    from oracle import gradsync_oracle as O
    from oracle import torch_port

    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=size)
    dctx = core.DistributedContext(rank=rank, size=size, local_rank=rank, local_size=size, cross_rank=0, cross_size=1,
                                   pub_port=zmq_ports[0], pull_port=zmq_ports[1])
    try:
        # ---- 1. reference path on gloo DDP vs the oracle --------------------------------------------------
        agg = 3
        torch.manual_seed(0)
        model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 2))
        ref = torch_port.RefContext(aggregation_frequency=agg)
        ddp = torch.nn.parallel.DistributedDataParallel(model)
        ref.models.append(ddp)
        opt = ref.wrap_optimizer(torch.optim.SGD(ddp.parameters(), lr=0.0))  # lr 0: keep weights, look at gradients
        params = list(model.parameters())
        raw = [[None] * len(params) for _ in range(agg)]
        g = torch.Generator().manual_seed(50 + rank)
        seen = {}

        def grab(ps):
            seen["synced"] = [p.grad.detach().numpy().copy() for p in ps]

        for m in range(agg):
            hooks = [p.register_hook(lambda gr, i=i, m=m: raw[m].__setitem__(i, gr.detach().numpy().copy())) for i, p in enumerate(params)]
            ref._current_batch_idx = m
            x, y = torch.randn(4, 6, generator=g), torch.randn(4, 2, generator=g)
            ref.backward(torch.nn.functional.mse_loss(ddp(x), y))
            ref.step_optimizer(opt, clip_grads=grab)
            for h in hooks:
                h.remove()
        everyone = [None] * size
        dist.all_gather_object(everyone, raw)
        for i in range(len(params)):
            want = O.synced_gradients([[everyone[r][m][i] for m in range(agg)] for r in range(size)], agg)
            np.testing.assert_allclose(seen["synced"][i], want, rtol=1e-6, atol=1e-8)

        # ---- 2. DistributedContext collectives between real processes vs torch.distributed -----------------
        payload = {"rank": rank, "v": [rank * 1.5, None]}
        mine = dctx.allgather(payload)
        theirs = [None] * size
        dist.all_gather_object(theirs, payload)
        assert mine == theirs
        assert dctx.broadcast("chief says hi" if rank == 0 else None) == "chief says hi"
        gathered = dctx.gather(rank)
        assert gathered == (list(range(size)) if rank == 0 else None)

        # ---- 3. training-metric averaging and a custom reducer on top of them -------------------------------
        per_batch = [{"loss": np.float32(rank + b), "acc": None if (rank == 1 and b == 0) else 0.25 * (rank + 1)} for b in range(3)]
        avg = _metric_utils._combine_and_average_training_metrics(dctx, per_batch)
        if rank == 0:
            assert [float(a["loss"]) for a in avg] == [0.5 + b for b in range(3)]
            assert [a["acc"] for a in avg] == [0.25, 0.375, 0.375]  # the None of rank 1, batch 0 is ignored
        rc = pytorch._PyTorchReducerContext(dctx.allgather)
        r = rc.wrap_reducer(lambda vals: sum(vals), name="total")
        for b in range(rank, 7, size):  # batches dealt round-robin
            r.update(b)
        assert rc.reduce_metrics(for_training=True) == {"total": 21}

        # ---- 4. the training loop around the path on 2 ranks: callbacks, validation, checkpoints -------------
        class Hooks(pytorch.PyTorchCallback):
            def __init__(self):
                self.workloads, self.validations, self.uploads, self.written = [], [], [], 0

            def on_training_workload_end(self, avg_metrics, batch_metrics):
                self.workloads.append((float(avg_metrics["loss"]), len(batch_metrics)))

            def on_validation_end(self, metrics):
                self.validations.append(float(metrics["v"]))

            def on_checkpoint_write_end(self, checkpoint_dir):
                self.written += 1

            def on_checkpoint_upload_end(self, uuid):
                self.uploads.append(uuid)

        class LoopTrial(pytorch.PyTorchTrial):
            def __init__(self, context):
                self.context = context  # no wrap_model: with several slots that brings up NCCL, and this is a CPU test
                self.hooks = Hooks()

            def build_callbacks(self):
                return {"hooks": self.hooks}

            def build_training_data_loader(self):
                return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(32).float()), batch_size=2)

            def build_validation_data_loader(self):
                return pytorch.DataLoader(torch.utils.data.TensorDataset(torch.arange(8).float()), batch_size=2)

            def train_batch(self, batch, epoch_idx, batch_idx):
                return {"loss": float(batch[0][0])}  # first record of this rank's batch

            def evaluate_batch(self, batch, batch_idx):
                return {"v": float(batch[0].sum())}

        lctx = pytorch.PyTorchTrialContext(distributed=dctx, hparams={"global_batch_size": 4}, slots_per_trial=size)
        ltrial = LoopTrial(lctx)
        trainer = pytorch.Trainer(ltrial, lctx)
        reports = trainer.fit(max_length=pytorch.Batch(4), validation_period=pytorch.Batch(2), reporting_period=pytorch.Batch(2),
                              checkpoint_policy="all", checkpoint_dir=os.path.join(out_dir, "ckpt"))
        h = ltrial.hooks
        # rank r sees global batches r, r+2, ...: batch k of rank r starts at record 2*(2k + r); mean over ranks = 4k + 1
        assert h.workloads == [(np.mean([1.0, 5.0]), 2), (np.mean([9.0, 13.0]), 2)]  # every rank gets the chief's averages
        # 4 validation batches dealt over 2 ranks; AVG over all of them = mean of (1, 5, 9, 13)
        assert h.validations == [7.0, 7.0]
        assert len(h.uploads) == 2 and h.uploads == dctx.broadcast(h.uploads)  # same uuids on every rank
        assert h.written == (2 if rank == 0 else 0)  # only the chief writes
        assert (len(reports), len(trainer.validation_reports)) == ((2, 2) if rank == 0 else (0, 0))
        assert sorted(os.listdir(os.path.join(out_dir, "ckpt"))) == sorted(h.uploads)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dctx.close()
        dist.destroy_process_group()


def test_world_size_2_over_gloo(tmp_path):
    import torch.multiprocessing as mp

    mp.spawn(_worker, args=(2, _free_port(), (_free_port(), _free_port()), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
