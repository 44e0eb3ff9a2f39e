"""Behaviour of the context + engine on cases the golden scenarios do not reach, checked against the
reference's own logic running on the same GPU (oracle/torch_port.RefContext = the reference's
step_optimizer lines over torch.optim), and size-independent properties at BASELINE.json's full bucket
sizes (1-2 GB), where an element-wise oracle would be too slow."""
import copy
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from torch.nn.utils import clip_grad_norm_ as imported_clip  # noqa: E402  (bound at import time, on purpose)

import determined_clone_b200 as det  # noqa: E402
from determined_clone_b200 import _lib as L, pytorch  # noqa: E402
from oracle import torch_port  # noqa: E402


class Net(torch.nn.Module):
    def __init__(self, unused: bool = False, dtype=torch.float32):
        super().__init__()
        self.a = torch.nn.Linear(12, 33)
        self.b = torch.nn.Linear(33, 5)
        self.unused = torch.nn.Parameter(torch.ones(7)) if unused else None
        self.to(dtype)

    def forward(self, x):
        return self.b(torch.tanh(self.a(x)))


def make_contexts(agg=1, **kw):
    ours = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, aggregation_frequency=agg, device=torch.device("cuda", 0), **kw)
    ref = torch_port.RefContext(aggregation_frequency=agg, device="cuda:0")
    return ours, ref


def run_both(build_opt, steps=6, agg=1, clip=None, losses=1, wrap_kw=None, net_kw=None, scheduler=None, after_step=None, rtol=1e-5):
    """Train the same net/data under both contexts; compare parameters after every optimizer step."""
    torch.manual_seed(3)
    base = Net(**(net_kw or {}))
    data = [(torch.randn(8, 12, device="cuda"), torch.randn(8, 5, device="cuda")) for _ in range(steps * agg)]
    ours, ref = make_contexts(agg)
    out = []
    for ctx in (ours, ref):
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(build_opt(model), **(wrap_kw or {}))
        sched = scheduler(opt) if scheduler else None
        snaps = []
        for i, (x, y) in enumerate(data):
            ctx._current_batch_idx = i
            x, y = x.to(next(model.parameters()).dtype), y.to(next(model.parameters()).dtype)
            pred = model(x)
            if losses == 1:
                ctx.backward(torch.nn.functional.mse_loss(pred.float(), y.float()))
            else:
                ctx.backward(torch.nn.functional.mse_loss(pred[:, :2].float(), y[:, :2].float()), retain_graph=True) if isinstance(ctx, pytorch.PyTorchTrialContext) else \
                    torch.nn.functional.mse_loss(pred[:, :2].float(), y[:, :2].float()).backward(retain_graph=True)
                ctx.backward(torch.nn.functional.l1_loss(pred[:, 2:].float(), y[:, 2:].float()))
            ctx.step_optimizer(opt, clip_grads=clip)
            if (i + 1) % agg == 0:
                if sched:
                    sched.step()
                if after_step:
                    after_step(ctx, model, opt, (i + 1) // agg)
                snaps.append([p.detach().float().cpu().numpy().copy() for p in model.parameters()])
        out.append((model, opt, snaps))
    (m1, o1, s1), (m2, o2, s2) = out
    for k, (a, b) in enumerate(zip(s1, s2)):
        for pa, pb in zip(a, b):
            np.testing.assert_allclose(pa, pb, rtol=rtol, atol=rtol * max(1e-3, float(np.abs(pb).max())), err_msg=f"after optimizer step {k}")
    ours.close()
    return out


def test_backward_passes_per_step_two_losses(lib):
    run_both(lambda m: torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9), agg=2, losses=2, wrap_kw={"backward_passes_per_step": 2})


def test_extra_backward_after_sync_raises(lib):
    ours, _ = make_contexts()
    model = ours.wrap_model(Net())
    opt = ours.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1))
    ours._current_batch_idx = 0
    loss = model(torch.randn(8, 12, device="cuda")).sum()
    ours.backward(loss, retain_graph=True)
    with pytest.raises(det.errors.InvalidExperimentException, match="backward_passes_per_step"):
        ours.backward(loss)
    ours.close()


def test_unused_parameter_is_skipped_like_torch(lib):
    out = run_both(lambda m: torch.optim.AdamW(m.parameters(), lr=1e-2, weight_decay=0.1), net_kw={"unused": True})
    (m1, o1, _), (m2, o2, _) = out
    assert torch.equal(m1.unused.detach().cpu(), torch.ones(7))  # no weight decay applied either
    assert float(o1.state[m1.a.weight]["step"]) == 6.0
    assert float(o1.state[m1.unused].get("step", 0.0)) == 0.0  # torch creates no state for it; ours never steps it


def test_average_aggregated_gradients_false(lib):
    run_both(lambda m: torch.optim.SGD(m.parameters(), lr=0.01), agg=3, wrap_kw={"average_aggregated_gradients": False})


def test_lr_scheduler_is_reread_every_step(lib):
    run_both(lambda m: torch.optim.SGD(m.parameters(), lr=0.2, momentum=0.5), steps=8,
             scheduler=lambda o: torch.optim.lr_scheduler.StepLR(o, step_size=2, gamma=0.5))


@pytest.mark.parametrize("opt", ["sgd_plain", "sgd_nesterov", "sgd_maximize", "adam_l2", "adam_amsgrad", "rmsprop"])
def test_optimizer_variants(lib, opt):
    build = {
        "sgd_plain": lambda m: torch.optim.SGD(m.parameters(), lr=0.05, weight_decay=1e-3),
        "sgd_nesterov": lambda m: torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9, nesterov=True),
        "sgd_maximize": lambda m: torch.optim.SGD(m.parameters(), lr=0.01, momentum=0.9, dampening=0.1, maximize=True),
        "adam_l2": lambda m: torch.optim.Adam(m.parameters(), lr=1e-2, weight_decay=0.05),
        "adam_amsgrad": lambda m: torch.optim.Adam(m.parameters(), lr=1e-2, amsgrad=True),  # not fused: torch's own step on views
        "rmsprop": lambda m: torch.optim.RMSprop(m.parameters(), lr=1e-3),  # not fused
    }[opt]
    run_both(build, clip=lambda ps: torch.nn.utils.clip_grad_norm_(ps, 0.1))


def test_clip_functions_of_every_shape(lib):
    norms = []

    def imported(ps):  # `from torch.nn.utils import clip_grad_norm_` at module import time
        norms.append(imported_clip(ps, 0.05))

    def by_value(ps):
        torch.nn.utils.clip_grad_value_(ps, 0.01)

    def replaces_grad(ps):  # assigns new tensors instead of editing in place
        for p in ps:
            p.grad = p.grad * 0.5

    def partial(ps):  # clips only some parameters: must fall through to torch untouched
        torch.nn.utils.clip_grad_norm_(list(ps)[:2], 0.01)

    for fn in (imported, by_value, replaces_grad, partial):
        run_both(lambda m: torch.optim.AdamW(m.parameters(), lr=1e-2), clip=fn, agg=2)
    assert len(norms) == 12 and all(torch.is_tensor(n) for n in norms)
    # ours (first 6) and the reference's (last 6) total norms agree
    np.testing.assert_allclose([float(n) for n in norms[:6]], [float(n) for n in norms[6:]], rtol=1e-5)


def test_load_state_dict_is_adopted(lib):
    saved = {}

    def after(ctx, model, opt, k):
        if k == 3:
            if "state" not in saved:  # first run (ours) saves; both runs then reload the same checkpoint
                saved["state"] = copy.deepcopy(opt.state_dict())
                saved["model"] = copy.deepcopy(model.state_dict())
            model.load_state_dict(saved["model"])
            opt.load_state_dict(copy.deepcopy(saved["state"]))

    for build in (lambda m: torch.optim.AdamW(m.parameters(), lr=1e-2, weight_decay=0.1),
                  lambda m: torch.optim.SGD(m.parameters(), lr=0.05, momentum=0.9)):
        saved.clear()
        run_both(build, steps=7, after_step=after)


def test_state_dict_round_trip_has_reference_layout(lib):
    ours, _ = make_contexts()
    model = ours.wrap_model(Net())
    opt = ours.wrap_optimizer(torch.optim.AdamW(model.parameters(), lr=1e-3))
    for i in range(2):
        ours._current_batch_idx = i
        ours.backward(model(torch.randn(8, 12, device="cuda")).pow(2).mean())
        ours.step_optimizer(opt)
    sd = opt.state_dict()
    assert set(sd["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"}
    assert float(sd["state"][0]["step"]) == 2.0 and sd["state"][0]["exp_avg"].shape == model.a.weight.shape
    assert list(model.state_dict().keys()) == ["a.weight", "a.bias", "b.weight", "b.bias"]  # no "module." prefix
    fresh = torch.optim.AdamW(Net().cuda().parameters(), lr=1e-3)
    fresh.load_state_dict(sd)  # a plain torch optimizer accepts it
    ours.close()


def test_mixed_dtype_optimizer(lib):
    torch.manual_seed(5)
    ours, ref = make_contexts()
    nets = []
    for ctx in (ours, ref):
        torch.manual_seed(5)
        m = Net()
        m.b.to(torch.bfloat16)
        nets.append(ctx.wrap_model(m))
    opts = [ctx.wrap_optimizer(torch.optim.AdamW(m.parameters(), lr=1e-2)) for ctx, m in zip((ours, ref), nets)]
    assert len(ours._engines[opts[0]].sets) == 2
    for i in range(4):
        x, y = torch.randn(8, 12, device="cuda"), torch.randn(8, 5, device="cuda")
        for ctx, m, o in zip((ours, ref), nets, opts):
            ctx._current_batch_idx = i
            h = torch.tanh(m.a(x)).to(torch.bfloat16)
            ctx.backward(torch.nn.functional.mse_loss(m.b(h).float(), y))
            ctx.step_optimizer(o, clip_grads=lambda ps: torch.nn.utils.clip_grad_norm_(ps, 0.5))
    # torch rounds each bf16 tensor's norm to bf16 before combining them, so its clip coefficient carries a
    # bf16-sized error that also scales the fp32 gradients; ours is computed in fp32 throughout
    for pa, pb in zip(nets[0].parameters(), nets[1].parameters()):
        tol = 2e-3 if pa.dtype == torch.float32 else 2e-2
        np.testing.assert_allclose(pa.detach().float().cpu().numpy(), pb.detach().float().cpu().numpy(), rtol=tol, atol=tol * 0.3)
    ours.close()


def test_unmanaged_context_steps_every_batch(lib):
    ours = pytorch.PyTorchTrialContext(managed_training=False, aggregation_frequency=4, device=torch.device("cuda", 0))
    model = Net().cuda()
    opt = ours.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1))
    before = model.a.weight.detach().clone()
    ours.backward(model(torch.randn(8, 12, device="cuda")).sum())
    ours.step_optimizer(opt)  # _should_communicate_and_update() is always True when unmanaged
    assert not torch.equal(before, model.a.weight)
    ours.close()


def test_mnist_tutorial_trial_ten_steps_match_reference_path(lib):
    """BASELINE.json configs[0]: the MNIST tutorial trial, single slot, fixed seed, 10 steps -- here under this
    repo's context vs the reference's step logic (torch_port.RefContext) on the same GPU, same weights, same
    batches.  Adadelta is not fused, so this is the flatten -> gradient views -> torch step() route."""
    from workloads.trials import MnistTrial

    torch.backends.cudnn.deterministic = True
    gen = torch.Generator().manual_seed(1234)
    batches = [tuple(t.cuda() for t in MnistTrial.synthetic_batch(64, gen)) for _ in range(10)]
    results = []
    for make in (lambda: pytorch.PyTorchTrialContext(hparams={"global_batch_size": 64}, device=torch.device("cuda", 0)),
                 lambda: torch_port.RefContext(device="cuda:0")):
        torch.manual_seed(17)
        ctx = make()
        trial = MnistTrial(ctx)
        losses = []
        for i, b in enumerate(batches):
            ctx._current_batch_idx = i
            losses.append(float(trial.train_batch(b, 0, i)["loss"]))
        results.append((losses, [p.detach().cpu().numpy().copy() for p in trial.model.parameters()]))
    (l1, p1), (l2, p2) = results
    np.testing.assert_allclose(l1, l2, rtol=1e-5)
    for a, b in zip(p1, p2):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-5 * float(np.abs(b).max()))
    torch.backends.cudnn.deterministic = False


# ---- size-independent properties at full bucket sizes ---------------------------------------------------
def _engine(n, dtype, agg, opt_fn):
    p = torch.nn.Parameter(torch.full((n,), 0.5, device="cuda", dtype=dtype))
    opt = opt_fn([p])
    eng = pytorch.GradSyncEngine(opt, torch.device("cuda", 0), None, aggregation_frequency=agg, average_aggregated_gradients=True)
    return p, opt, eng


def _window(eng, p, grads, clip=None):
    for k, g in enumerate(grads):
        eng.begin_pass(final=k == len(grads) - 1)
        p.grad = g
        eng.end_pass()
    present = eng.finish_window()
    req = None
    if clip is not None:
        req = pytorch._gradsync.ClipRequest()
        req.max_norm = clip
        eng.total_norm()
    eng.fused_step(present, req)
    eng.reset_window()


@pytest.mark.parametrize("mb,dtype", [(1, torch.float32), (2048, torch.float32), (2048, torch.bfloat16)])
def test_full_size_bucket_properties(lib, mb, dtype):
    n = mb * (1 << 20) // (4 if dtype == torch.float32 else 2)
    # (1) cancellation: accumulating g and -g gives an exactly zero gradient -> only the decoupled decay acts
    p, opt, eng = _engine(n, dtype, 2, lambda ps: torch.optim.AdamW(ps, lr=0.5, weight_decay=0.5, eps=1e-8))
    g = (torch.randn(1 << 20, device="cuda") * 1e-2).to(dtype).repeat(n // (1 << 20) or 1)[:n].contiguous()
    _window(eng, p, [g, -g])
    torch.cuda.synchronize()
    assert torch.all(eng.sets[0].wire == 0)
    assert torch.all(p.detach() == 0.5 * (1 - 0.5 * 0.5))  # p * (1 - lr*wd) exactly; m = v = 0
    assert torch.all(opt.state[p]["exp_avg"] == 0) and torch.all(opt.state[p]["exp_avg_sq"] == 0)
    # (2) idempotence: lr = 0, wd = 0 leaves parameters bit-identical, whatever the gradient
    opt.param_groups[0]["lr"], opt.param_groups[0]["weight_decay"] = 0.0, 0.0
    before = p.detach().clone()
    _window(eng, p, [g, g], clip=1.0)
    torch.cuda.synchronize()
    assert torch.equal(before, p.detach())
    # (3) the averaged gradient of (g, g) is g again (scale 1/2 is exact), and the norm is linear in a scalar
    wire = eng.sets[0].wire
    assert torch.equal(wire, g)
    norm1 = float(eng._norm)
    _window(eng, p, [g * 4, g * 4], clip=1.0)
    torch.cuda.synchronize()
    np.testing.assert_allclose(float(eng._norm), 4 * norm1, rtol=2e-6 if dtype == torch.float32 else 1e-3)
    want = float(torch.linalg.vector_norm(g[: 1 << 20].double())) * 4 * np.sqrt(max(n // (1 << 20), 1)) if n >= (1 << 20) else float(torch.linalg.vector_norm(g.double())) * 4
    np.testing.assert_allclose(float(eng._norm), want, rtol=1e-5 if dtype == torch.float32 else 2e-3)
    eng.close()
    del p, opt, eng, g, wire, before
    torch.cuda.empty_cache()


class _ValTrial(pytorch.PyTorchTrial):
    """evaluate_batch returns device scalars; three reducer kinds at once."""

    def __init__(self, context):
        self.context = context
        self.model = context.wrap_model(torch.nn.Linear(4, 1))
        self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=0.1))
        g = torch.Generator().manual_seed(3)
        self.data = torch.utils.data.TensorDataset(torch.randn(22, 4, generator=g), torch.randn(22, 1, generator=g))

    def build_training_data_loader(self):
        return pytorch.DataLoader(self.data, batch_size=4)

    def build_validation_data_loader(self):
        return pytorch.DataLoader(self.data, batch_size=4)  # 6 batches, the last one short

    def train_batch(self, batch, epoch_idx, batch_idx):
        return {"loss": self.model(batch[0]).mean()}

    def evaluate_batch(self, batch, batch_idx=0):
        err = (self.model(batch[0]) - batch[1]).abs()
        return {"mean_err": err.mean(), "max_err": err.max(), "n": torch.tensor(float(len(err)), device=err.device)}

    def evaluation_reducer(self):
        return {"mean_err": pytorch.Reducer.AVG, "max_err": pytorch.Reducer.MAX, "n": pytorch.Reducer.SUM}


def test_validation_metrics_reduced_on_device_match_host_route(lib):
    outs = []
    for device_metrics in (True, False):
        torch.manual_seed(1)
        ctx = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 4}, device=torch.device("cuda", 0))
        trial = _ValTrial(ctx)
        ctl = pytorch._PyTorchTrialController(trial, ctx, max_length=1, device_metrics=device_metrics)
        ctl._set_data_loaders()
        outs.append(ctl._validate())
        ctx.close()
    dev, host = outs
    assert set(dev) == set(host) == {"mean_err", "max_err", "n"}
    assert float(dev["n"]) == float(host["n"]) == 22.0  # a count: exact
    for k in dev:
        np.testing.assert_allclose(float(dev[k]), float(host[k]), rtol=1e-6)


def test_grad_scaler_route_matches_reference(lib):
    """AMP with a GradScaler passed to step_optimizer (reference _pytorch_context.py:899-921): the scaler works on
    the engine's gradient views; the step stays in torch so that inf/nan skipping keeps its semantics."""
    torch.manual_seed(9)
    base = Net()
    data = [(torch.randn(8, 12, device="cuda"), torch.randn(8, 5, device="cuda")) for _ in range(5)]
    results = []
    for which in ("ours", "ref"):
        ours, ref = make_contexts()
        ctx = ours if which == "ours" else ref
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9))
        scaler = torch.amp.GradScaler("cuda", init_scale=1024.0)
        for i, (x, y) in enumerate(data):
            ctx._current_batch_idx = i
            loss = torch.nn.functional.mse_loss(model(x), y)
            ctx.backward(scaler.scale(loss))
            if which == "ours":
                ctx.step_optimizer(opt, scaler=scaler)
            else:
                scaler.step(opt)
                opt.zero_grad()
            scaler.update()
        results.append([p.detach().cpu().numpy().copy() for p in model.parameters()])
        ours.close()
    for a, b in zip(*results):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-7)


def test_checkpoint_files_interoperate_with_the_reference_layout(lib, tmp_path):
    """state_dict.pth written here loads into plain torch objects (what the reference's _load does), and a
    checkpoint in the reference's layout -- DDP-prefixed model keys, torch.optim state -- loads here and training
    continues exactly as plain torch continues from it."""
    torch.manual_seed(4)
    base = Net()
    data = [(torch.randn(8, 12, device="cuda"), torch.randn(8, 5, device="cuda")) for _ in range(8)]

    def train(ctx, model, opt, rng):
        for i in rng:
            ctx._current_batch_idx = i
            ctx.backward(torch.nn.functional.mse_loss(model(data[i][0]), data[i][1]))
            ctx.step_optimizer(opt)

    # (1) reference-side: plain torch trains 4 batches and saves with the reference's keys, DDP-style names
    ref = torch_port.RefContext(device="cuda:0")
    rmodel = ref.wrap_model(copy.deepcopy(base))
    ropt = ref.wrap_optimizer(torch.optim.AdamW(rmodel.parameters(), lr=1e-2, weight_decay=0.1))
    train(ref, rmodel, ropt, range(4))
    ckpt = tmp_path / "from_reference"
    ckpt.mkdir()
    torch.save({"models_state_dict": [{"module." + k: v for k, v in rmodel.state_dict().items()}],
                "optimizers_state_dict": [ropt.state_dict()], "lr_schedulers_state_dict": [], "callbacks": {},
                "rng_state": {"cpu_rng_state": torch.random.get_rng_state(), "np_rng_state": np.random.get_state(),
                              "random_rng_state": __import__("random").getstate()}}, ckpt / "state_dict.pth")
    train(ref, rmodel, ropt, range(4, 8))  # the reference continues

    # (2) ours: load that checkpoint into a fresh trial and continue
    ours = pytorch.PyTorchTrialContext(hparams={"global_batch_size": 8}, device=torch.device("cuda", 0))
    model = ours.wrap_model(Net())
    opt = ours.wrap_optimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.1))
    ctl = pytorch._PyTorchTrialController(None, ours, max_length=0)  # type: ignore
    ctl._load(ckpt)
    train(ours, model, opt, range(4, 8))
    for a, b in zip(model.parameters(), rmodel.parameters()):
        np.testing.assert_allclose(a.detach().cpu().numpy(), b.detach().cpu().numpy(), rtol=1e-5, atol=1e-7)
    assert float(opt.state[model.a.weight]["step"]) == 8.0

    # (3) ours saves; plain torch loads and agrees
    ctl.batches_trained = 8
    ctl._save(tmp_path / "from_ours")
    saved = torch.load(tmp_path / "from_ours" / "state_dict.pth", weights_only=False)
    assert set(saved) >= {"models_state_dict", "optimizers_state_dict", "lr_schedulers_state_dict", "rng_state"}
    fresh = Net().cuda()
    fresh.load_state_dict(saved["models_state_dict"][0])
    fopt = torch.optim.AdamW(fresh.parameters(), lr=1e-2, weight_decay=0.1)
    fopt.load_state_dict(saved["optimizers_state_dict"][0])
    for a, b in zip(fresh.parameters(), rmodel.parameters()):
        np.testing.assert_allclose(a.detach().cpu().numpy(), b.detach().cpu().numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(fopt.state[fresh.a.weight]["exp_avg"].cpu().numpy(), ropt.state[rmodel.a.weight]["exp_avg"].cpu().numpy(), rtol=1e-5, atol=1e-9)
    ctl2 = pytorch._PyTorchTrialController(None, ours, max_length=0)  # type: ignore
    ctl2._load(tmp_path / "from_ours")
    assert ctl2.batches_trained == 8  # integer: exact
    ours.close()


def test_fp16_wire_compression(lib):
    """optimizations.gradient_compression (Horovod's Compression.fp16, reference _pytorch_context.py:357-366):
    fp32 gradients travel as fp16; the fused step reads the fp16 wire directly.  Checked against torch stepping
    on the same fp16-rounded averaged gradients."""
    torch.manual_seed(2)
    n = 10000
    p = torch.nn.Parameter(torch.randn(n, device="cuda"))
    q = torch.nn.Parameter(p.detach().clone())
    opt = torch.optim.SGD([p], lr=0.1, momentum=0.9, weight_decay=1e-3)
    ref = torch.optim.SGD([q], lr=0.1, momentum=0.9, weight_decay=1e-3)
    eng = pytorch.GradSyncEngine(opt, torch.device("cuda", 0), None, aggregation_frequency=2, average_aggregated_gradients=True,
                                 fp16_compression=True)
    assert eng.sets[0].wire.dtype == torch.float16
    for step in range(4):
        g1, g2 = torch.randn(n, device="cuda") * 0.1, torch.randn(n, device="cuda") * 0.1
        _window(eng, p, [g1.clone(), g2.clone()])
        q.grad = ((g1 + g2) * 0.5).half().float()
        ref.step()
        torch.cuda.synchronize()
        np.testing.assert_allclose(p.detach().cpu().numpy(), q.detach().cpu().numpy(), rtol=2e-6, atol=1e-7)
    eng.close()


def test_tiny_gpt2_with_tied_embeddings_matches_reference_path(lib):
    """BASELINE.json configs[3] in miniature: a GPT-2 causal LM (tied input/output embedding: one parameter used
    twice in the graph), AdamW with a decay and a no-decay group, aggregation_frequency 2, clip_grad_norm_ 1.0 --
    ours vs the reference's step logic on the same GPU, fp32, 6 optimizer steps."""
    transformers = pytest.importorskip("transformers")

    def build():
        torch.manual_seed(21)
        cfg = transformers.GPT2Config(vocab_size=211, n_positions=32, n_embd=64, n_layer=2, n_head=4,
                                      resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, use_cache=False)
        return transformers.GPT2LMHeadModel(cfg)

    g = torch.Generator().manual_seed(5)
    batches = [torch.randint(0, 211, (4, 32), generator=g).cuda() for _ in range(12)]
    results = []
    for which in ("ours", "ref"):
        ours, ref = make_contexts(agg=2)
        ctx = ours if which == "ours" else ref
        model = ctx.wrap_model(build())
        assert model.lm_head.weight is model.transformer.wte.weight  # tied
        decay = [p for p in model.parameters() if p.ndim >= 2]
        no_decay = [p for p in model.parameters() if p.ndim < 2]
        opt = ctx.wrap_optimizer(torch.optim.AdamW([{"params": decay, "weight_decay": 0.1}, {"params": no_decay, "weight_decay": 0.0}],
                                                   lr=1e-3, betas=(0.9, 0.95)))
        norms, losses = [], []
        for i, tok in enumerate(batches):
            ctx._current_batch_idx = i
            loss = model(input_ids=tok, labels=tok).loss
            losses.append(float(loss.detach()))
            ctx.backward(loss)
            ctx.step_optimizer(opt, clip_grads=lambda ps: norms.append(float(torch.nn.utils.clip_grad_norm_(ps, 1.0))))
        results.append((losses, norms, [p.detach().cpu().numpy().copy() for p in model.parameters()]))
        ours.close()
    (l1, n1, p1), (l2, n2, p2) = results
    assert len(n1) == len(n2) == 6
    np.testing.assert_allclose(l1, l2, rtol=2e-5)
    np.testing.assert_allclose(n1, n2, rtol=2e-5)
    # The key bias of attention has a mathematically zero gradient (softmax is invariant to it): what reaches
    # the optimizer is run-to-run rounding noise of the model's own backward kernels, which Adam divides by
    # sqrt(v) + eps and turns into moves of up to lr per step.  Those elements cannot agree to 1e-5 between any
    # two runs; they are bounded by a small fraction of the distance Adam can move at all (lr * steps).
    for a, b in zip(p1, p2):
        np.testing.assert_allclose(a, b, rtol=2e-5, atol=max(2e-5 * float(np.abs(b).max()), 2e-3 * 1e-3 * 6))
