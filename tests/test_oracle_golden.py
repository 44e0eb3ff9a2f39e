"""Pin the CPU oracle against the reference: fixtures recorded from the unmodified reference harness
(tests/golden/make_golden.py) and the known answers of the reference's own tests."""
import numpy as np
import pytest

from golden_io import GROUP_OF, N_PARAMS, SCENARIOS, Golden, assert_close, tolerances
from oracle import gradsync_oracle as O

FUSED = [n for n, sc in SCENARIOS.items() if sc["hp"]["optimizer"] in ("sgd", "adam", "adamw")]


@pytest.mark.parametrize("name", FUSED)
def test_oracle_reproduces_reference_run(name):
    g = Golden(name)
    tol = tolerances(g.dtype)
    if g.dtype == "float32":
        # fp32 is pinned element by element, tighter than the product tolerance
        def close(a, d, what=""):
            np.testing.assert_allclose(a, d, **tol, err_msg=what)
    else:
        def close(a, d, what=""):
            assert_close(a, d, g.dtype, what)
    opt = O.OptimizerOracle(g.hp["optimizer"], g.init_params(), GROUP_OF, g.groups(), g.dtype)
    for s in range(g.steps):
        out = O.train_window(opt, g.raw(s), g.agg, g.clip, average=g.average)
        for i in range(N_PARAMS):
            close(out["avg_grads"][i], g.per_step("avg_grads", s)[i], f"avg grad step {s} param {i}")
            close(out["clipped_grads"][i], g.per_step("clipped_grads", s)[i], f"clipped step {s} param {i}")
            close(opt.params[i], g.per_step("params", s)[i], f"param step {s} param {i}")
            if g.hp["optimizer"] == "sgd":
                close(opt.buf[i], g.state(s, i, "momentum_buffer"))
            else:
                close(opt.m[i], g.state(s, i, "exp_avg"))
                close(opt.v[i], g.state(s, i, "exp_avg_sq"))
                assert int(g.state(s, i, "step")) == opt.steps[i] == s + 1  # integer: exact
        if g.clip is not None and g.clip[0] == "norm":
            close(out["total_norm"], g.z["total_norm"][s])


@pytest.mark.parametrize("name", list(SCENARIOS))
def test_reported_counters_and_metrics(name):
    """Integer counters are exact; the chief's averaged loss equals the oracle's average of the
    per-rank losses."""
    g = Golden(name)
    assert g.z["reported_steps"].tolist() == [(s + 1) * g.agg for s in range(g.steps)]
    for s in range(g.steps):
        per_rank = {"loss": [[np.float32(v) for v in g.loss(r)[s * g.agg:(s + 1) * g.agg]] for r in range(g.slots)]}
        per_batch = O.average_training_metrics(per_rank, [g.agg] * g.slots)
        got = O.make_metrics(per_batch)["avg_metrics"]["loss"]
        np.testing.assert_allclose(got, g.z["reported_avg_loss"][s], rtol=1e-6)
        np.testing.assert_allclose([b["loss"] for b in per_batch], g.z["reported_batch_loss"][s], rtol=1e-6)


def test_should_communicate_and_update():
    # harness/determined/pytorch/_pytorch_context.py:711-716
    assert [O.should_communicate_and_update(b, 4) for b in range(8)] == [False, False, False, True] * 2
    assert all(O.should_communicate_and_update(b, 1) for b in range(5))
    assert O.should_communicate_and_update(None, 4, managed=False)
    with pytest.raises(RuntimeError):
        O.should_communicate_and_update(None, 4)


def test_simple_reduce_metrics_known_answers():
    # harness/tests/experiment/pytorch/test_reducer.py:18-29
    metrics = np.array([0.25, 0.5, 0.75, 1, 25.5, 1.9])
    assert np.around(O.simple_reduce_metrics(O.AVG, metrics), decimals=2) == 4.98
    assert O.simple_reduce_metrics(O.SUM, metrics) == 29.9
    assert O.simple_reduce_metrics(O.MIN, metrics) == 0.25
    assert O.simple_reduce_metrics(O.MAX, metrics) == 25.5
    num_batches = [1, 2, 5, 4, 5, 6]
    assert np.around(O.simple_reduce_metrics(O.AVG, metrics, num_batches), decimals=2) == 6.43
    with pytest.raises(RuntimeError):
        O.simple_reduce_metrics(O.AVG, metrics, [1, 2])


def test_average_training_metrics_known_answers():
    # harness/tests/experiment/pytorch/test_metric_utils.py:13-86
    assert O.process_combined_metrics_and_batches([({}, 2), ({}, 2)]) == ({}, [2, 2])
    combined = [({}, 0), ({}, 0), ({"loss1": [1, 2], "loss2": [-1, -2]}, 2), ({"loss1": [3, 4], "loss2": [-3, -4]}, 2)]
    lists, batches = O.process_combined_metrics_and_batches(combined)
    assert lists == {"loss1": [[1, 2], [3, 4]], "loss2": [[-1, -2], [-3, -4]]}
    assert batches == [2, 2]
    assert O.average_training_metrics(lists, batches) == [{"loss1": 2, "loss2": -2}, {"loss1": 3, "loss2": -3}]
    arr = {"a": [[np.array(1.0), np.array(3.0)], [np.array(2.0), np.array(5.0)]]}
    out = O.average_training_metrics(arr, [2, 2])
    assert isinstance(out[0]["a"], np.ndarray) and out[0]["a"].shape == () and out[0]["a"] == 1.5 and out[1]["a"] == 4.0
    # None values are ignored
    out = O.average_training_metrics({"a": [[1.0, None], [3.0, 4.0]]}, [2, 2])
    assert out == [{"a": 2.0}, {"a": 4.0}]


def test_identity_closed_form():
    # harness/tests/experiment/pytorch/test_pytorch_trial.py:1604-1642 (first values quoted in SURVEY 8c)
    w = O.identity_expected_weights()
    assert len(w) == 48
    np.testing.assert_allclose(w[:4], [0.0, 0.00375, 0.00375, 0.0130649375], rtol=1e-12)


def test_bf16_round_matches_torch():
    import torch

    x = np.random.default_rng(0).standard_normal(10000).astype(np.float32) * 3
    x[:4] = [0.0, -0.0, 1.00390625, 3.3895314e38]
    want = torch.from_numpy(x).to(torch.bfloat16).float().numpy()
    np.testing.assert_array_equal(O.bf16_round(x), want)


@pytest.mark.parametrize("name", [n for n, sc in SCENARIOS.items() if sc["slots"] == 1])
def test_torch_port_reproduces_reference_run(name):
    """oracle/torch_port.py (the timed CPU baseline: the reference's step_optimizer lines over torch) is pinned
    too: driven with the fixtures' initial weights and data it must retrace the real reference harness's
    single-slot runs -- losses, synchronized gradients, total norms and parameters after every optimizer step."""
    import torch

    from golden_model import GoldenNet, make_data, make_optimizer
    from oracle import torch_port

    g = Golden(name)
    torch.set_num_threads(1)
    dtype = getattr(torch, g.dtype)
    model = GoldenNet().to(dtype)
    with torch.no_grad():
        for p, init in zip(model.parameters(), g.init_params()):
            p.copy_(torch.from_numpy(init).to(dtype))
    ctx = torch_port.RefContext(aggregation_frequency=g.agg, hparams=g.hp)
    model = ctx.wrap_model(model)
    opt = ctx.wrap_optimizer(make_optimizer(model, g.hp), average_aggregated_gradients=g.average)
    x, y = make_data(g.steps * g.agg, g.hp["global_batch_size"])
    x, y = x.to(dtype), y.to(dtype)
    bs = ctx.get_per_slot_batch_size()
    seen = {"avg": [], "norm": []}

    def clip(params):
        params = list(params)
        seen["avg"].append([p.grad.detach().float().numpy().copy() for p in params])
        if g.clip is not None:
            kind, val = g.clip
            if kind == "norm":
                seen["norm"].append(float(torch.nn.utils.clip_grad_norm_(params, val)))
            else:
                torch.nn.utils.clip_grad_value_(params, val)

    tight = dict(rtol=1e-6, atol=1e-8) if g.dtype == "float32" else dict(rtol=1e-2, atol=1e-3)
    for i in range(g.steps * g.agg):
        ctx._current_batch_idx = i
        xb, yb = x[i * bs:(i + 1) * bs], y[i * bs:(i + 1) * bs]
        loss = torch.nn.functional.mse_loss(model(xb).float(), yb.float())
        np.testing.assert_allclose(float(loss), g.loss(0)[i], **tight)
        ctx.backward(loss)
        ctx.step_optimizer(opt, clip_grads=clip)
        if (i + 1) % g.agg == 0:
            s = (i + 1) // g.agg - 1
            for p, want in zip(model.parameters(), g.per_step("params", s)):
                np.testing.assert_allclose(p.detach().float().numpy(), want, **tight)
    opt_order = [0, 4, 1, 2, 3, 5]
    for s in range(g.steps):
        want = g.per_step("avg_grads", s)
        for j, i in enumerate(opt_order):
            np.testing.assert_allclose(seen["avg"][s][j], want[i], **tight)
    if g.clip is not None and g.clip[0] == "norm":
        np.testing.assert_allclose(seen["norm"], g.z["total_norm"], **tight)
