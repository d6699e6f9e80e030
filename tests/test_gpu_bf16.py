"""bf16 tensor-core mode (tcgen05 path) against the oracle with the same rounding points
(oracle/mlp_numpy.py emulate_bf16=True). Predictions bit-exact on bf16-representable inputs and weights (any
mismatch is reported with its top-2 logit margin); loss and updated weights within 1e-2 relative."""
import numpy as np
import pytest
import torch

from oracle import mlp_numpy as onp
from tests.golden_util import load_case, replay, assert_params_close
from tests.test_gpu_parity import build_model, seeded_init, synthetic, net_params, EngineAdapter, OPT

pytestmark = pytest.mark.gpu


def check_predictions(got, want, oracle, x, tag):
    bad = np.nonzero(got != want)[0]
    if len(bad) == 0:
        return
    z, _, _ = oracle._forward(x, None)
    top2 = np.sort(z[bad], axis=1)[:, -2:]
    margin = (top2[:, 1] - top2[:, 0])
    # a flip is only acceptable as an fp32 summation-order tie: margin at round-off level of the logits
    assert (margin < 1e-5 * np.abs(z[bad]).max(axis=1)).all(), \
        f'{tag}: {len(bad)} prediction mismatches, margins {margin[:8]}'


@pytest.mark.parametrize('optimizer,lr', [('sgd', 1e-2), ('adam', 1e-3)])
def test_bf16_batch4096_against_emulating_oracle(optimizer, lr):
    dims, p, B, steps = [784, 256, 10], 0.1, 4096, 3
    x, y = synthetic(B * steps + B)
    x = x.bfloat16().float()
    init = seeded_init(dims, bf16=True)
    model, net = build_model(dims, p, 'relu', optimizer, lr, init, 'bf16', max_batch=B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, opt_kind=OPT[optimizer], lr=lr, emulate_bf16=True)
    g = torch.Generator().manual_seed(42)
    for s in range(steps):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        mask = (torch.rand(B, dims[1], generator=g) >= p).to(torch.uint8)
        # evaluate first (pure), so mismatching rows can be examined under the pre-step weights
        pr, l = model.fit(xb.cuda(), yb.cuda(), masks=[mask.cuda()])
        opr, ol = oracle.fit(xb.numpy(), yb.numpy(), [mask.numpy()])
        assert (pr.cpu().numpy() != opr).mean() <= 1e-3, f'step {s}: too many prediction mismatches'
        assert l.item() == pytest.approx(float(ol), rel=1e-2)
    xb, yb = x[steps * B:], y[steps * B:]
    pr, l = model.evaluate(xb.cuda(), yb.cuda())
    opr, ol = oracle.evaluate(xb.numpy(), yb.numpy())
    assert (pr.cpu().numpy() != opr).mean() <= 2e-3
    assert l.item() == pytest.approx(float(ol), rel=1e-2)
    assert_params_close(net_params(net), oracle.parameters(), 1e-2, lr=lr, steps=steps, adam=optimizer == 'adam')


def test_bf16_first_step_predictions_bit_exact():
    """Before any update both sides hold identical bf16-representable operands: argmax must agree exactly,
    modulo fp32 summation-order ties, which are reported with their margins."""
    dims, B = [784, 256, 10], 8192
    x, y = synthetic(B, seed=77)
    x = x.bfloat16().float()
    init = seeded_init(dims, bf16=True)
    model, _ = build_model(dims, 0.0, 'relu', 'sgd', 0.0, init, 'bf16', max_batch=B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_SGD, lr=0.0, emulate_bf16=True)
    pr, l = model.evaluate(x.cuda(), y.cuda())
    opr, ol = oracle.evaluate(x.numpy(), y.numpy())
    check_predictions(pr.cpu().numpy(), opr, oracle, x.numpy(), 'evaluate')
    assert l.item() == pytest.approx(float(ol), rel=1e-4)


def test_bf16_golden_within_1e2_of_reference_run():
    """The reference's own fp32 run (golden) vs the bf16 engine: loss within 1e-2."""
    case = load_case('tinysys_adam_b32')
    model, net = build_model(case.dims, case.dropout, case.activation, case.optimizer, case.lr, case.init_params(), 'bf16',
                             max_batch=case.batch)
    losses, preds, _ = replay(case, EngineAdapter(model), onp.OracleMetrics())
    np.testing.assert_allclose(losses, case.raw['batch_loss'], rtol=1e-2)
    assert (preds == case.raw['batch_pred']).mean() > 0.97


def test_bf16_deep_net_and_bf16_inputs():
    dims, p, B = [784, 128, 64, 10], 0.1, 1024
    x, y = synthetic(B, seed=8)
    xb16 = x.flatten(1).bfloat16()
    init = seeded_init(dims, bf16=True)
    model, net = build_model(dims, p, 'relu', 'adam', 1e-3, init, 'bf16', max_batch=B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, opt_kind=onp.OPT_ADAM, lr=1e-3, emulate_bf16=True)
    g = torch.Generator().manual_seed(3)
    for s in range(2):
        masks = [(torch.rand(B, h, generator=g) >= p).to(torch.uint8) for h in dims[1:-1]]
        pr, l = model.fit(xb16.cuda(), y.cuda(), masks=[m.cuda() for m in masks])      # bf16 inputs straight from the loader
        opr, ol = oracle.fit(xb16.float().numpy(), y.numpy(), [m.numpy() for m in masks])
        assert (pr.cpu().numpy() != opr).mean() <= 2e-3
        assert l.item() == pytest.approx(float(ol), rel=1e-2)
    assert_params_close(net_params(net), oracle.parameters(), 1e-2, lr=1e-3, steps=2)


def test_bf16_wide_hidden_layers_config4_shape_class():
    """North-star config 4 shape class (wide hidden layers, generalised MLP) at a size the oracle finishes in seconds:
    784 -> 1024 -> 1024 -> 10 exercises the 256-wide tcgen05 tiles, the MN-major dX / dW contractions without split-K
    and the layer-by-layer path (the fused kernel only takes one hidden layer)."""
    dims, p, B = [784, 1024, 1024, 10], 0.1, 1024
    x, y = synthetic(B, seed=21)
    x = x.bfloat16().float()
    init = seeded_init(dims, bf16=True)
    model, net = build_model(dims, p, 'relu', 'sgd', 0.05, init, 'bf16', max_batch=B)
    assert not model.engine.spec.dims == [784, 256, 10]
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, opt_kind=onp.OPT_SGD, lr=0.05, emulate_bf16=True)
    g = torch.Generator().manual_seed(5)
    for s in range(2):
        masks = [(torch.rand(B, h, generator=g) >= p).to(torch.uint8) for h in dims[1:-1]]
        pr, l = model.fit(x.cuda(), y.cuda(), masks=[m.cuda() for m in masks])
        opr, ol = oracle.fit(x.numpy(), y.numpy(), [m.numpy() for m in masks])
        assert (pr.cpu().numpy() != opr).mean() <= 2e-3
        assert l.item() == pytest.approx(float(ol), rel=1e-2)
    assert_params_close(net_params(net), oracle.parameters(), 1e-2, lr=0.05, steps=2, adam=False)


@pytest.mark.parametrize('B', [4096, 12000])      # a slab is a 4-CTA cluster / a single CTA (12000 rows = 94 slabs, ragged)
def test_bf16_inkernel_cast_matches_cast_kernel(monkeypatch, B):
    """fp32 batches converted inside the fused kernel (TSB200_INKERNEL_CAST=1 when the handle is created) must give the
    same bits as the cast kernel in front of it: predictions, loss records and the weights after three steps."""
    dims, p, steps = [784, 256, 10], 0.1, 3
    x, y = synthetic(B * steps + 1000, seed=5)
    init = seeded_init(dims, bf16=True)
    outs = []
    for flag in ('0', '1'):
        monkeypatch.setenv('TSB200_INKERNEL_CAST', flag)
        model, net = build_model(dims, p, 'relu', 'sgd', 1e-2, init, 'bf16', max_batch=B)
        g = torch.Generator().manual_seed(42)
        got = []
        for s in range(steps):
            mask = (torch.rand(B, dims[1], generator=g) >= p).to(torch.uint8)
            pr, l = model.fit(x[s * B:(s + 1) * B].cuda(), y[s * B:(s + 1) * B].cuda(), masks=[mask.cuda()])
            got += [pr.cpu(), l.cpu()]
        pr, l = model.evaluate(x[steps * B:].cuda(), y[steps * B:].cuda())     # ragged last slab: 1000 rows
        got += [pr.cpu(), l.cpu()] + [q.detach().cpu().clone() for q in net.parameters()]
        outs.append(got)
    for a, b in zip(*outs):
        assert torch.equal(a, b)


@pytest.mark.parametrize('precision,tol', [('bf16', 1e-2), ('f32', 1e-5)])
def test_many_row_blocks_two_level_partial_sum(precision, tol):
    """20,000 rows = 157 row blocks: more than the optimizer sums directly, so the last-layer / bias gradient partials
    go through the grouped pre-reduction first (and, in bf16 mode, a slab is one CTA with on-chip input conversion)."""
    dims, p, B, steps = [784, 256, 10], 0.1, 20000, 2
    x, y = synthetic(B * steps)
    bf16 = precision == 'bf16'
    if bf16:
        x = x.bfloat16().float()
    init = seeded_init(dims, bf16=bf16)
    model, net = build_model(dims, p, 'relu', 'sgd', 1e-2, init, precision, max_batch=B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, opt_kind=OPT['sgd'], lr=1e-2, emulate_bf16=bf16)
    g = torch.Generator().manual_seed(42)
    for s in range(steps):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        mask = (torch.rand(B, dims[1], generator=g) >= p).to(torch.uint8)
        pr, l = model.fit(xb.cuda(), yb.cuda(), masks=[mask.cuda()])
        opr, ol = oracle.fit(xb.numpy(), yb.numpy(), [mask.numpy()])
        assert (pr.cpu().numpy() != opr).mean() <= (1e-3 if bf16 else 0.0)
        assert l.item() == pytest.approx(float(ol), rel=tol)
    assert_params_close(net_params(net), oracle.parameters(), tol, lr=1e-2, steps=steps, adam=False)


@pytest.mark.parametrize('dims', [[784, 512, 10], [784, 256, 128, 10]])
def test_bf16_gelu_against_emulating_oracle(dims):
    """GELU in the tensor-core mode: layer-by-layer tcgen05 path with the pre-activations kept in bf16 for gelu'
    (the oracle rounds them at the same point), dropout masks supplied; loss and weights within 1e-2."""
    p, B, steps = 0.1, 2048, 3
    x, y = synthetic(B * steps + B)
    x = x.bfloat16().float()
    init = seeded_init(dims, bf16=True)
    model, net = build_model(dims, p, 'gelu', 'sgd', 1e-2, init, 'bf16', max_batch=B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, activation=onp.ACT_GELU, opt_kind=OPT['sgd'], lr=1e-2,
                           emulate_bf16=True)
    g = torch.Generator().manual_seed(42)
    for s in range(steps):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        masks = [(torch.rand(B, h, generator=g) >= p).to(torch.uint8) for h in dims[1:-1]]
        pr, l = model.fit(xb.cuda(), yb.cuda(), masks=[k.cuda() for k in masks])
        opr, ol = oracle.fit(xb.numpy(), yb.numpy(), [k.numpy() for k in masks])
        assert (pr.cpu().numpy() != opr).mean() <= 2e-3
        assert l.item() == pytest.approx(float(ol), rel=1e-2)
    pr, l = model.evaluate(x[steps * B:].cuda(), y[steps * B:].cuda())
    opr, ol = oracle.evaluate(x[steps * B:].numpy(), y[steps * B:].numpy())
    assert (pr.cpu().numpy() != opr).mean() <= 2e-3
    assert l.item() == pytest.approx(float(ol), rel=1e-2)
    assert_params_close(net_params(net), oracle.parameters(), 1e-2, lr=1e-2, steps=steps, adam=False)
