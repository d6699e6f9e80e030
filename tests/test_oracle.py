"""Pin the CPU oracle against outputs of the unmodified reference (tests/golden/*.npz).

The reference's own tests hold no vector for fit/evaluate/Metrics, so the
golden files were produced by running the reference itself (tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import mlp_numpy as onp
from oracle import ref_torch
from tests.golden_util import CASES, load_case, replay, golden_phase_results, assert_params_close

ACT = {'relu': onp.ACT_RELU, 'gelu': onp.ACT_GELU}
OPT = {'adam': onp.OPT_ADAM, 'sgd': onp.OPT_SGD}


def make_oracle(case, emulate_bf16=False):
    params = case.init_params()
    return onp.OracleMLP(weights=params[0::2], biases=params[1::2], dropout=case.dropout,
                         activation=ACT[case.activation], opt_kind=OPT[case.optimizer], lr=case.lr,
                         emulate_bf16=emulate_bf16)


@pytest.mark.parametrize('name', CASES)
def test_numpy_oracle_matches_reference_run(name):
    case = load_case(name)
    engine = make_oracle(case)
    losses, preds, phases = replay(case, engine, onp.OracleMetrics())
    # predictions bit-exact, losses and weights to fp32 round-off (tolerance 1e-5 relative)
    assert np.array_equal(preds, case.raw['batch_pred'])
    np.testing.assert_allclose(losses, case.raw['batch_loss'], rtol=1e-5, atol=1e-6)
    steps = case.epochs * len(case.batches('train'))
    assert_params_close(engine.parameters(), case.final_params(), 1e-5, lr=case.lr, steps=steps,
                        adam=case.optimizer == 'adam')
    for got, (loss, acc) in zip(phases, golden_phase_results(case)):
        assert got['accuracy'] == pytest.approx(acc, abs=1e-7)   # counts are exact; golden stored as f32
        assert got['loss'] == pytest.approx(loss, rel=1e-5)
    if case.optimizer == 'adam':
        assert engine.step == int(case.raw['opt_step'])
        assert_params_close([engine.exp_avg[0], engine.exp_avg_sq[0]],
                            [case.raw['exp_avg_0'], case.raw['exp_avg_sq_0']], 1e-3, adam=False, what='adam state')


@pytest.mark.parametrize('name', CASES)
def test_torch_restatement_matches_reference_run(name):
    """oracle/ref_torch.py draws its own dropout masks from torch's CPU generator, like the
    reference; seeded identically it must reproduce the recorded run."""
    case = load_case(name)
    torch.manual_seed(0)
    net = ref_torch.make_layers(case.dims, case.dropout, case.activation)
    with torch.no_grad():
        for p, init in zip(net.parameters(), case.init_params()):
            p.copy_(torch.from_numpy(init))
    opt = (torch.optim.Adam(net.parameters(), lr=case.lr) if case.optimizer == 'adam'
           else torch.optim.SGD(net.parameters(), lr=case.lr))
    model, metrics = ref_torch.CpuClassifier(net, opt), ref_torch.CpuMetrics()

    def loader(split):
        return [(torch.from_numpy(x), torch.from_numpy(y)) for x, y in case.batches(split)]

    torch.manual_seed(42)
    got = []
    for _ in range(case.epochs):
        r = ref_torch.train(model, loader('train'), metrics); metrics.reset()
        got.append((r['loss'].item(), r['accuracy'].item()))
        r = ref_torch.validate(model, loader('val'), metrics); metrics.reset()
        got.append((r['loss'].item(), r['accuracy'].item()))
    for (gl, ga), (wl, wa) in zip(got, golden_phase_results(case)):
        assert ga == pytest.approx(wa, abs=1e-7)
        assert gl == pytest.approx(wl, rel=1e-6)
    for p, want in zip(net.parameters(), case.final_params()):
        np.testing.assert_allclose(p.detach().numpy(), want, rtol=1e-6, atol=1e-7)


def test_mask_predraw_matches_dropout():
    """The recorded masks are what nn.Dropout drew: relu(dropout(h)) == mask * relu(h) / (1 - p)."""
    torch.manual_seed(7)
    h = torch.randn(16, 64)
    mask = ref_torch.draw_dropout_mask((16, 64), 0.1)
    out = torch.nn.functional.dropout(h, 0.1, training=True)
    assert torch.equal(out, h * (mask.float() / 0.9))


def test_bf16_emulation_stays_within_tolerance():
    """bf16 operand rounding (the tensor-core mode) must keep loss within 1e-2 of the fp32 reference run."""
    case = load_case('tinysys_adam_b32')
    from tests.parity_util import LockstepF32
    engine = make_oracle(case, emulate_bf16=True)
    walker = LockstepF32(engine, engine.parameters, make_oracle(case))
    losses, preds, _ = replay(case, walker, onp.OracleMetrics())
    np.testing.assert_allclose(losses, case.raw['batch_loss'], rtol=1e-2)
    # exact or explained, never a rate: a row may leave the fp32 reference's prediction only where its fp32 top-2 margin
    # is below the bf16 rounding noise of its logits (tests/parity_util.py)
    assert (preds != case.raw['batch_pred']).sum() == len(walker.flips)
    print(f'{len(walker.flips)} of {len(preds)} rows flipped; (margin, bound): {walker.flips[:5]}')


def test_bf16_round_ties_to_even():
    vals = np.array([1.0, 1.00390625, 1.005859375, 1.01171875, -3.14159, 0.0], dtype=np.float32)
    want = torch.from_numpy(vals).bfloat16().float().numpy()
    assert np.array_equal(onp.bf16_round(vals), want)


def test_metrics_semantics_mean_of_batch_means():
    m = onp.OracleMetrics()
    m.update(1, np.float32(1.0), np.array([1, 2, 3]), np.array([1, 2, 0]))
    m.update(2, np.float32(3.0), np.array([5]), np.array([5]))      # short batch weighs the same
    r = m.compute()
    assert r['loss'] == 2.0 and r['accuracy'] == 0.75
    m.reset()
    assert m.compute()['loss'] == 0.0


OPT_VARIANTS = {
    'sgd_wd': ('sgd', dict(lr=0.05, weight_decay=1e-2)),
    'sgd_momentum': ('sgd', dict(lr=0.05, momentum=0.9)),
    'sgd_momentum_dampening_wd': ('sgd', dict(lr=0.05, momentum=0.8, dampening=0.3, weight_decay=1e-3)),
    'sgd_nesterov': ('sgd', dict(lr=0.05, momentum=0.9, nesterov=True)),
    'adam_wd': ('adam', dict(lr=1e-3, weight_decay=1e-2)),
    'adam_amsgrad': ('adam', dict(lr=1e-3, amsgrad=True)),
    'adamw': ('adamw', dict(lr=1e-3, weight_decay=1e-1)),
    'adamw_amsgrad': ('adamw', dict(lr=1e-3, weight_decay=1e-1, amsgrad=True, betas=(0.8, 0.99))),
}


def torch_optimizer(kind, params, hyper):
    return {'sgd': torch.optim.SGD, 'adam': torch.optim.Adam, 'adamw': torch.optim.AdamW}[kind](params, **hyper)


def oracle_hyper(kind, hyper):
    h = dict(hyper)
    h['opt_kind'] = onp.OPT_SGD if kind == 'sgd' else onp.OPT_ADAM
    if kind == 'adamw':
        h['decoupled_weight_decay'] = True
    return h


@pytest.mark.parametrize('name', sorted(OPT_VARIANTS))
def test_oracle_optimizer_variants_match_torch_optim(name):
    """Pins the restated SGD (weight decay, momentum, dampening, Nesterov) and Adam / AdamW (weight decay, amsgrad)
    updates to torch.optim's single-tensor CPU path on the same gradient sequence (5 steps)."""
    kind, hyper = OPT_VARIANTS[name]
    g = torch.Generator().manual_seed(7)
    shapes = [(16, 24), (16,), (10, 16), (10,)]
    params = [torch.nn.Parameter(torch.randn(s, generator=g)) for s in shapes]
    opt = torch_optimizer(kind, params, dict(hyper, foreach=False))
    oracle = onp.OracleMLP(weights=[params[0].detach().numpy(), params[2].detach().numpy()],
                           biases=[params[1].detach().numpy(), params[3].detach().numpy()], **oracle_hyper(kind, hyper))
    for _ in range(5):
        grads = [torch.randn(s, generator=g) * 0.1 for s in shapes]
        for p, gr in zip(params, grads):
            p.grad = gr.clone()
        opt.step()
        oracle._optimizer_step([gr.numpy().copy() for gr in grads])
        for p, q in zip(params, oracle.parameters()):
            np.testing.assert_allclose(q, p.detach().numpy(), rtol=2e-6, atol=1e-7)
