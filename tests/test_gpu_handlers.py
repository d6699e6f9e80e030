"""GPU test of the whole drop-in: the train / validate / iterate handlers (same flow as the reference's
training.py) driving B200Classifier + B200Metrics + the loaders, checked against the CPU oracle run through the
oracle's own handler loop."""
import numpy as np
import pytest
import torch
from torch import nn

from oracle import mlp_numpy as onp
from tests.test_gpu_parity import build_model, seeded_init, synthetic, net_params
from tests.golden_util import assert_params_close

pytestmark = pytest.mark.gpu


def oracle_epoch(oracle, xt, yt, xv, yv, B):
    m = onp.OracleMetrics()
    out = []
    for b, i in enumerate(range(0, len(xt), B), 1):
        p, l = oracle.fit(xt[i:i + B], yt[i:i + B], None)
        m.update(b, l, p, yt[i:i + B])
    out.append(m.compute()); m.reset()
    for b, i in enumerate(range(0, len(xv), B), 1):
        p, l = oracle.evaluate(xv[i:i + B], yv[i:i + B])
        m.update(b, l, p, yv[i:i + B])
    out.append(m.compute()); m.reset()
    return out


@pytest.mark.parametrize('loader_kind', ['device', 'pinned'])
def test_iterate_handler_end_to_end(loader_kind):
    from torchsystem_b200 import B200Metrics, DeviceLoader, PinnedLoader, training
    from torchsystem_b200.compat import Consumer
    dims, B = [784, 256, 10], 512
    x, y = synthetic(B * 5 + 100)                       # ragged last validation batch
    xt, yt, xv, yv = x[:B * 3], y[:B * 3], x[B * 3:], y[B * 3:]
    init = seeded_init(dims)
    model, net = build_model(dims, 0.0, 'relu', 'adam', 1e-3, init, 'f32', max_batch=B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_ADAM, lr=1e-3)
    metrics = B200Metrics(device='cuda:0', log_every=2)
    mk = (lambda a, b: DeviceLoader(a, b, B, device='cuda:0')) if loader_kind == 'device' else \
         (lambda a, b: PinnedLoader(a, b, B, device='cuda:0', prefetch=2))
    loaders = [('train', mk(xt, yt)), ('validation', mk(xv, yv))]

    events = []
    consumer = Consumer()

    @consumer.handler
    def on_phase(e: training.Trained | training.Validated):
        events.append((type(e).__name__, e.model.epoch, e.results['loss'].item(), e.results['accuracy'].item(), e.model.phase))

    @consumer.handler
    def on_iterated(e: training.Iterated):
        events.append(('Iterated', e.model.epoch))

    training.producer.consumers.clear()
    training.producer.register(consumer)
    training.provider.override(training.device, lambda: 'cuda:0')

    want = []
    for epoch in range(2):
        training.service.handle('iterate', model, loaders, metrics)
        want += oracle_epoch(oracle, xt.numpy(), yt.numpy(), xv.numpy(), yv.numpy(), B)
    assert model.epoch == 2
    names = [e[0] for e in events]
    assert names == ['Trained', 'Validated', 'Iterated'] * 2
    assert events[0][4] == 'train' and events[1][4] == 'evaluation' and events[2] == ('Iterated', 1)
    got = [e for e in events if e[0] != 'Iterated']
    for g, w in zip(got, want):
        assert g[3] == pytest.approx(w['accuracy'], abs=1e-7)       # correct counts are exact
        assert g[2] == pytest.approx(w['loss'], rel=1e-5)
    assert_params_close(net_params(net), oracle.parameters(), 1e-5, lr=1e-3, steps=6)
    # the network the consumers see is a live, picklable torch module (storage.py:84-86)
    import io, pickle
    buf = io.BytesIO(); torch.save(model.nn, buf)


def test_id_and_checkpoint_roundtrip():
    """Repository.store/restore contract (repository.py:13-17): save nn, load_state_dict into a fresh aggregate,
    phase change re-syncs -> same evaluation."""
    from torchsystem_b200 import compat
    dims, B = [784, 64, 10], 256
    x, y = synthetic(B)
    model, net = build_model(dims, 0.0, 'relu', 'sgd', 0.1, seeded_init(dims), 'f32', max_batch=B)
    model.phase = 'train'
    for _ in range(3):
        model.fit(x.cuda(), y.cuda())
    state = {k: v.clone() for k, v in model.nn.state_dict().items()}
    p1, l1 = model.evaluate(x.cuda(), y.cuda())
    other, net2 = build_model(dims, 0.0, 'relu', 'sgd', 0.1, seeded_init(dims, seed=5), 'f32', max_batch=B)
    other.nn.load_state_dict(state)
    other.phase = 'evaluation'
    p2, l2 = other.evaluate(x.cuda(), y.cuda())
    assert torch.equal(p1, p2) and l1.item() == l2.item()
    with pytest.raises(AttributeError):
        _ = model.id                     # not registered: same behaviour as gethash in the reference
    cls = type('MLP', (type(model.nn),), {})
    relu = type('ReLU', (nn.ReLU,), {})
    compat.register(cls); compat.register(relu)
    reg, _ = build_model([784, 256, 10], 0.1, 'relu', 'sgd', 0.1, seeded_init([784, 256, 10]), 'f32', max_batch=B)
    object.__setattr__(reg, 'nn', None)
    reg.nn = cls(784, 256, 10, dropout=0.1, activation=relu())
    assert reg.id == '302adcec37ba79ed6a33d58956391d20'        # examples/tinysys/notebook.ipynb:34


def test_main_equivalent_runs_end_to_end():
    """examples/main_b200.py is the runnable counterpart of the reference's examples/tinysys/main.py:49-71 (which cannot run on
    torch >= 2.6): two epochs at the shipped batch size 32; the events arrive in the reference's order with finite results, the
    id is the registry hash of the shipped aggregate (examples/tinysys/notebook.ipynb:34), accuracy improves, epoch counts up."""
    import importlib.util, os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'examples', 'main_b200.py')
    spec = importlib.util.spec_from_file_location('main_b200', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    classifier, history = mod.main(['--epochs', '2', '--train-samples', '8192', '--val-samples', '2048'])
    assert [h[0] for h in history] == ['train', 'validation'] * 2
    assert all(np.isfinite(h[2]) and 0.0 <= h[3] <= 1.0 for h in history)
    assert classifier.epoch == 2
    assert history[3][3] > history[1][3] > 0.15                      # validation accuracy: above chance and rising
    assert classifier.id == '302adcec37ba79ed6a33d58956391d20'
