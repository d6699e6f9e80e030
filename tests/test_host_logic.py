"""CPU tests of the host side: network inspection, identity hashes, the stand-in orchestration layer against
the real TorchSystem (when /root/reference is present), handler call sequences, sharding."""
import os
import subprocess
import sys
import textwrap
from unittest.mock import MagicMock

import pytest
import torch
from torch import nn

import __graft_entry__ as entry

entry.build()

from torchsystem_b200 import MLP, training                      # noqa: E402
from torchsystem_b200 import compat                             # noqa: E402
from torchsystem_b200.modules import inspect_mlp                # noqa: E402
from torchsystem_b200.parallel import shard_bounds              # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = '/root/reference'
have_reference = os.path.isdir(os.path.join(REFERENCE, 'torchsystem'))


def test_inspect_reference_shaped_mlp():
    spec = inspect_mlp(MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU()))
    assert spec.dims == [784, 256, 10] and spec.dropout == 0.1 and spec.activation == 'relu'
    spec = inspect_mlp(MLP(784, [4096, 4096, 4096, 4096], 10, dropout=0.0, activation=nn.ReLU()))
    assert spec.dims == [784, 4096, 4096, 4096, 4096, 10]
    seq = nn.Sequential(nn.Linear(8, 16), nn.Dropout(0.2), nn.GELU(), nn.Linear(16, 4))
    spec = inspect_mlp(seq)
    assert spec.dims == [8, 16, 4] and spec.activation == 'gelu' and spec.dropout == 0.2


@pytest.mark.parametrize('bad', [
    nn.Sequential(nn.Linear(8, 16), nn.ReLU(), nn.Dropout(0.2), nn.Linear(16, 4)),     # dropout after activation
    nn.Sequential(nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 4)),
    nn.Sequential(nn.Linear(8, 4)),
    nn.Sequential(nn.Linear(8, 16), nn.ReLU(), nn.Linear(8, 4)),
    nn.Sequential(nn.Linear(8, 16), nn.GELU(approximate='tanh'), nn.Linear(16, 4)),
])
def test_inspect_rejects_what_the_engine_cannot_run(bad):
    with pytest.raises(TypeError):
        inspect_mlp(bad)


def test_mlp_matches_reference_layout_and_state_dict_names():
    net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
    assert list(net.state_dict()) == ['input_layer.weight', 'input_layer.bias', 'output_layer.weight', 'output_layer.bias']
    net.eval()
    x = torch.randn(4, 784)
    want = net.output_layer(torch.relu(net.input_layer(x)))
    assert torch.equal(net(x), want)


def test_identity_hash_goldens():
    """examples/tinysys/notebook.ipynb:34 and docs/registry.md:50 (reproduced with the stand-in registry)."""
    class _MLP(MLP):
        pass
    _MLP.__name__ = 'MLP'
    compat.register(_MLP)
    relu = type('ReLU', (nn.ReLU,), {})
    compat.register(relu)
    assert compat.gethash(_MLP(784, 256, 10, dropout=0.1, activation=relu())) == '302adcec37ba79ed6a33d58956391d20'
    assert compat.gethash(_MLP(784, 256, 10, dropout=0.5, activation=relu())) == 'af51a51a38f7ad81f9523360fafe7068'


def test_named_registration_matches_the_reference_registry():
    """`@register('name')` (torchsystem/registry/accessors.py:150-174): the name replaces the class name in gethash. The
    stand-in must produce the id the reference's own registry produces (compared directly when /root/reference is here)."""
    import importlib.util, os, sys
    def build(register):
        @register('encoder-v2')
        class Net:
            def __init__(self, width: int, depth: int = 2):
                pass

        @register
        class Plain:
            def __init__(self, width: int):
                pass
        return Net(64, depth=3), Plain(8)
    from hashlib import md5
    from json import dumps
    named, plain = build(compat.register)
    assert compat.gethash(named) == md5(('encoder-v2' + dumps({'width': 64, 'depth': 3})).encode()).hexdigest()
    assert compat.gethash(plain) == md5(('Plain' + dumps({'width': 8})).encode()).hexdigest()
    if os.path.isdir('/root/reference/torchsystem'):
        sys.path.insert(0, '/root/reference')
        try:
            from torchsystem.registry import register as ref_register, gethash as ref_gethash
            r_named, r_plain = build(ref_register)
            assert ref_gethash(r_named) == compat.gethash(named) and ref_gethash(r_plain) == compat.gethash(plain)
        finally:
            sys.path.remove('/root/reference')


def test_shard_bounds_equal_and_disjoint():
    spans = [shard_bounds(1_048_576 + 123, r, 8, 4096) for r in range(8)]
    assert all(hi - lo == spans[0][1] - spans[0][0] for lo, hi in spans)
    assert all((hi - lo) % 4096 == 0 for lo, hi in spans)
    assert all(spans[i][1] == spans[i + 1][0] for i in range(7))


# ------------------------------------------------------------------ stand-in orchestration semantics
@pytest.mark.skipif(compat.HAVE_TORCHSYSTEM, reason='real torchsystem on the path: stand-ins not in use')
def test_standin_depends_service_consumer():
    opened = []

    def dep():
        opened.append('open'); yield 'resource'; opened.append('close')

    def other():
        return 'override'

    provider = compat.Provider()
    service = compat.Service(provider=provider)

    @service.handler
    def do_thing(a, b=compat.Depends(dep)):
        return a, b

    assert service.handle('do-thing', 1) == (1, 'resource') and opened == ['open', 'close']
    provider.override(dep, other)
    assert do_thing(2) == (2, 'override')
    assert do_thing(2, 'explicit') == (2, 'explicit')
    with pytest.raises(KeyError):
        service.handle('missing')

    @compat.event
    class ModelTrained:
        value: int

    @compat.event
    class ModelValidated:
        value: int

    seen = []
    consumer = compat.Consumer()

    @consumer.handler
    def on_any(e: ModelTrained | ModelValidated):
        seen.append(e.value)

    producer = compat.Producer()
    producer.register(consumer)
    producer.dispatch(ModelTrained(1)); producer.dispatch(ModelValidated(2)); producer.dispatch('ignored')
    assert seen == [1, 2] and 'model-trained' in consumer.handlers


@pytest.mark.skipif(compat.HAVE_TORCHSYSTEM, reason='real torchsystem on the path')
def test_standin_aggregate_hooks():
    calls = []

    class A(compat.Aggregate):
        def __init__(self):
            super().__init__()
            self.epoch = 0

        def onepoch(self): calls.append(('epoch', self.epoch))
        def onphase(self): calls.append(('phase', self.phase))

    a = A()
    a.phase = 'train'; a.phase = 'evaluation'; a.epoch += 1
    assert calls == [('phase', 'train'), ('phase', 'evaluation'), ('epoch', 1)] and not a.training


def _trace_handlers(mod, n_batches=3):
    """Run iterate() of a handlers module over fakes; return the sequence of observable calls."""
    log = []
    model = MagicMock()
    model.epoch = 0
    model.fit.side_effect = lambda i, t: (log.append(('fit', i, t)) or ('p', 'l'))
    model.evaluate.side_effect = lambda i, t: (log.append(('evaluate', i, t)) or ('p', 'l'))
    metrics = MagicMock()
    metrics.update.side_effect = lambda b, l, p, t: log.append(('update', b, l, p, t))
    metrics.compute.side_effect = lambda: (log.append(('compute',)) or {'loss': 0, 'accuracy': 1})
    metrics.reset.side_effect = lambda: log.append(('reset',))

    class T:
        def __init__(self, name): self.name = name
        def to(self, device): log.append(('to', self.name, device)); return self.name

    loader = [(T(f'x{i}'), T(f'y{i}')) for i in range(n_batches)]
    consumer = (compat.Consumer if mod is training else __import__('torchsystem.services', fromlist=['Consumer']).Consumer)()

    def on_trained(e: mod.Trained): log.append(('Trained', e.results['accuracy']))
    def on_validated(e: mod.Validated): log.append(('Validated', e.results['accuracy']))
    def on_iterated(e: mod.Iterated): log.append(('Iterated', e.model.epoch))
    for h in (on_trained, on_validated, on_iterated):
        consumer.handler(h)
    mod.producer.consumers.clear()
    mod.producer.register(consumer)
    mod.provider.override(mod.device, lambda: 'cuda:7')
    mod.service.handle('iterate', model, [('train', loader), ('validation', loader)], metrics)
    phases = [c.args[0] for c in type(model).phase.mock_calls] if hasattr(type(model), 'phase') else None
    return log, model.epoch, phases


def test_handlers_trace_without_reference():
    log, epoch, _ = _trace_handlers(training)
    kinds = [e[0] for e in log]
    assert kinds.count('fit') == 3 and kinds.count('evaluate') == 3 and kinds.count('update') == 6
    assert kinds[-1] == 'Iterated' and epoch == 1
    assert log[0] == ('to', 'x0', 'cuda:7') and ('update', 1, 'l', 'p', 'y0') in log
    assert kinds.index('Trained') < kinds.index('evaluate') and kinds.count('reset') == 2


@pytest.mark.skipif(not have_reference, reason='needs /root/reference (build container only)')
def test_handlers_trace_equals_unmodified_reference_handlers():
    """Same observable call sequence as /root/reference/examples/tinysys/tinysys/services/training.py."""
    code = textwrap.dedent(f'''
        import sys, json
        sys.path[:0] = [{ROOT!r}, {REFERENCE!r}, {REFERENCE + '/examples/tinysys'!r}]
        from tests.test_host_logic import _trace_handlers
        from torchsystem_b200 import compat, training as mine
        assert compat.HAVE_TORCHSYSTEM
        from tinysys.services import training as theirs
        a = _trace_handlers(theirs)
        b = _trace_handlers(mine)
        assert a[0] == b[0] and a[1] == b[1], (a, b)
        # B200Classifier plugs into the real Aggregate
        import torchsystem
        from torchsystem_b200 import B200Classifier
        assert issubclass(B200Classifier, torchsystem.Aggregate)
        print('same', len(a[0]))
    ''')
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 0 and out.stdout.startswith('same'), out.stderr[-2000:]


@pytest.mark.skipif(not have_reference, reason='needs /root/reference (build container only)')
def test_reference_suite_still_passes_with_this_package_imported(tmp_path):
    """The reference's own unit tests (DI, services, events, aggregate, registry) with torchsystem_b200 loaded
    first: the package must not disturb the orchestration layer. test_compiler.py fails on torch >= 2.6 with or
    without this package and is excluded."""
    conftest = tmp_path / 'conftest.py'
    conftest.write_text(f'import sys; sys.path[:0] = [{ROOT!r}, {REFERENCE!r}]\nimport torchsystem_b200\n')
    import shutil
    shutil.copytree(os.path.join(REFERENCE, 'tests'), tmp_path / 'tests', ignore=shutil.ignore_patterns('conftest.py', '__pycache__'))
    out = subprocess.run([sys.executable, '-m', 'pytest', '-q', '-p', 'no:cacheprovider', '--deselect',
                          'tests/test_compiler.py', str(tmp_path / 'tests'), '-c', str(conftest)],
                         capture_output=True, text=True, cwd=tmp_path)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
    assert '17 passed' in out.stdout, out.stdout[-500:]


def test_device_loader_batches_shuffle_and_shards():
    """Loader protocol (tinysys/domain.py:26-29) of DeviceLoader, exercised on the CPU device: ragged last batch,
    drop_last, a seeded permutation that keeps inputs and targets paired, equal disjoint rank shards."""
    from torchsystem_b200 import DeviceLoader
    x = torch.arange(103, dtype=torch.float32).reshape(103, 1).repeat(1, 4)
    y = torch.arange(103)
    sizes = [xb.shape[0] for xb, _ in DeviceLoader(x, y, 32, device='cpu')]
    assert sizes == [32, 32, 32, 7] and len(DeviceLoader(x, y, 32, device='cpu')) == 4
    assert [xb.shape[0] for xb, _ in DeviceLoader(x, y, 32, device='cpu', drop_last=True)] == [32, 32, 32]
    # unshuffled: the batch views are made once and handed out again every epoch (same storage, same order, still views)
    plain = DeviceLoader(x.clone(), y, 32, device='cpu')
    e1, e2 = list(plain), list(plain)
    assert all(a is b and c is d for (a, c), (b, d) in zip(e1, e2)) and len(e1) == 4
    assert e1[1][0].data_ptr() == plain.inputs[32:64].data_ptr() and torch.equal(e1[3][1], plain.targets[96:])
    plain.inputs[40] = 7.0                                   # a view: sees later writes to the dataset
    assert float(e1[1][0][8].flatten()[0]) == 7.0
    a = [(xb, yb) for xb, yb in DeviceLoader(x, y, 32, device='cpu', shuffle=True, seed=7)]
    b = [(xb, yb) for xb, yb in DeviceLoader(x, y, 32, device='cpu', shuffle=True, seed=7)]
    assert all(torch.equal(p[0], q[0]) and torch.equal(p[1], q[1]) for p, q in zip(a, b))
    allx = torch.cat([p[0] for p in a]); ally = torch.cat([p[1] for p in a])
    assert torch.equal(allx[:, 0].long(), ally) and sorted(ally.tolist()) == list(range(103)) and not torch.equal(ally, y)
    shards = [DeviceLoader(x, y, 8, device='cpu', rank=r, world=4) for r in range(4)]
    seen = [torch.cat([yb for _, yb in s]) for s in shards]
    assert all(len(s) == 24 for s in seen) and len(set(torch.cat(seen).tolist())) == 96
    xb, yb = next(iter(DeviceLoader(x, y, 8, device='cpu', dtype=torch.bfloat16)))
    assert xb.dtype == torch.bfloat16 and yb.dtype == torch.int64 and xb.to('cpu') is xb      # .to(device) is a no-op


def test_layer_buckets_tile_the_flat_gradient():
    """All-reduce buckets are whole layers, tile [0, total) without gaps, and merge small layers (config 4's model:
    784 -> 4096 x 4 -> 10 gives three buckets of 16-20 M elements; the shipped model is one bucket)."""
    import math
    from torchsystem_b200.parallel import layer_buckets

    def layout(dims):
        offs, o = [], 0
        for l in range(len(dims) - 1):
            offs.append(o); o += math.ceil(dims[l] * dims[l + 1] / 1024) * 1024
            offs.append(o); o += math.ceil(dims[l + 1] / 1024) * 1024
        return offs, o

    offs, total = layout([784, 4096, 4096, 4096, 4096, 10])
    b = layer_buckets(offs, total)
    assert [x[:2] for x in b] == [(0, 4), (4, 2), (6, 4)]
    assert b[0][2] == 0 and all(b[i][2] + b[i][3] == b[i + 1][2] for i in range(len(b) - 1)) and b[-1][2] + b[-1][3] == total
    assert all(x[3] >= 16 << 20 for x in b)
    offs, total = layout([784, 256, 10])
    assert layer_buckets(offs, total) == [(0, 4, 0, total)]


def test_optimizer_config_maps_torch_optimizers():
    """torch.optim.SGD / Adam / AdamW hyper-parameters as the fused optimizer takes them; anything else is refused."""
    import torch
    from torchsystem_b200.classifier import _optimizer_config
    w = [torch.nn.Parameter(torch.zeros(3))]
    kind, h = _optimizer_config(torch.optim.SGD(w, lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-4))
    assert kind == 'sgd' and h == dict(lr=0.1, weight_decay=1e-4, momentum=0.9, dampening=0.0, nesterov=True)
    kind, h = _optimizer_config(torch.optim.Adam(w, lr=1e-3, amsgrad=True))
    assert kind == 'adam' and h['amsgrad'] and not h['decoupled_weight_decay'] and h['betas'] == (0.9, 0.999)
    kind, h = _optimizer_config(torch.optim.AdamW(w, lr=1e-3, weight_decay=0.05))
    assert kind == 'adam' and h['decoupled_weight_decay'] and h['weight_decay'] == 0.05
    with pytest.raises(NotImplementedError):
        _optimizer_config(torch.optim.SGD(w, lr=0.1, maximize=True))
    with pytest.raises(TypeError):
        _optimizer_config(torch.optim.RMSprop(w, lr=0.1))
