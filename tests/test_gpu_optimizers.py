"""The fused optimizer's torch.optim variants against the oracle, fp32 mode, through B200Classifier
and the C ABI: SGD with weight decay / momentum / dampening / Nesterov, Adam with weight decay / amsgrad, AdamW.
The oracle's update rules are pinned to torch.optim in tests/test_oracle.py (CPU)."""
import numpy as np
import pytest
import torch
from torch import nn

from oracle import mlp_numpy as onp
from tests.golden_util import assert_params_close
from tests.test_gpu_parity import seeded_init, synthetic, net_params
from tests.test_oracle import OPT_VARIANTS, torch_optimizer, oracle_hyper

pytestmark = pytest.mark.gpu


def make(kind, hyper, dims, init, precision, B):
    from torchsystem_b200 import B200Classifier, MLP
    net = MLP(dims[0], dims[1], dims[-1], dropout=0.1, activation=nn.ReLU())
    with torch.no_grad():
        for p, w in zip([q for lin in net.linears() for q in (lin.weight, lin.bias)], init):
            p.copy_(torch.from_numpy(np.asarray(w)))
    opt = torch_optimizer(kind, net.parameters(), hyper)
    return B200Classifier(net, nn.CrossEntropyLoss(), opt, precision=precision, device='cuda:0', max_batch=B), net, opt


@pytest.mark.parametrize('name', sorted(OPT_VARIANTS))
def test_optimizer_variant_matches_oracle_f32(name):
    kind, hyper = OPT_VARIANTS[name]
    dims, B, steps = [784, 256, 10], 512, 4
    x, y = synthetic(B * steps)
    init = seeded_init(dims)
    model, net, opt = make(kind, hyper, dims, init, 'f32', B)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=0.1, **oracle_hyper(kind, hyper))
    g = torch.Generator().manual_seed(42)
    for s in range(steps):
        xb, yb = x[s * B:(s + 1) * B], y[s * B:(s + 1) * B]
        mask = (torch.rand(B, dims[1], generator=g) >= 0.1).to(torch.uint8)
        pr, l = model.fit(xb.cuda(), yb.cuda(), masks=[mask.cuda()])
        opr, ol = oracle.fit(xb.numpy(), yb.numpy(), [mask.numpy()])
        assert np.array_equal(pr.cpu().numpy(), opr)
        assert l.item() == pytest.approx(float(ol), rel=1e-5)
    adam = kind != 'sgd'
    assert_params_close(net_params(net), oracle.parameters(), 1e-5, lr=hyper['lr'], steps=steps, adam=adam)
    # the torch optimizer's state dict shows the live state (aliases of the engine's flat buffers)
    st = opt.state[next(iter(net.parameters()))]
    if kind == 'sgd' and hyper.get('momentum', 0):
        np.testing.assert_allclose(st['momentum_buffer'].cpu().numpy(), oracle.momentum_buffer[0], rtol=1e-4, atol=1e-7)
    if adam and hyper.get('amsgrad'):
        np.testing.assert_allclose(st['max_exp_avg_sq'].cpu().numpy(), oracle.max_exp_avg_sq[0], rtol=1e-4, atol=1e-12)


def test_momentum_resume_takes_over_existing_state():
    """An optimizer that already holds momentum buffers (a resumed run) is continued, not restarted."""
    kind, hyper = 'sgd', dict(lr=0.05, momentum=0.9)
    dims, B = [784, 256, 10], 256
    x, y = synthetic(B * 2)
    init = seeded_init(dims)
    from torchsystem_b200 import MLP
    ref = MLP(dims[0], dims[1], dims[-1], dropout=0.0, activation=nn.ReLU())
    with torch.no_grad():
        for p, w in zip([q for lin in ref.linears() for q in (lin.weight, lin.bias)], init):
            p.copy_(torch.from_numpy(np.asarray(w)))
    crit = nn.CrossEntropyLoss()
    ropt = torch.optim.SGD(ref.parameters(), **hyper)
    for s in range(2):                       # two eager steps on the CPU
        ropt.zero_grad()
        crit(ref(x[s * B:(s + 1) * B].flatten(1)), y[s * B:(s + 1) * B]).backward()
        if s == 0:                           # after step 0: hand the half-trained net + optimizer to the engine
            ropt.step()
            import copy
            net2 = copy.deepcopy(ref)
            opt2 = torch.optim.SGD(net2.parameters(), **hyper)
            for p_new, p_old in zip(net2.parameters(), ref.parameters()):
                opt2.state[p_new] = {'momentum_buffer': ropt.state[p_old]['momentum_buffer'].clone()}
        else:
            ropt.step()
    from torchsystem_b200 import B200Classifier
    model = B200Classifier(net2, crit, opt2, precision='f32', device='cuda:0', max_batch=B)
    model.fit(x[B:2 * B].cuda(), y[B:2 * B].cuda())
    for a, b in zip(net2.parameters(), ref.parameters()):
        np.testing.assert_allclose(a.detach().cpu().numpy(), b.detach().numpy(), rtol=1e-4, atol=2e-6)
