"""CPU side of ops/linear.py: the PyTorch fallback of a link equals the textbook composition,
the analytic GELU derivative used by the dGELU epilogue's oracle equals autograd, and the MLP
built from links equals a plain nn.Sequential."""
import torch
import torch.nn.functional as F

from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.ops import linear


def test_dgelu_formula_matches_autograd():
    z = torch.linspace(-6, 6, 2001, dtype=torch.float64, requires_grad=True)
    F.gelu(z, approximate="tanh").sum().backward()
    assert torch.allclose(linear._dgelu(z.detach()).double(), z.grad, atol=1e-6)


def test_link_fallback_matches_composition():
    torch.manual_seed(0)
    x = torch.randn(5, 3, 16, requires_grad=True)
    w1, b1 = torch.randn(24, 16, requires_grad=True), torch.randn(24, requires_grad=True)
    w2, b2 = torch.randn(8, 24, requires_grad=True), torch.randn(8, requires_grad=True)
    z, act = linear.link(x, None, w1, b1, gelu_in=False, gelu_out=True)
    assert act is None  # the fallback lets the next link apply its own GELU
    y, _ = linear.link(z, act, w2, b2, gelu_in=True, gelu_out=False)
    ref = F.linear(F.gelu(F.linear(x, w1, b1), approximate="tanh"), w2, b2)
    assert torch.allclose(y, ref, atol=1e-6)
    gy = torch.autograd.grad(y.sum(), [x, w1, b1, w2, b2])
    gr = torch.autograd.grad(ref.sum(), [x, w1, b1, w2, b2])
    for a, b in zip(gy, gr):
        assert torch.allclose(a, b, atol=1e-5)
    zr, ar = linear.link_reference(x, w1, b1, gelu_in=False, gelu_out=True)
    assert torch.allclose(zr, z) and torch.allclose(ar, F.gelu(z, approximate="tanh"))


def test_mlp_of_links_equals_sequential():
    torch.manual_seed(1)
    cfg = MLPConfig(d_in=12, hidden=(20, 28), d_out=6)
    m = MLP(cfg)
    seq = []
    for i, layer in enumerate(m.layers):
        seq.append(layer)
        if i + 1 < len(m.layers):
            seq.append(torch.nn.GELU(approximate="tanh"))
    seq = torch.nn.Sequential(*seq)
    x, y = torch.randn(7, 12), torch.randn(7, 6)
    assert torch.allclose(m(x), seq(x), atol=1e-6)
    assert torch.allclose(mse_loss(m(x), y), F.mse_loss(seq(x), y))


def test_take_sink_counts_uses_only_when_a_gradient_is_wanted():
    class Sink:
        uses = 0

        def acquire(self):
            self.uses += 1
            return self

    p = torch.nn.Parameter(torch.zeros(2))
    assert linear.take_sink(p, True) is None  # no engine attached
    p._psb_sink = Sink()
    assert linear.take_sink(p, False) is None and p._psb_sink.uses == 0
    assert linear.take_sink(p, True) is p._psb_sink and p._psb_sink.uses == 1
    assert linear.take_sink(None, True) is None
