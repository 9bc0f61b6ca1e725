"""Host logic of the Python mirror that can run without a GPU: the autograd node that makes eval()
differentiable with respect to the table (its wiring, not its arithmetic: the plugin and the library are
duck-typed stand-ins that record what reaches them)."""
from types import SimpleNamespace

import pytest

torch = pytest.importorskip("torch")


def test_eval_autograd_node_wiring(monkeypatch):
    from mitsuba_customization_b200 import bsdf
    cleared = []

    class FakeLib:
        def merl_grad_clear_counts(self, h, p, s):
            cleared.append(p.value)
            return 0

    monkeypatch.setattr(bsdf._cabi, "lib", lambda: FakeLib())
    monkeypatch.setattr(bsdf, "_stream", lambda d: None)

    class Plug:
        info = SimpleNamespace(ncell=5, d_cells=None)
        _h = None
        seen = None

        def _enabled(self, ctx):
            return True

        def eval(self, ctx, si, wo, active):
            return torch.ones(3, 4)

        def backward(self, si, wo, g, active=True, grad_cells=None):
            Plug.seen = (si, wo, g.clone(), active)
            grad_cells[:, :3] += g.sum()
            grad_cells[:, 3] += 4
            return grad_cells

    cells = torch.zeros(5, 4, requires_grad=True)
    out = bsdf._EvalOp.apply(cells, Plug(), "ctx", "si", "wo", "mask")
    assert out.requires_grad and out.shape == (3, 4)
    (out * 2).sum().backward()
    assert Plug.seen[0] == "si" and Plug.seen[1] == "wo" and Plug.seen[3] == "mask"
    assert torch.equal(Plug.seen[2], torch.full((3, 4), 2.0))           # grad_out reached merl_backward's wrapper
    assert cells.grad.shape == (5, 4) and bool((cells.grad[:, :3] == 24).all())
    assert len(cleared) == 1                                              # the hit-count lane is handed to the library to clear


def test_eval_ad_refuses_a_foreign_tensor():
    from mitsuba_customization_b200 import bsdf
    plug = bsdf.MERL.__new__(bsdf.MERL)
    plug.info = SimpleNamespace(ncell=5, d_cells=1234)
    with pytest.raises(bsdf.MerlError):
        plug.eval_ad(None, None, None, torch.zeros(5, 4))
