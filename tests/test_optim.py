
import torch

from b200bench.config import OptimConfig
from b200bench.optim import FlatAdamW, warmup_lr


def test_flat_adamw_matches_torch():
    torch.manual_seed(0)
    w0 = torch.randn(1000)
    cfg = OptimConfig(lr=1e-3, weight_decay=0.01)
    opt = FlatAdamW(w0, cfg)
    work = w0.clone()
    p = torch.nn.Parameter(w0.clone())
    ref = torch.optim.AdamW([p], lr=1e-3, weight_decay=0.01)
    for _ in range(5):
        g = torch.randn(1000)
        p.grad = g.clone()
        ref.step()
        opt.step(g, work)
    assert torch.allclose(work, p.data, atol=1e-6, rtol=1e-5)


def test_grad_scale_and_warmup():
    cfg = OptimConfig(lr=1e-4, warmup_steps=5)
    assert warmup_lr(0, cfg) == 0.0
    assert warmup_lr(5, cfg) == 1e-4
    assert 0 < warmup_lr(2, cfg) < 1e-4
    assert abs(warmup_lr(3, cfg, log=False) - 1e-4 * 3 / 5) < 1e-12
    w = torch.ones(8)
    a, b = FlatAdamW(w, OptimConfig()), FlatAdamW(w, OptimConfig())
    wa, wb = w.clone(), w.clone()
    a.step(torch.full((8,), 4.0), wa, grad_scale=0.25)
    b.step(torch.full((8,), 1.0), wb)
    assert torch.allclose(wa, wb)
