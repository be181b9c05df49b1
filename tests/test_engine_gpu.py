"""Native engine vs the torch fp32 oracle: same weights, same batch, dropout 0 -> same loss, the same gradient
for EVERY parameter tensor (biases, LayerNorm vectors and position embeddings included), the same update
direction of the fp32 master weights; loss goes down over a few steps."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(strategy="ddp", use_graph=False, dropout=0.0, B=2, T=128, model_cfg=None, grad_accum=1, mode="reference"):
    from b200bench.config import StrategyConfig, get_model_config
    from b200bench.engine import NativeEngine
    from b200bench.parallel.collectives import DistCtx
    from b200bench.parallel.torch_engine import TorchEngine

    dev = torch.device("cuda:0")
    cfg = dict(model_cfg) if model_cfg is not None else get_model_config("T", T)
    sc, sc2 = StrategyConfig(), StrategyConfig()
    sc.grad_accum_mode = sc2.grad_accum_mode = mode
    nat = NativeEngine(cfg, strategy, sc, DistCtx(), dev, dropout=dropout, batch=B, seq_len=T, use_graph=use_graph,
                       grad_accum=grad_accum)
    ref = TorchEngine(cfg, strategy, sc2, DistCtx(), dev, torch.float32, dropout=dropout, grad_accum=grad_accum)
    ref.load_state_dict(nat.state_dict())
    return nat, ref


def _check_against_oracle(nat, ref, min_cos_matrix=0.99, min_cos_vector=0.97):
    """After ONE optimizer step from identical weights: Adam's first moment is 0.1 * g, so it carries the gradient
    of every tensor; the master weights moved by about lr * sign(g)."""
    Lt = nat.layout
    g_nat, g_ref = nat.exp_avg.float(), ref.opt.exp_avg.float()
    w_nat, w_ref = nat.master.float(), ref.opt.master.float()
    bad = []
    for s in Lt:
        a, b = g_nat[s.offset:s.offset + s.numel], g_ref[s.offset:s.offset + s.numel]
        cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
        need = min_cos_matrix if len(s.shape) == 2 else min_cos_vector
        # relative size of the gradient: a wrong scale (missing 1/n_valid, a doubled accumulation) changes the norm
        ratio = (a.norm() / b.norm().clamp_min(1e-20)).item()
        if not (cos >= need and 0.9 < ratio < 1.1):
            bad.append(f"{s.name}: cos {cos:.4f} (need {need}), |g| ratio {ratio:.3f}")
    assert not bad, "gradient mismatch:\n  " + "\n  ".join(bad)
    # update direction of the fp32 master weights, weighted by |g| so that entries whose true gradient is ~0
    # (e.g. the key bias: softmax is invariant to it) do not vote with a random sign
    w0 = ref._w0                       # both engines started from these values (the native bf16 weights, as fp32)
    d_nat, d_ref = w_nat - w0, w_ref - w0
    used = slice(0, Lt.total)
    wgt = g_ref[used].abs()
    agree = ((torch.sign(d_nat[used]) == torch.sign(d_ref[used])).float() * wgt).sum() / wgt.sum()
    assert agree.item() > 0.99, f"weighted update-sign agreement {agree.item():.4f}"
    # size of the update: one Adam step moves a weight by ~lr (bias-corrected m / sqrt(v) = +-1)
    rel = (d_nat[used] - d_ref[used]).abs().sum() / d_ref[used].abs().sum()
    assert rel.item() < 0.15, f"relative update error {rel.item():.4f}"


def _one_step(nat, ref, B, T, V, seed=0, steps=1):
    torch.manual_seed(seed)
    ref._w0 = ref.opt.master.clone()
    for i in range(steps):
        tok = torch.randint(0, V, (B, T), device="cuda", dtype=torch.int32)
        l_nat = nat.train_step(tok, tok).clone()
        l_ref = ref.train_step(tok, tok)
        assert abs(l_nat.item() - float(l_ref)) < 0.02 * max(1.0, abs(float(l_ref))), (i, l_nat.item(), float(l_ref))


def test_loss_and_update_match_oracle():
    nat, ref = _mk()
    _one_step(nat, ref, 2, 128, 512)
    _check_against_oracle(nat, ref)
    nat.close()


def test_tier_a_width_matches_oracle():
    """d = 1024, 16 heads, V = 32000 (two blocks, T = 256): the CTA-pair 256 x 256 GEMM tiles, the register-cached
    cross-entropy row and the wide LayerNorm / column-sum paths that the tiny tier never reaches."""
    cfg = dict(vocab_size=32000, n_embd=1024, n_head=16, n_layer=2, block_size=256)
    nat, ref = _mk(B=1, T=256, model_cfg=cfg)
    _one_step(nat, ref, 1, 256, 32000, seed=3)
    _check_against_oracle(nat, ref, min_cos_matrix=0.985, min_cos_vector=0.96)
    nat.close()


def test_grad_accumulation_matches_oracle():
    """Two micro-steps per optimizer step (--grad-accum-mode uniform): accumulate flags of every wgrad / bias / LN /
    embedding gradient path, and the 1/accum scale."""
    nat, ref = _mk(grad_accum=2, mode="uniform")
    assert nat.accum == 2 and ref.accum == 2
    _one_step(nat, ref, 2, 128, 512, seed=5, steps=2)
    _check_against_oracle(nat, ref)
    nat.close()


@pytest.mark.parametrize("use_graph", [False, True])
def test_loss_decreases(use_graph):
    torch.manual_seed(1)
    nat, _ = _mk(use_graph=use_graph, dropout=0.1)
    tok = torch.randint(0, 512, (2, 128), device="cuda", dtype=torch.int32)
    losses = [nat.train_step(tok, tok).item() for _ in range(30)]
    assert losses[-1] < losses[0] - 0.05
    assert nat.kernel_launches_per_step() > 20
    nat.close()
