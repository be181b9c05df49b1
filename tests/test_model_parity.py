"""TinyGPT vs the reference class: state_dict keys/shapes, parameter counts, logits, loss and
gradients (fp32, dropout 0) — SURVEY §7.5 "Oracle"."""
import pytest
import torch

from b200bench.config import get_model_config
from b200bench.models import TinyGPT, build_layout


def _pair(reference_module, cfg):
    torch.manual_seed(0)
    ref = reference_module.TinyGPT(dropout=0.0, **cfg)
    mine = TinyGPT(dropout=0.0, verbose=False, **cfg)
    sd = ref.state_dict()
    missing, unexpected = mine.load_state_dict(sd, strict=True), None
    return ref.eval(), mine.eval()


def test_state_dict_keys_and_shapes(reference_module):
    cfg = dict(vocab_size=128, n_embd=64, n_head=4, n_layer=2, block_size=16)
    ref = reference_module.TinyGPT(**cfg)
    mine = TinyGPT(verbose=False, **cfg)
    a, b = ref.state_dict(), mine.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert len(a) == 29
    for k in a:
        assert tuple(a[k].shape) == tuple(b[k].shape), k
    assert mine.lm_head.weight.data_ptr() == mine.transformer["wte"].weight.data_ptr()


@pytest.mark.parametrize("tier,seq,expected", [("A", 2048, 236_406_784), ("B", 2048, 1_681_199_104),
                                               ("B", 4096, 1_685_393_408)])
def test_param_counts(tier, seq, expected):
    layout = build_layout(**get_model_config(tier, seq))
    assert layout.num_params() == expected


def test_param_count_matches_module():
    cfg = get_model_config("T", 32)
    m = TinyGPT(verbose=False, **cfg)
    assert m.num_params() == build_layout(**cfg).num_params()


def test_forward_backward_matches_reference(reference_module):
    cfg = dict(vocab_size=96, n_embd=64, n_head=4, n_layer=2, block_size=24)
    ref, mine = _pair(reference_module, cfg)
    idx = torch.randint(0, 96, (3, 24))
    tgt = idx.clone()
    tgt[0, :5] = -1                                    # exercise ignore_index
    lr, lossr = ref(idx, tgt)
    lm, lossm = mine(idx, tgt)
    assert torch.allclose(lr, lm, atol=2e-5, rtol=1e-4)
    assert abs(lossr.item() - lossm.item()) < 1e-5
    lossr.backward()
    lossm.backward()
    gr = dict(ref.named_parameters())
    for name, p in mine.named_parameters():
        assert torch.allclose(gr[name].grad, p.grad, atol=3e-5, rtol=1e-3), name


def test_attention_is_bidirectional_by_default(reference_module):
    """The reference passes no mask (TH:127): changing the LAST token changes position-0 logits."""
    cfg = dict(vocab_size=64, n_embd=32, n_head=2, n_layer=1, block_size=8)
    m = TinyGPT(dropout=0.0, verbose=False, **cfg).eval()
    a = torch.randint(0, 64, (1, 8))
    b = a.clone()
    b[0, -1] = (b[0, -1] + 1) % 64
    assert not torch.allclose(m(a)[0][0, 0], m(b)[0][0, 0])
    mc = TinyGPT(dropout=0.0, causal=True, verbose=False, **cfg).eval()
    assert torch.allclose(mc(a)[0][0, 0], mc(b)[0][0, 0])
