"""The closed-form memory model against the peaks measured on B200 (round 2: profiles/round2/bench_n*.json,
`peak_vram_gb` of the headline and of every `strategies` entry; sequence-length and tier-B rows: round-1 sweep)."""
import pytest

from b200bench.analysis.memory import estimate_memory
from b200bench.config import get_model_config

# tier, strategy, world size, seq_len, measured peak_vram_gb (allocator peak + peer buffers)
MEASURED = [
    ("A", "ddp", 1, 2048, 5.0676), ("A", "ddp", 2, 2048, 5.0713), ("A", "ddp", 8, 2048, 5.0796),
    ("A", "ddp", 1, 512, 4.0782), ("A", "ddp", 1, 8192, 9.0001), ("B", "ddp", 1, 2048, 31.4816),
    ("B", "ddp", 1, 8192, 45.4171),
    ("A", "zero2", 1, 2048, 5.5405), ("A", "zero2", 2, 2048, 3.8901), ("A", "zero2", 4, 2048, 3.0645),
    ("A", "zero2", 8, 2048, 2.6513),
    ("A", "fsdp", 2, 2048, 4.1286), ("A", "fsdp", 4, 2048, 3.1833), ("A", "fsdp", 8, 2048, 2.7109),
    ("A", "zero3", 2, 2048, 4.1286), ("A", "zero3", 8, 2048, 2.7109),
    ("7B", "zero3", 1, 2048, 127.534), ("7B", "zero3", 2, 2048, 75.9841), ("7B", "zero3", 4, 2048, 49.6649),
    ("7B", "zero3", 8, 2048, 36.4947),
]


@pytest.mark.parametrize("tier,strategy,ws,seq,measured", MEASURED)
def test_model_matches_measured_peak(tier, strategy, ws, seq, measured):
    est = estimate_memory(get_model_config(tier, seq), strategy, ws, 1, seq)
    assert est["total"] / 1e9 == pytest.approx(measured, rel=0.02)


def test_components_behave():
    cfg = get_model_config("A", 2048)
    ddp, z2, z3 = (estimate_memory(cfg, s, 8, 1, 2048) for s in ("ddp", "zero2", "zero3"))
    assert z2["master_and_adam_fp32"] == pytest.approx(ddp["master_and_adam_fp32"] / 8, rel=1e-3)
    assert z3["weights_bf16"] < z2["weights_bf16"] / 7 and "gather_staging_bf16" in z3
    assert ddp["activations"] == z2["activations"] == z3["activations"]          # no activation sharding
    # no O(T^2) term: 4x the sequence length is < 4.1x the activations (the reference keeps T x T)
    a1 = estimate_memory(get_model_config("A", 2048), "ddp", 1, 1, 2048)["activations"]
    a4 = estimate_memory(get_model_config("A", 8192), "ddp", 1, 1, 8192)["activations"]
    assert 3.9 < a4 / a1 < 4.1
    # re-sharding after forward needs two staging blocks instead of all of them
    keep = estimate_memory(cfg, "fsdp", 8, 1, 2048, keep_gathered=True)["gather_staging_bf16"]
    two = estimate_memory(cfg, "fsdp", 8, 1, 2048, keep_gathered=False)["gather_staging_bf16"]
    assert keep - two == pytest.approx(14 * 12 * 1024 * 1024 * 2, rel=1e-3)
