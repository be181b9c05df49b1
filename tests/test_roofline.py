"""Closed-form work / roofline model (CPU only)."""
import pytest

from b200bench.analysis.roofline import NVLINK_GBS, measured_peaks, roofline, step_work
from b200bench.config import get_model_config


def test_flops_and_bytes_closed_form():
    cfg = get_model_config("A", 2048)
    w = step_work(cfg, "ddp", 8, 1, 2048)
    d, L, V, T = 1024, 16, 32000, 2048
    assert w["gemm_flops"] == pytest.approx(6 * (12 * d * d * L + V * d) * T)
    assert w["attention_flops"] == pytest.approx(12 * T * T * d * L)
    # ddp: full optimizer on every rank, two-shot all-reduce moves 2 (N-1)/N S in each direction
    assert w["hbm_optimizer_bytes"] == pytest.approx(28 * 236.4e6, rel=0.02)
    assert w["nvlink_bytes_in"] == pytest.approx(2 * 7 / 8 * 2 * 236.4e6, rel=0.02) == w["nvlink_bytes_out"]
    assert step_work(cfg, "ddp", 1, 1, 2048)["nvlink_bytes_in"] == 0


def test_strategy_ordering():
    cfg = get_model_config("A", 2048)
    ddp = step_work(cfg, "ddp", 8, 1, 2048)
    z2 = step_work(cfg, "zero2", 8, 1, 2048, grad_accum=4)
    z3_keep = step_work(cfg, "zero3", 8, 1, 2048, grad_accum=4, keep_gathered=True)
    z3_reshard = step_work(cfg, "zero3", 8, 1, 2048, grad_accum=4, keep_gathered=False)
    assert z2["hbm_optimizer_bytes"] == pytest.approx(ddp["hbm_optimizer_bytes"] / 8 / 4, rel=1e-3)
    assert z2["nvlink_bytes_in"] == pytest.approx(ddp["nvlink_bytes_in"] / 4, rel=1e-3)
    assert z3_reshard["nvlink_bytes_in"] > 4 * z3_keep["nvlink_bytes_in"]      # gathers every micro-step, twice
    assert ddp["flops"] == z2["flops"] == z3_keep["flops"]


def test_roofline_uses_measured_peaks_and_reports_fraction():
    pk = measured_peaks()
    assert pk["hbm_gbs"] > 1000 and pk["bf16_tflops_sustained"] > 100
    r = roofline(get_model_config("A", 2048), "ddp", 8, 1, 2048, measured_ms=9.60)
    assert r["compute_ms"] == pytest.approx(r["flops"] / (pk["bf16_tflops_sustained"] * 1e12) * 1e3)
    assert r["nvlink_ms"] == pytest.approx(r["nvlink_bytes_in"] / (NVLINK_GBS * 1e9) * 1e3)
    assert r["roofline_ms"] <= r["serial_ms"]
    assert 0 < r["fraction_of_roofline"] < 1 and r["model_tflops_achieved"] == pytest.approx(r["flops"] / 9.6e-3 / 1e12)
