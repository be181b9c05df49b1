"""Data source, timers and the clock sampler (CPU only)."""
import sys
import time
import types

import torch

from b200bench.utils.clocks import ClockSampler
from b200bench.utils.data import PinnedBatchSource, SyntheticDataset
from b200bench.utils.timer import StepTimer

CPU = torch.device("cpu")


def test_synthetic_dataset_is_seeded_like_the_reference():
    # TH:138-150: seed then randint(0, vocab, (size, seq_len)) int64; two constructions give the same tokens
    a = SyntheticDataset(1000, 16, size=50, seed=42, verbose=False)
    b = SyntheticDataset(1000, 16, size=50, seed=42, verbose=False)
    c = SyntheticDataset(1000, 16, size=50, seed=43, verbose=False)
    assert len(a) == 50 and a[0].dtype == torch.int64 and a[0].shape == (16,)
    assert torch.equal(a.data, b.data) and not torch.equal(a.data, c.data)
    assert int(a.data.min()) >= 0 and int(a.data.max()) < 1000


def test_ranks_draw_disjoint_samples_and_reshuffle_per_epoch():
    ds = SyntheticDataset(1000, 8, size=20, seed=1, verbose=False)
    srcs = [PinnedBatchSource(ds, batch=2, rank=r, world_size=2, device=CPU) for r in range(2)]
    seen = [set(), set()]
    for r, s in enumerate(srcs):
        assert s.bytes_per_step == 2 * 8 * 4                    # int32 tokens: what is really copied per step
        for _ in range(5):                                      # 5 batches of 2 = this rank's half of the data
            for row in s.next():
                seen[r].add(tuple(row.tolist()))
    rows = {tuple(x.tolist()) for x in ds.data.to(torch.int32)}
    assert seen[0] | seen[1] == rows and not (seen[0] & seen[1])
    first_epoch_order = srcs[0]._order.clone()
    srcs[0].next()                                              # wraps: new epoch, new permutation
    assert srcs[0].epoch == 1 and not torch.equal(first_epoch_order, srcs[0]._order)


def test_batch_one_is_a_view_of_the_host_copy():
    ds = SyntheticDataset(100, 8, size=4, seed=0, verbose=False)
    s = PinnedBatchSource(ds, batch=1, rank=0, world_size=1, device=CPU)
    t = s.next_host()
    assert t.shape == (1, 8) and t.dtype == torch.int32
    assert t.untyped_storage().data_ptr() == s.host.untyped_storage().data_ptr()


def test_host_timer_keeps_only_post_warmup_steps():
    t = StepTimer("device", CPU)            # no CUDA device: falls back to the reference's host timer
    assert t.mode == "host"
    for step in range(4):
        t.start()
        time.sleep(0.002)
        t.stop(keep=step >= 2)
    assert len(t.times()) == 2 and all(x >= 0.002 for x in t.times())
    assert t.host_times() == t.times()


def _fake_pynvml(reason_bits):
    m = types.ModuleType("pynvml")
    m.NVML_CLOCK_SM = 1
    m.nvmlInit = lambda: None
    m.nvmlDeviceGetHandleByIndex = lambda i: i
    m.nvmlDeviceGetMaxClockInfo = lambda h, k: 1965
    clocks = iter([1965, 1950, 1965, 1965, 1200] + [1965] * 1000)
    m.nvmlDeviceGetClockInfo = lambda h, k: next(clocks)
    m.nvmlDeviceGetPowerUsage = lambda h: 750_000
    m.nvmlDeviceGetCurrentClocksEventReasons = lambda h: reason_bits
    return m


def test_clock_sampler_reports_median_and_reasons(monkeypatch):
    monkeypatch.setitem(sys.modules, "pynvml", _fake_pynvml(0x4 | 0x1))       # sw_power_cap + idle
    s = ClockSampler(0, period=0.005).start()
    time.sleep(0.08)
    out = s.stop()
    assert out["sm_mhz"] == 1965 and out["sm_max_mhz"] == 1965 and out["samples"] >= 5
    assert out["reasons"] == ["sw_power_cap"]                                  # gpu_idle is not a throttle reason
    assert out["power_w_max"] == 750.0
    monkeypatch.setitem(sys.modules, "pynvml", _fake_pynvml(0x8 | 0x40))
    s = ClockSampler(0, period=0.005).start()
    time.sleep(0.03)
    assert set(s.stop()["reasons"]) == {"hw_slowdown", "hw_thermal_slowdown"}


def test_clock_sampler_without_nvml_never_raises(monkeypatch):
    broken = types.ModuleType("pynvml")

    def boom():
        raise RuntimeError("NVML not available")
    broken.nvmlInit = boom
    monkeypatch.setitem(sys.modules, "pynvml", broken)
    out = ClockSampler(0, period=0.005).start().stop()
    assert out["sm_mhz"] is None and out["reasons"] == [] and "error" in out
