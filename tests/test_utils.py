"""Host-side utilities used by bench.py."""
import datetime
import tempfile
import time

from gru_mpi_cuda_b200.utils.clocks import ClockSampler


def _fake_sampler(rows):
    c = ClockSampler(0)
    c.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
    for t, clk, pw, cap in rows:
        stamp = datetime.datetime.fromtimestamp(t).strftime("%Y/%m/%d %H:%M:%S.%f")[:-3]
        c.tmp.write(f"{stamp}, {clk}, 1965, {pw}, Not Active, Not Active, Not Active, "
                    f"{'Active' if cap else 'Not Active'}\n")
    c.tmp.flush()
    return c


def test_clock_summary_uses_the_timed_window():
    now = time.time()
    rows = [(now + 0.02 * i, 1965 if i < 5 else 1320 + i, 200 if i < 5 else 950, i >= 5) for i in range(12)]
    s = _fake_sampler(rows).summary(now + 0.09, now + 0.3)
    assert s["under_load"] == "inside the timed region"
    assert s["samples"] == 7 and s["samples_total"] == 12
    assert 1320 <= s["sm_mhz"] <= 1340 and s["sm_max_mhz"] == 1965
    assert s["reasons"] == ["sw_power_cap"]


def test_clock_summary_falls_back_to_power_when_no_window_matches():
    now = time.time()
    rows = [(now + 0.02 * i, 1965 if i < 5 else 1300, 200 if i < 5 else 950, i >= 5) for i in range(10)]
    s = _fake_sampler(rows).summary(now + 100, now + 200)
    assert s["under_load"] == "top half of observed power"
    assert s["sm_mhz"] == 1300 and s["samples"] == 5


def test_clock_summary_without_nvidia_smi():
    c = ClockSampler(0)
    assert c.summary()["samples"] == 0
