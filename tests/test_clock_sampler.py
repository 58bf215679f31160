"""Clock sampling used by bench.py (tenzing_b200/utils/clocks.py), with a stand-in for nvidia-smi."""
import os
import sys
import time

from tenzing_b200.utils.clocks import ClockSampler, parse_smi_lines

FAKE_SMI = r"""
import sys, time
rows = ["0, 1965, 1965, 512.3, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
        "0, 1920, 1965, 699.0, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
        "0, 1965, 1965, 640.1, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active"]
time.sleep(0.15)            # nvidia-smi takes a while to come up
i = 0
while True:
    print(rows[i % len(rows)], flush=True)
    i += 1
    time.sleep(0.02)
"""


def test_parse_smi_lines_reads_clocks_and_reasons():
    s = parse_smi_lines(["0, 1965, 1965, 500.0, 0x0, Not Active, Not Active, Not Active, Not Active",
                         "0, 1500, 1965, 700.0, 0x48, Active, Active, Not Active, Not Active",
                         "garbage", "0, [N/A], 1965, 1, 0, a, b, c, d"])
    assert s == [(1965.0, 1965.0, set()), (1500.0, 1965.0, {"hw_slowdown", "hw_thermal_slowdown"})]


def test_start_waits_for_the_first_sample_so_a_short_region_is_covered():
    sampler = ClockSampler(0, smi_cmd=[sys.executable, "-c", FAKE_SMI], use_nvml=False)
    t0 = time.time()
    sampler.start(wait_first_sample_s=5.0)
    waited = time.time() - t0
    assert 0.1 < waited < 4.9            # came back after the first line, not after the timeout
    # the "timed region": long enough for three more lines even on a loaded machine
    deadline = time.time() + 10.0
    while time.time() < deadline and os.path.getsize(sampler.path) < sampler._smi_ready + 3 * 60:
        time.sleep(0.02)
    c = sampler.stop()
    assert c["source"] == "nvidia-smi" and c["samples"] >= 3
    assert c["sm_mhz"] == 1965.0 and c["sm_max_mhz"] == 1965.0 and c["sm_min_mhz"] == 1920.0
    assert c["reasons"] == ["sw_power_cap"]


def test_without_any_source_the_result_is_explicitly_empty():
    sampler = ClockSampler(0, smi_cmd=["/nonexistent/nvidia-smi"], use_nvml=False)
    t0 = time.time()
    sampler.start(wait_first_sample_s=2.0)
    assert time.time() - t0 < 1.0        # nothing alive to wait for
    assert sampler.stop() == {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0,
                              "source": None}
    # NVML is absent on a CPU-only box: asking for it must not raise either
    s2 = ClockSampler(0, smi_cmd=["/nonexistent/nvidia-smi"], use_nvml=True).start(0.2)
    assert s2.stop()["samples"] == 0
