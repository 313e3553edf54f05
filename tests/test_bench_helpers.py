"""bench.py's clock sampler on CPU, with a stand-in for pynvml: the first two queries are over before
`wait_ready` returns, `mark` drops what was seen before a timed region and forces a sample at once, and a
region shorter than the sampling interval still holds a sample."""
import importlib.util
import os
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench_for_sampler_test", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


class FakeNvml:
    NVML_CLOCK_SM = 0

    def __init__(self):
        self.calls = 0
        self.lock = threading.Lock()
        self.reason_bits = 0

    def nvmlDeviceGetClockInfo(self, h, kind):
        with self.lock:
            self.calls += 1
            return 1000 + self.calls

    def nvmlDeviceGetCurrentClocksEventReasons(self, h):
        return self.reason_bits


def test_ready_mark_and_short_region():
    nv = FakeNvml()
    s = bench.ClockSampler(0, interval=5.0, nv=nv, handle=object())  # interval far longer than the "region"
    s.start()
    s.wait_ready(2.0)
    assert nv.calls >= 2, "the first two queries happen before any timed work"
    before = nv.calls
    nv.reason_bits = 0x4
    s.mark()
    time.sleep(0.2)  # a region much shorter than the interval
    out = s.finish()
    assert out["samples"] >= 1, "mark() must force a sample inside the region"
    assert out["samples"] <= 3
    assert out["sm_mhz"] > 1000 + before - 1, "samples from before the mark are dropped"
    assert out["reasons"] == ["sw_power_cap"]


def test_periodic_samples_and_reasons_reset_at_mark():
    nv = FakeNvml()
    nv.reason_bits = 0x8  # a slowdown seen only before the timed region must not be reported for it
    s = bench.ClockSampler(0, interval=0.01, nv=nv, handle=object())
    s.start()
    s.wait_ready(2.0)
    nv.reason_bits = 0
    time.sleep(0.05)
    s.mark()
    time.sleep(0.15)
    out = s.finish()
    assert out["samples"] >= 5
    assert out["reasons"] == []


def test_without_nvml_reports_the_error_and_no_numbers():
    s = bench.ClockSampler(0, interval=0.01, nv=None)
    if s.nv is None:  # the container has no NVML-capable driver
        s.start()
        s.wait_ready(1.0)
        s.mark()
        out = s.finish()
        assert out["samples"] == 0 and out["sm_mhz"] is None and "error" in out


def test_step_form_autotune_is_a_no_op_on_one_gpu_or_when_the_form_is_given():
    import types

    for world, overlap in ((1, "auto"), (8, "on"), (8, "off")):
        args = types.SimpleNamespace(aggregate="auto", overlap=overlap, transport="peer")
        assert bench.tune_overlap("c1", args, 0, world) is None
        assert not hasattr(args, "overlap_tuned")
    # a small workload (C3: 25.6 M parameters) is never pipelined, so there is nothing to time
    args = types.SimpleNamespace(aggregate="auto", overlap="auto", transport="peer")
    assert bench.tune_overlap("c3", args, 0, 8) is None
