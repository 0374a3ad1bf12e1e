"""Oracles, timing, clock sampling and structured logging."""
import ctypes as C
import json
import math
import subprocess
import os
import threading
import time

from .. import _native


def oracle_sin(a, b, n, rule="left"):
    """Closed-form Riemann sum of sin, long-double evaluation (native)."""
    lib, _ = _native.load()
    return float(lib.bq_oracle_sin_sum(float(a), float(b), int(n), _native.RULES[rule]))


def oracle_fsum(fn, a, b, n, rule="left"):
    """Exactly rounded Riemann sum of a scalar Python function (math.fsum); n <= ~1e7."""
    h = (b - a) / n
    off = {"left": 0.0, "right": 1.0, "mid": 0.5, "trap": 0.0}[rule]
    s = math.fsum(fn(math.fma(i + off, h, a)) if hasattr(math, "fma") else fn(a + (i + off) * h)
                  for i in range(n))
    val = h * s
    if rule == "trap":
        val += 0.5 * h * (fn(b) - fn(a))
    return val


def oracle_profile(n, a=0.0, b=1800.0, rule="left", table=None):
    """Exact (math.fsum) Riemann sum of the lerped profile table."""
    from ..models import default_profile
    tab = table if table is not None else default_profile()

    def f(x):
        i = min(max(int(x), 0), len(tab) - 2)
        return tab[i] + (tab[i + 1] - tab[i]) * (x - i)
    return oracle_fsum(f, a, b, n, rule)


class ClockSampler:
    """Samples SM clocks, power and throttle reasons while a timed region runs.

    NVML (the library nvidia-smi itself reads) is polled in-process every few
    milliseconds, so even a sub-second timed region collects a useful number of
    samples; if the NVML binding is missing, `nvidia-smi` is spawned instead.
    """

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0, period_ms=5):
        self.gpu_index = gpu_index
        self.period = period_ms / 1000.0
        self.samples = []
        self.source = None
        self._stop = threading.Event()
        self._thr = None

    def _nvml_handle(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.gpu_index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[self.gpu_index])
                except (ValueError, IndexError):
                    idx = self.gpu_index
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(idx)
        except Exception:
            return None, None

    def _poll(self):
        nv, h = self._nvml_handle()
        if h is not None:
            self.source = "nvml"
            R = nv
            bits = [("hw_slowdown", getattr(R, "nvmlClocksThrottleReasonHwSlowdown", 0x8)),
                    ("hw_thermal_slowdown", getattr(R, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40)),
                    ("sw_thermal_slowdown", getattr(R, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20)),
                    ("sw_power_cap", getattr(R, "nvmlClocksThrottleReasonSwPowerCap", 0x4))]
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            while not self._stop.is_set():
                try:
                    sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    self.samples.append([str(sm), str(mx), str(pw)] +
                                        ["Active" if mask & b else "Not Active" for _, b in bits])
                except Exception:
                    pass
                self._stop.wait(self.period)
            return
        self.source = "nvidia-smi"
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu_index),
                                      "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits"],
                                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.samples.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        self._thr = threading.Thread(target=self._poll, daemon=True)
        self._thr.start()
        return self

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=10)
        return self.summary()

    def summary(self):
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for row in self.samples:
            try:
                sm.append(float(row[0]))
                mx.append(float(row[1]))
                power.append(float(row[2]))
                for name, v in zip(names, row[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "reasons": sorted(reasons),
                "samples": len(sm), "source": self.source}


def json_line(**kw):
    return json.dumps(kw, sort_keys=False)


class WallTimer:
    def __enter__(self):
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        self.seconds = time.perf_counter() - self.t0
        return False
