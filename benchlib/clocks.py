"""SM clock / throttle sampling through NVML during the timed regions, and the roofline denominators."""
from __future__ import annotations

import json
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent


class ClockSampler(threading.Thread):
    """Samples SM clock / power / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index, period_s=0.002):
        super().__init__(daemon=True)
        self.index, self.period, self.rows, self._stop_ev = index, period_s, [], threading.Event()
        self.max_mhz, self.ok = None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False

    def sample(self):
        nv = self.nv
        mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        try:
            reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        try:
            power = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
        except Exception:
            power = None
        self.rows.append((float(mhz), int(reasons), power, time.perf_counter()))

    def run(self):
        if not self.ok:
            return
        errors = 0
        while not self._stop_ev.is_set() and errors < 100:
            try:
                self.sample()
            except Exception:
                errors += 1  # a transient NVML error must not end the sampling of the timed region
            time.sleep(self.period)

    def mark(self):
        """A point in time: summary(since=mark) covers the samples taken after it."""
        return time.perf_counter()

    def summary(self, since=None):
        rows = [r for r in self.rows if since is None or r[3] >= since]
        if not self.ok or not rows:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        nv = self.nv
        bits = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        seen = set()
        for _, r, _, _ in rows:
            for name, bit in bits.items():
                if r & bit:
                    seen.add(name)
        pw = [p for _, _, p, _ in rows if p is not None]
        return {"sm_mhz": float(np.median([m for m, _, _, _ in rows])), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(seen), "samples": len(rows), "power_w_max": max(pw) if pw else None}

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        return self.summary()


def load_peaks():
    """(hbm GB/s, bf16 sustained TFLOP/s, source) from the driver-written MEASURED_PEAKS.json, else the fallback
    B200_PROFILING.md states."""
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            d = json.loads(p.read_text())
            return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", 1400.0)), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, 1400.0, "fallback (B200_PROFILING.md: 6.65 TB/s, 1.4 PFLOP/s sustained)"


def load_traffic(name="k1_traffic.json"):
    """Counters of a committed ncu --set full capture of the traversal kernel (profiles/k1_traffic*.json), if any."""
    p = ROOT / "profiles" / name
    if p.exists():
        try:
            return json.loads(p.read_text())
        except Exception:
            return None
    return None
