"""Samples SM clocks and throttle reasons through NVML while a timed region runs.

A polling thread (a few ms period) is used instead of ``nvidia-smi -lms`` because the timed
region of a communication benchmark is only milliseconds long.
"""
from __future__ import annotations

import statistics
import threading
import time

_REASON_BITS = {
    "hw_slowdown": 0x8,            # nvmlClocksThrottleReasonHwSlowdown
    "sw_power_cap": 0x4,           # nvmlClocksThrottleReasonSwPowerCap
    "sw_thermal_slowdown": 0x20,   # nvmlClocksThrottleReasonSwThermalSlowdown
    "hw_thermal_slowdown": 0x40,   # nvmlClocksThrottleReasonHwThermalSlowdown
}


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: float = 2.0):
        self.gpu_index, self.period = gpu_index, period_ms / 1e3
        self._stop = threading.Event()
        self._thread = None
        self._sm, self._reasons = [], set()
        self._max = None
        self.summary = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}

    def _loop(self, nv, handle):
        while not self._stop.is_set():
            try:
                self._sm.append(nv.nvmlDeviceGetClockInfo(handle, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(handle)
                for name, bit in _REASON_BITS.items():
                    if mask & bit:
                        self._reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            import os

            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.gpu_index
            if visible:
                ids = [v for v in visible.split(",") if v.strip() != ""]
                if idx < len(ids) and ids[idx].strip().isdigit():
                    idx = int(ids[idx])
            handle = nv.nvmlDeviceGetHandleByIndex(idx)
            self._max = nv.nvmlDeviceGetMaxClockInfo(handle, nv.NVML_CLOCK_SM)
            self._thread = threading.Thread(target=self._loop, args=(nv, handle), daemon=True)
            self._thread.start()
        except Exception:
            self._thread = None
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        self.summary = {"sm_mhz": statistics.median(self._sm) if self._sm else None, "sm_max_mhz": self._max,
                        "reasons": sorted(self._reasons), "samples": len(self._sm)}
        return False
